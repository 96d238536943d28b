"""GPU parity tests of the scan / compact library, through the C ABI (include/b200_sc.h).

Mirrors the reference's self-checking harness (stream_compaction/src/main.cpp:29-140): the CPU
result is the truth, every GPU variant is compared element-wise, POT and NPOT (= POT-3), last POT
element 0.  The bar is BIT-EXACT (int32, two's-complement wrap).  Full-size cases (2^28) use the
numpy closed form of the same definition (exclusive cumsum mod 2^32; x[x != 0]) which
test_oracle_sc.py ties to the oracle and the reference's golden vectors."""
import os

import numpy as np
import pytest

from oracle_bindings import ScOracle, gen_array

pytestmark = pytest.mark.gpu

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "sc_golden.npz"))


@pytest.fixture(scope="module")
def sc():
    from cuda_pathtracer_b200 import stream_compaction as m
    return m


@pytest.fixture(scope="module")
def oracle():
    return ScOracle()


def np_exclusive_scan(a):
    c = np.cumsum(a.astype(np.int64))
    out = np.empty(a.size, dtype=np.int64)
    out[0:1] = 0
    out[1:] = c[:-1]
    return out.astype(np.uint64).astype(np.uint32).view(np.int32)


VARIANTS = ["Naive", "Efficient", "Thrust", "SinglePassTile", "SinglePassRing", "SinglePassWave"]
SIZES = [1, 2, 3, 4, 5, 31, 32, 33, 63, 64, 65, 255, 256, 257, 4095, 4096, 4097, 8191, 12289,
         (1 << 16) - 3, 1 << 16, (1 << 20) - 3, 1 << 20]


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("n", SIZES)
def test_scan_vs_oracle(sc, oracle, variant, n):
    pot = 1 << int(np.ceil(np.log2(max(n, 2))))
    a = gen_array(pot, 50)[:n].copy()
    o = np.full(n, -7, dtype=np.int32)
    getattr(sc, variant).scan(n, o, a)
    assert np.array_equal(o, oracle.scan(a))


@pytest.mark.parametrize("n", SIZES)
def test_compact_vs_oracle(sc, oracle, n):
    pot = 1 << int(np.ceil(np.log2(max(n, 2))))
    a = gen_array(pot, 4)[:n].copy()
    want = oracle.compact_without_scan(a)
    o = np.full(n, -7, dtype=np.int32)
    cnt = sc.Efficient.compact(n, o, a)
    assert cnt == want.size
    assert np.array_equal(o[:cnt], want)
    assert np.all(o[cnt:] == -7)          # nothing written past the count


@pytest.mark.parametrize("variant", VARIANTS)
def test_golden_vectors(sc, variant):
    a = GOLD["scan_in"]
    for src, want in ((a, GOLD["scan_pot"]), (a[:-3].copy(), GOLD["scan_npot"]),
                      (GOLD["wrap_in"], GOLD["wrap_scan"])):
        o = np.zeros_like(src)
        getattr(sc, variant).scan(src.size, o, src)
        assert np.array_equal(o, want)
    b = GOLD["compact_in"]
    for src, want in ((b, GOLD["compact_pot"]), (b[:-3].copy(), GOLD["compact_npot"])):
        o = np.zeros_like(src)
        cnt = sc.Efficient.compact(src.size, o, src)
        assert cnt == want.size and np.array_equal(o[:cnt], want)


def test_edge_cases(sc):
    e = np.zeros(0, dtype=np.int32)
    sc.Thrust.scan(0, e, e)
    assert sc.Efficient.compact(0, e, e) == 0
    z = np.zeros(10000, dtype=np.int32)
    o = np.ones_like(z)
    assert sc.Efficient.compact(z.size, o, z) == 0            # everything removed
    ones = np.ones(10000, dtype=np.int32)
    assert sc.Efficient.compact(ones.size, o, ones) == 10000  # nothing removed
    assert np.array_equal(o, ones)
    for v in VARIANTS:
        getattr(sc, v).scan(ones.size, o, ones)
        assert np.array_equal(o, np.arange(10000, dtype=np.int32))
    neg = np.array([5, -3, 0, 7, -2**31, 1], dtype=np.int32)   # negatives are "non-zero"
    o = np.zeros_like(neg)
    assert sc.Efficient.compact(neg.size, o, neg) == 5
    assert np.array_equal(o[:5], neg[neg != 0])


def test_unaligned_pointers(sc, oracle):
    """host arrays at 4-byte (not 16-byte) alignment, and device pointers at odd offsets"""
    import torch
    base = gen_array(1 << 14, 50)
    a = base[1:9001]
    assert a.ctypes.data % 16 != 0
    o = np.zeros(9000, dtype=np.int32)
    sc.Thrust.scan(9000, o, np.ascontiguousarray(a))
    assert np.array_equal(o, oracle.scan(a))
    d = torch.from_numpy(base).cuda()
    dout = torch.zeros(1 << 14, dtype=torch.int32, device="cuda")
    for v in VARIANTS:
        dout.zero_()
        if v == "SinglePassRing":                              # bulk copies need 16-byte alignment: asked for by name, it refuses
            with pytest.raises(Exception, match="10001|argument"):
                getattr(sc, v).scan(9000, dout[3:], d[1:])
            continue
        getattr(sc, v).scan(9000, dout[3:], d[1:])             # device pointers, both misaligned
        torch.cuda.synchronize()
        assert np.array_equal(dout[3:9003].cpu().numpy(), oracle.scan(base[1:9001]))
    dout.zero_()
    cnt = sc.Efficient.compact(9000, dout[1:], d[5:])
    want = oracle.compact_without_scan(base[5:9005])
    assert cnt == want.size and np.array_equal(dout[1:1 + cnt].cpu().numpy(), want)


def test_device_ctx_seed_total_and_reduce(sc):
    import torch
    ctx = sc.Context()
    n = (1 << 18) + 5
    a = torch.randint(-1000, 1000, (n,), dtype=torch.int32, device="cuda")
    out = torch.empty_like(a)
    seed = torch.tensor([123456], dtype=torch.int32, device="cuda")
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    want = np_exclusive_scan(a.cpu().numpy())
    for variant in (sc.LOOKBACK, sc.LOOKBACK_TILE, sc.LOOKBACK_RING, sc.LOOKBACK_WAVE, sc.EFFICIENT, sc.NAIVE):
        out.zero_(); total.zero_()
        ctx.scan(n, out, a, variant, d_seed=seed, d_total=total)
        torch.cuda.synchronize()
        assert ctx.last_launches() >= 1
        assert np.array_equal(out.cpu().numpy(), (want.astype(np.int64) + 123456).astype(np.int32))
        assert int(total.item()) == int(np.int32(np.int64(a.sum().item()) + 123456))
    ctx.reduce(n, a, total)
    torch.cuda.synchronize()
    assert int(total.item()) == int(a.sum().item())
    ctx.close()


def test_repeated_launches_reuse_workspace(sc):
    """descriptor epochs: many back-to-back launches of different sizes on one context"""
    import torch
    ctx = sc.Context()
    rng = np.random.RandomState(1)
    for _ in range(40):
        n = int(rng.randint(1, 300000))
        a = torch.randint(0, 4, (n,), dtype=torch.int32, device="cuda")
        out = torch.empty_like(a)
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        ctx.scan(n, out, a)
        ring = torch.empty_like(a)
        ctx.scan(n, ring, a, sc.LOOKBACK_RING)                 # shares the ticket counters with the compaction kernel
        assert torch.equal(ring, out)
        ctx.compact(n, out, a, cnt)
        torch.cuda.synchronize()
        h = a.cpu().numpy()
        k = int(cnt.item())
        assert k == int((h != 0).sum())
        assert np.array_equal(out[:k].cpu().numpy(), h[h != 0])
    ctx.close()


@pytest.mark.parametrize("delta", [0, 3])
def test_full_size_2p28(sc, delta):
    """BASELINE size (2^28, and 2^28-3): bit-exact against the closed form, data resident on the
    device (torch RNG instead of glibc rand so the test stays fast); int32 wrap exercised because
    the running sum passes 2^31."""
    import torch
    n = (1 << 28) - delta
    g = torch.Generator(device="cuda").manual_seed(565)
    a = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g)
    a[-1] = 0
    out = torch.empty_like(a)
    ctx = sc.Context()
    want = torch.cumsum(a, 0, dtype=torch.int64)
    want = (want - a).to(torch.int32)                 # exclusive, wraps like int32
    for variant in (sc.LOOKBACK, sc.LOOKBACK_TILE, sc.LOOKBACK_RING, sc.EFFICIENT):
        out.zero_()
        ctx.scan(n, out, a, variant)
        torch.cuda.synchronize()
        assert torch.equal(out, want)
    del want
    b = a % 4
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx.compact(n, out, b, cnt)
    torch.cuda.synchronize()
    keep = b[b != 0]
    assert int(cnt.item()) == keep.numel()
    assert torch.equal(out[: keep.numel()], keep)
    ctx.close()


@pytest.mark.parametrize("delta", [0, 3])
def test_max_size_2p30(sc, delta):
    """Largest size of BASELINE config 3 (2^30 and 2^30-3 ints, 4 GiB per array): bit-exact against the
    closed form, and the size-independent properties -- out[i+1] - out[i] == in[i] (mod 2^32) for every
    i, total == sum, compaction keeps order and is idempotent."""
    import torch
    n = (1 << 30) - delta
    g = torch.Generator(device="cuda").manual_seed(566)
    a = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g)
    out = torch.empty_like(a)
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx = sc.Context()
    ctx.scan(n, out, a, sc.LOOKBACK, d_total=total)
    torch.cuda.synchronize()
    assert int(out[0].item()) == 0
    assert torch.equal(out[1:] - out[:-1], a[:-1])                  # int32 arithmetic wraps on both sides
    assert int(total.item()) == int(np.int64(a.sum(dtype=torch.int64).item()).astype(np.int32))
    want = torch.cumsum(a, 0, dtype=torch.int64)
    want -= a
    assert torch.equal(out, want.to(torch.int32))
    del want
    a %= 4
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx.compact(n, out, a, cnt)
    torch.cuda.synchronize()
    k = int(cnt.item())
    assert k == int((a != 0).sum().item())
    assert torch.equal(out[:k], a[a != 0])
    once = out[:k].clone()
    ctx.compact(k, a, once, cnt)                                     # compacting a compacted array changes nothing
    torch.cuda.synchronize()
    assert int(cnt.item()) == k and torch.equal(a[:k], once)
    ctx.close()


def test_scan_is_linear_and_shift_consistent(sc):
    """scan(x + y) == scan(x) + scan(y) (mod 2^32); scan of a shifted array is the shifted scan minus its head"""
    import torch
    n = (1 << 22) + 77
    g = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    y = torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int64, device="cuda", generator=g).to(torch.int32)
    ctx = sc.Context()
    sx, sy, sxy, sh = (torch.empty_like(x) for _ in range(4))
    for variant in (sc.LOOKBACK, sc.LOOKBACK_TILE, sc.LOOKBACK_RING, sc.LOOKBACK_WAVE, sc.EFFICIENT):
        ctx.scan(n, sx, x, variant)
        ctx.scan(n, sy, y, variant)
        ctx.scan(n, sxy, x + y, variant)
        ctx.scan(n - 5, sh, x[5:].clone(), variant)     # (a copy: 16-byte aligned, which the ring kernel asks for)
        torch.cuda.synchronize()
        assert torch.equal(sxy, sx + sy)
        assert torch.equal(sh[: n - 5], sx[5:] - sx[5])
    ctx.close()


@pytest.mark.parametrize("n", [1, 3, 4, 5, 8191, 8192, 8193, 8195, 7 * 8192, 7 * 8192 + 1, 148 * 8192 - 3, 148 * 8192 * 3 + 2,
                               (1 << 23) - 3, 1 << 23, (1 << 24) + 1])
def test_ring_scan_ragged_sizes(sc, n):
    """the TMA-ring kernel by name at sizes around its stage (8192 ints), ring (7 stages) and grid (148 CTAs) boundaries,
    with a seed and a total, against the closed form; LOOKBACK itself takes this kernel from 2^24 ints on"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(n)
    a = torch.randint(-50, 50, (n,), dtype=torch.int32, device="cuda", generator=g)
    out = torch.full((n + 8,), -7, dtype=torch.int32, device="cuda")
    seed = torch.tensor([-99], dtype=torch.int32, device="cuda")
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx = sc.Context()
    want = (torch.cumsum(a, 0, dtype=torch.int64) - a - 99).to(torch.int32)
    for variant in (sc.LOOKBACK_RING, sc.LOOKBACK):
        out.fill_(-7); total.zero_()
        ctx.scan(n, out, a, variant, d_seed=seed, d_total=total)
        torch.cuda.synchronize()
        assert torch.equal(out[:n], want)
        assert bool((out[n:] == -7).all())                      # nothing written past n
        assert int(total.item()) == int(np.int64(a.sum(dtype=torch.int64).item()) - 99)
    ctx.close()


@pytest.mark.parametrize("n", [1, 3, 4, 5, 8191, 8192, 8193, 8195, 7 * 8192 + 1, 148 * 8192 - 3, 148 * 8192 * 3 + 2, (1 << 24) + 1])
@pytest.mark.parametrize("keep_mod", [1, 2, 4, 1000, 0])
def test_ring_compact_ragged_sizes(sc, n, keep_mod):
    """the TMA-ring compaction (forced with set_ring_min_n(1)) at sizes around its stage / ring / grid boundaries, at
    several densities (keep_mod 1: everything kept, 0: nothing kept), into 16-byte aligned and misaligned outputs, against
    x[x != 0]; the register-tile kernel (ring off) must give the same"""
    import torch
    g = torch.Generator(device="cuda").manual_seed(n + keep_mod)
    a = torch.randint(1, 1 << 30, (n,), dtype=torch.int32, device="cuda", generator=g)
    if keep_mod == 0:
        a.zero_()
    elif keep_mod > 1:
        a = torch.where(a % keep_mod == 0, a, torch.zeros_like(a)) if keep_mod == 1000 else a % keep_mod
    want = a[a != 0]
    k = want.numel()
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    ctx = sc.Context()
    for ring_min in (1, 2**31 - 1):
        ctx.set_ring_min_n(ring_min)
        for off in (0, 1, 2, 3):
            out = torch.full((n + 12,), -7, dtype=torch.int32, device="cuda")
            cnt.fill_(-1)
            ctx.compact(n, out[off:], a, cnt)
            torch.cuda.synchronize()
            assert int(cnt.item()) == k
            assert torch.equal(out[off:off + k], want)
            assert bool((out[:off] == -7).all()) and bool((out[off + k:] == -7).all())   # nothing outside the run
    ctx.close()


def test_ring_kernels_interleaved_on_one_context(sc):
    """ring scan, ring compact and the register-tile kernels back to back on one context (shared descriptors, epochs and
    ticket counters), sizes that leave CTAs idle and sizes that do not"""
    import torch
    ctx = sc.Context()
    ctx.set_ring_min_n(1)
    rng = np.random.RandomState(3)
    for it in range(30):
        n = int(rng.choice([5, 8192 * 3 + 1, 8192 * 148, 8192 * 400 + 77, 1 << 22]))
        a = torch.randint(0, 4, (n,), dtype=torch.int32, device="cuda")
        out, out2 = torch.empty_like(a), torch.empty_like(a)
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        ctx.scan(n, out, a)
        ctx.compact(n, out2, a, cnt)
        ctx.scan(n, out, a, sc.LOOKBACK_TILE)
        torch.cuda.synchronize()
        assert torch.equal(out, (torch.cumsum(a, 0, dtype=torch.int64) - a).to(torch.int32))
        keep = a[a != 0]
        assert int(cnt.item()) == keep.numel() and torch.equal(out2[: keep.numel()], keep)
    ctx.close()


def test_wave_scan_sizes_and_limit(sc):
    """the single-wave kernel by name: every tile shape (896 threads x 1 / 2 / 4 / 8 vectors), ragged ends, misaligned
    pointers, up to its limit of #SMs x 28672 ints; one element more is refused by name and served by LOOKBACK"""
    import torch
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    limit = sms * 28672
    ctx = sc.Context()
    big = torch.randint(-50, 50, (limit + 9,), dtype=torch.int32, device="cuda")
    ref = torch.cumsum(big, 0, dtype=torch.int64)
    out = torch.empty(limit + 16, dtype=torch.int32, device="cuda")
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    for n in (1, 2, 3583, 3584, 3585, sms * 3584 - 1, sms * 3584, sms * 3584 + 1, sms * 7168 + 5, sms * 14336 - 3,
              sms * 14336 + 4, 1 << 22, limit - 1, limit):
        for off in (0, 1):
            a = big[off:off + n]
            want = (ref[off:off + n] - a - (ref[off - 1] if off else 0)).to(torch.int32)
            for variant in (sc.LOOKBACK_WAVE, sc.LOOKBACK):
                out.fill_(-7)
                ctx.scan(n, out[off:], a, variant, d_total=total)
                torch.cuda.synchronize()
                assert torch.equal(out[off:off + n], want), (n, off, variant)
                assert bool((out[off + n:] == -7).all()) and bool((out[:off] == -7).all())
                assert int(total.item()) == int((want[-1] + a[-1]).item())
    with pytest.raises(Exception, match="10001|argument"):
        ctx.scan(limit + 1, out, big[: limit + 1], sc.LOOKBACK_WAVE)
    ctx.scan(limit + 1, out, big[: limit + 1], sc.LOOKBACK)
    torch.cuda.synchronize()
    assert torch.equal(out[: limit + 1], (ref[: limit + 1] - big[: limit + 1]).to(torch.int32))
    ctx.close()
