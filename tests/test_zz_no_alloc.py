"""b200_scan_i32_noalloc (include/b200_sc.h): the allocation-free variant of the single-pass look-back scan, its scratch
in a file-scope static __device__ array of the library.  No counterpart in the reference -- added
for callers that cannot call cudaMalloc.  Checked through the C ABI against the existing entry point
(b200_scan_i32_dev) and the CPU oracle, BIT FOR BIT (the tolerance of every scan test here: int32, exact), on calls in a
row that reuse the shared scratch.  This file is named to be collected last."""
import ctypes

import numpy as np
import pytest

from oracle_bindings import ScOracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch():
    import torch as t
    return t


def test_noalloc_scan_matches_existing_entry_point_twice_in_a_row(torch):
    from cuda_pathtracer_b200 import stream_compaction as sc
    lib = ctypes.CDLL(sc._lib()._name)                       # the C ABI itself
    lib.b200_scan_i32_noalloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.b200_scan_i32_noalloc.restype = ctypes.c_int
    stream = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device="cuda").manual_seed(565)
    sizes = [(1 << 22) - 3, (1 << 22) - 3, 4099, 1 << 24, 1, 12289]      # same size twice, then smaller / larger: stale scratch
    ins = [torch.randint(-50, 50, (n,), dtype=torch.int32, device="cuda", generator=g) for n in sizes]
    for a in ins:
        a[::5] = 2 ** 30                                     # the running sum wraps (int32)
    outs = [torch.full((n + 4,), -7, dtype=torch.int32, device="cuda") for n in sizes]
    wants = [torch.empty_like(a) for a in ins]
    ctx = sc.Context()
    for a, w in zip(ins, wants):
        ctx.scan(a.numel(), w, a, sc.LOOKBACK, stream=stream)            # the existing entry point
    assert lib.b200_scan_i32_noalloc(16, outs[0].data_ptr(), ins[0].data_ptr(), stream) == 0   # set-up call (symbol address)
    torch.cuda.synchronize()
    outs[0].fill_(-7)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for a, o in zip(ins, outs):                                          # back to back on one stream, no sync in between
        assert lib.b200_scan_i32_noalloc(a.numel(), o.data_ptr(), a.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before                   # nothing was allocated on the device
    oracle = ScOracle()
    for a, o, w in zip(ins, outs, wants):
        n = a.numel()
        assert torch.equal(o[:n], w), f"n = {n}: differs from b200_scan_i32_dev"
        assert bool((o[n:] == -7).all()), f"n = {n}: wrote past the end"
        if n <= (1 << 22):
            assert np.array_equal(o[:n].cpu().numpy(), oracle.scan(a.cpu().numpy())), f"n = {n}: differs from the oracle"
    ctx.close()


def test_noalloc_scan_limits(torch):
    from cuda_pathtracer_b200 import stream_compaction as sc
    from cuda_pathtracer_b200 import B200Error
    a = torch.ones(8, dtype=torch.int32, device="cuda")
    o = torch.zeros(8, dtype=torch.int32, device="cuda")
    with pytest.raises(B200Error) as e:                                  # beyond the fixed scratch: a clear error, nothing enqueued
        sc.scan_noalloc(sc.NOALLOC_MAX_N + 1, o, a)
    assert e.value.code == sc.E_TOOLARGE and "NOALLOC_MAX_N" in str(e.value)
    with pytest.raises(B200Error):
        sc.scan_noalloc(-1, o, a)
    sc.scan_noalloc(0, o, a)                                             # empty: ok, nothing happens
    sc.scan_noalloc(8, o, a)
    torch.cuda.synchronize()
    assert o.tolist() == list(range(8))
    # the largest size the scratch serves
    n = sc.NOALLOC_MAX_N
    big = torch.ones(n, dtype=torch.int32, device="cuda")
    out = torch.empty_like(big)
    sc.scan_noalloc(n, out, big)
    torch.cuda.synchronize()
    assert int(out[-1]) == n - 1 and int(out[n // 2]) == n // 2 and int(out[0]) == 0
