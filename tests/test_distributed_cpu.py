"""N > 1 host logic (cuda_pathtracer_b200/distributed.py) under gloo, world_size 2 and 3, on the CPU.

The compute callables are the ORACLE here (this is a test; the product passes the B200 library's
functions, see bench.py and tests/test_distributed_gpu.py): what is being checked is the
sharding, the single exchange step of each path and that the assembled result is bit-identical
(scan / compact) or equal up to float re-association (render) to the single-process result."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cuda_pathtracer_b200 import distributed as D
from oracle_bindings import PtOracle, ScOracle, gen_array
from pt_cases import gui_from_flags, load_scene


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc_o = ScOracle()
        # ---- scan: contiguous shards, one all-gather of shard totals ---------------------------------
        a = gen_array(1 << 20, 50)[:n].copy()
        a[::7] = 2**30                                         # make the running sum wrap
        b, e = D.scan_shard(n, rank, world)
        local = a[b:e]

        def reduce_fn(x):
            return D.wrap_i32(int(x.astype(np.int64).sum()))

        def scan_fn(x, seed):
            return (sc_o.scan(x).astype(np.int64) + seed).astype(np.uint64).astype(np.uint32).view(np.int32)

        out, grand = D.sharded_scan(local, rank, world, dist, reduce_fn, scan_fn)
        np.save(os.path.join(out_dir, f"scan{rank}.npy"), out)
        # ---- compact: one all-gather of kept counts --------------------------------------------------
        c = gen_array(1 << 20, 4)[:n].copy()
        kept, off, total = D.sharded_compact(c[b:e], rank, world, dist, sc_o.compact_without_scan)
        np.save(os.path.join(out_dir, f"compact{rank}.npy"), kept)
        np.save(os.path.join(out_dir, f"meta{rank}.npy"), np.array([b, e, grand, off, total], dtype=np.int64))
        # ---- ... and the optional exchange that leaves the global output evenly sharded ------------------
        kept2, off2, total2, counts = D.sharded_compact_counts(c[b:e], rank, world, dist, sc_o.compact_without_scan)
        mine, my_b, total3 = D.rebalance_compacted(torch.from_numpy(kept2.copy()), counts, rank, world, dist,
                                                   lambda m: torch.empty(m, dtype=torch.int32))
        assert (off2, total2, total3) == (off, total, total)
        np.save(os.path.join(out_dir, f"rebal{rank}.npy"), mine.numpy())
        np.save(os.path.join(out_dir, f"rebalmeta{rank}.npy"), np.array([my_b], dtype=np.int64))
        # ---- render: iteration split, one all-reduce of the SUM image ----------------------------------
        scene, gui, pt_o = load_scene("cornell_tiny"), gui_from_flags([]), PtOracle()

        def render_fn(first, count, stride):
            img, _ = pt_o.render(scene, gui, first, count, stride=stride, threads=1)
            return torch.from_numpy(img)

        img = D.render_split(render_fn, dist.all_reduce, 1, 7, rank, world)
        if rank == 0:
            np.save(os.path.join(out_dir, "img.npy"), img.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, (1 << 20) - 3), (3, 100003), (2, 5)])
def test_sharded_paths_match_single_process(tmp_path, world, n):
    mp.spawn(_worker, args=(world, free_port(), n, str(tmp_path)), nprocs=world, join=True)
    a = gen_array(1 << 20, 50)[:n].copy()
    a[::7] = 2**30
    c = gen_array(1 << 20, 4)[:n].copy()
    sc_o = ScOracle()
    metas = [np.load(tmp_path / f"meta{r}.npy") for r in range(world)]
    # shards tile [0, n) in rank order
    assert metas[0][0] == 0 and metas[-1][1] == n and all(metas[r][1] == metas[r + 1][0] for r in range(world - 1))
    scan = np.concatenate([np.load(tmp_path / f"scan{r}.npy") for r in range(world)])
    assert np.array_equal(scan, sc_o.scan(a))                                   # bit-exact incl. wrap-around
    want_total = D.wrap_i32(int(a.astype(np.int64).sum()))
    assert all(m[2] == want_total for m in metas)
    want = sc_o.compact_without_scan(c)
    got = np.zeros(want.size, dtype=np.int32)
    for r in range(world):
        k = np.load(tmp_path / f"compact{r}.npy")
        got[metas[r][3]:metas[r][3] + k.size] = k
        assert metas[r][4] == want.size
    assert np.array_equal(got, want)
    # rebalanced: rank d holds exactly scan_shard(total, d, world) of the global output
    for r in range(world):
        ob, oe = D.scan_shard(want.size, r, world)
        assert int(np.load(tmp_path / f"rebalmeta{r}.npy")[0]) == ob
        assert np.array_equal(np.load(tmp_path / f"rebal{r}.npy"), want[ob:oe])
    scene, gui = load_scene("cornell_tiny"), gui_from_flags([])
    ref, _ = PtOracle().render(scene, gui, 1, 7, threads=1)
    assert np.allclose(np.load(tmp_path / "img.npy"), ref, rtol=2e-6, atol=1e-6)


def test_rebalance_plan_moves_every_survivor_exactly_once():
    rng = np.random.RandomState(3)
    for world in (1, 2, 3, 8):
        for _ in range(20):
            counts = [int(rng.randint(0, 50)) if rng.rand() < 0.8 else 0 for _ in range(world)]
            total = sum(counts)
            sent = np.zeros(total, dtype=np.int32)
            got = np.zeros(total, dtype=np.int32)
            offs = np.concatenate([[0], np.cumsum(counts)])
            for r in range(world):
                sends, recvs, my_b, my_e, t = D.rebalance_plan(counts, r, world)
                assert t == total and (my_b, my_e) == D.scan_shard(total, r, world)
                for d, b, e in sends:
                    sent[offs[r] + b:offs[r] + e] += 1
                    db, de = D.scan_shard(total, d, world)
                    assert db <= offs[r] + b and offs[r] + e <= de        # the piece lies inside d's shard
                for src, b, e in recvs:
                    got[my_b + b:my_b + e] += 1
                    assert offs[src] <= my_b + b and my_b + e <= offs[src + 1]
            assert (sent == 1).all() and (got == 1).all()


def test_iteration_shards_partition_the_iterations():
    for world in (1, 2, 3, 8):
        for n in (0, 1, 5, 8, 5000):
            seen = []
            for rank in range(world):
                first, count, stride = D.iteration_shard(11, n, rank, world)
                seen += [first + k * stride for k in range(count)]
            assert sorted(seen) == list(range(11, 11 + n))


def test_scan_shards_are_aligned_and_cover():
    for world in (1, 2, 4, 8):
        for n in (0, 1, 3, 4, 1023, (1 << 20) - 3, 1 << 28):
            edges = [D.scan_shard(n, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            for (b0, e0), (b1, e1) in zip(edges, edges[1:]):
                assert e0 == b1 and b1 % 4 == 0 or b1 == n
    pre, total = D.exclusive_prefix_i32([2**31 - 1, 1, -5])
    assert pre == [0, 2**31 - 1, -2**31] and total == D.wrap_i32(2**31 - 5)
