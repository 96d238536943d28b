"""Multi-GPU check + timing of the two sharded paths on N B200s of one node (one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29511 benchmarks/multi_gpu_check.py [--lg 28] [--scene cornell_small] [--iters 32]

* scan / compact: the array is sharded contiguously (cuda_pathtracer_b200.distributed.scan_shard);
  every rank reduces its shard (b200_reduce_i32_dev), ONE NCCL all-gather of the per-rank totals,
  the exclusive prefix of the totals becomes the device-side seed of the rank's single-pass scan.
  Everything is stream-ordered on the device (no host round trip inside the timed region).
  Checked bit for bit against torch.cumsum (int32 wrap) of the whole array.
* compact, rebalanced: the kept elements are then moved into evenly sized shards of the global output with
  point-to-point NCCL copies (cuda_pathtracer_b200.distributed.rebalance_compacted).
* render: iterations split with stride N, ONE NCCL all-reduce of the float SUM image; compared with
  the same iterations rendered on rank 0 alone.
Timing: CUDA events on the launching stream, max over ranks.  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

from cuda_pathtracer_b200 import distributed as D  # noqa: E402
from cuda_pathtracer_b200 import pathtracer as pt  # noqa: E402
from cuda_pathtracer_b200 import stream_compaction as sc  # noqa: E402


def timed(fn, iters, stream, world):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    times = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    times.sort()
    t = torch.tensor([times[len(times) // 2]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lg", type=int, default=28)
    ap.add_argument("--npot", type=int, default=3)
    ap.add_argument("--scene", default="cornell_small")
    ap.add_argument("--iters", type=int, default=32)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    out = {"n_gpus": world}

    # ---- sharded scan / compact -------------------------------------------------------------------
    n = (1 << args.lg) - args.npot
    g = torch.Generator(device="cuda")
    g.manual_seed(565)
    full = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g)
    full[::11] = 2 ** 30                                             # the running sum wraps
    b, e = D.scan_shard(n, rank, world)
    shard = full[b:e].clone()
    ctx = sc.Context()
    res = torch.empty_like(shard)
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    totals = torch.zeros(world, dtype=torch.int32, device="cuda")
    seed = torch.zeros(1, dtype=torch.int32, device="cuda")

    def sharded_scan():
        if world > 1:
            ctx.reduce(e - b, shard, total, stream=sptr)
            dist.all_gather_into_tensor(totals, total)
            # exclusive prefix of the totals before this rank, int32 wrap-around
            seed.copy_(totals[:rank].sum(dtype=torch.int32).reshape(1) if rank else torch.zeros_like(seed))
            ctx.scan(e - b, res, shard, sc.LOOKBACK, d_seed=seed, stream=sptr)
        else:
            ctx.scan(e - b, res, shard, sc.LOOKBACK, stream=sptr)

    ms = timed(sharded_scan, args.reps, stream, world)
    want = torch.cumsum(full, 0, dtype=torch.int32) - full           # exclusive, wraps like the reference
    ok = torch.equal(res, want[b:e])
    out["scan"] = {"n": n, "ms": ms, "gbs": 8.0 * n / ms / 1e6, "bit_exact": bool(ok)}
    del want

    flags = (full % 4).to(torch.int32)
    fshard = flags[b:e].clone()
    kept = torch.empty_like(fshard)
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    counts = torch.zeros(world, dtype=torch.int32, device="cuda")

    def sharded_compact():
        ctx.compact(e - b, kept, fshard, cnt, stream=sptr)
        if world > 1:
            dist.all_gather_into_tensor(counts, cnt)

    ms = timed(sharded_compact, args.reps, stream, world)
    k = int(cnt[0])
    want_c = flags[flags != 0]
    offs = counts.cpu().tolist() if world > 1 else [k]
    off = sum(offs[:rank])
    ok = torch.equal(kept[:k], want_c[off:off + k]) and sum(offs) == want_c.numel()
    out["compact"] = {"n": n, "ms": ms, "gbs": (4.0 * n + 4.0 * want_c.numel()) / ms / 1e6, "bit_exact": bool(ok),
                      "kept": int(want_c.numel())}
    # ... and, optionally, the survivors moved into evenly sized destination shards (NCCL send / recv over NVLink:
    # a source shard overlaps at most two destination shards)
    if world > 1:
        cl = counts.cpu().tolist()

        def rebalance():
            return D.rebalance_compacted(kept[:k], cl, rank, world, dist,
                                         lambda m: torch.empty(m, dtype=torch.int32, device="cuda"))

        mine, my_b, tot = rebalance()
        ob, oe = D.scan_shard(want_c.numel(), rank, world)
        ok_r = my_b == ob and tot == want_c.numel() and torch.equal(mine, want_c[ob:oe])
        ms_r = timed(rebalance, args.reps, stream, world)
        out["compact_rebalance"] = {"ms": ms_r, "bit_exact": bool(ok_r),
                                    "moved_gbs": 4.0 * want_c.numel() / ms_r / 1e6}
    del full, flags, want_c

    # ---- iteration-split render -----------------------------------------------------------------------
    from pt_cases import load_scene
    scene = load_scene(args.scene)
    r = pt.Renderer(scene)
    gui = pt.GuiDataContainer()
    r.set_options(gui)
    ptr, nfl = r.sum_device_ptr()

    class _Dev:
        def __init__(self, p, n_):
            self.__cuda_array_interface__ = {"shape": (n_,), "typestr": "<f4", "data": (p, False), "version": 2}

    img = torch.as_tensor(_Dev(ptr, nfl), device=f"cuda:{local}")

    def render_fn(first, count, stride):
        r.render(first, count, stride, stream=sptr, sync=False)
        return img

    got = D.render_split(render_fn, dist.all_reduce, 1, args.iters, rank, world).clone()
    torch.cuda.synchronize()
    r.reset()
    r.render(1, args.iters, 1)
    solo = torch.from_numpy(r.read_sum().reshape(-1)).cuda()
    out["render"] = {"scene": args.scene, "iters": args.iters,
                     "max_abs_diff_vs_single_gpu": float((got - solo).abs().max()),
                     "allclose": bool(torch.allclose(got, solo, rtol=2e-6, atol=1e-6))}
    r.close()
    oks = torch.tensor([int(out["scan"]["bit_exact"] and out["compact"]["bit_exact"] and out["render"]["allclose"]
                            and out.get("compact_rebalance", {"bit_exact": True})["bit_exact"])],
                       device="cuda")
    if world > 1:
        dist.all_reduce(oks, op=dist.ReduceOp.MIN)
    out["all_ranks_ok"] = bool(int(oks[0]))
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if out["all_ranks_ok"] else 2)


if __name__ == "__main__":
    main()
