"""Multi-GPU host logic (one process per GPU, torch.distributed): only where the paths shard.

* Path tracer -- iterations are independent (the seed depends on (iter, index, depth) only,
  path_tracer/src/pathtrace.cu:57-62), so rank g of N traces iterations first+g, first+g+N, ...
  on a replicated scene into a private accumulation image, and ONE all-reduce(sum) of the
  3*W*H floats closes the job.  No collective inside the loop.
* Scan -- contiguous shards; every rank reduces its shard to one int, ONE all-gather of those
  totals, and each rank scans its shard seeded with the exclusive prefix of the totals before it
  (int32 wrap-around adds are associative, so the result is bit-exact).
* Compact -- same shards; ONE all-gather of the per-shard kept counts gives every shard's offset
  in the global output.

The compute callables are injected, so the same logic runs on the B200 library (NCCL, device
tensors) and -- in the CPU test-suite -- under gloo with the oracle standing in for the kernels.
Nothing here imports the oracle.
"""
import numpy as np

INT32_MASK = 0xFFFFFFFF


def iteration_shard(first_iter, n_iters, rank, world):
    """(first, count, stride) of the iterations rank `rank` traces out of first_iter .. first_iter+n_iters-1"""
    if n_iters < 0 or world < 1 or not 0 <= rank < world:
        raise ValueError("bad shard request")
    count = (n_iters - rank + world - 1) // world if n_iters > rank else 0
    return first_iter + rank, count, world


def scan_shard(n, rank, world, align=4):
    """[begin, end) of rank's contiguous shard of n ints; boundaries are multiples of `align` ints
    (16 bytes) so that every shard keeps the 128-bit vector path"""
    if n < 0 or world < 1 or not 0 <= rank < world:
        raise ValueError("bad shard request")
    per = -(-n // world)
    per = -(-per // align) * align
    begin = min(n, rank * per)
    return begin, min(n, begin + per)


def wrap_i32(x):
    x &= INT32_MASK
    return x - (1 << 32) if x & 0x80000000 else x


def exclusive_prefix_i32(totals):
    """exclusive prefix of per-rank totals with int32 two's-complement wrap-around"""
    out, run = [], 0
    for t in totals:
        out.append(wrap_i32(run))
        run = (run + int(t)) & INT32_MASK
    return out, wrap_i32(run)


def _all_gather_ints(dist, value, world, device):
    import torch
    mine = torch.tensor([int(value)], dtype=torch.int32, device=device)
    if dist is None or world == 1:
        return [int(mine[0])]
    got = [torch.zeros(1, dtype=torch.int32, device=device) for _ in range(world)]
    dist.all_gather(got, mine)
    return [int(g[0]) for g in got]


def sharded_scan(local_in, rank, world, dist, reduce_fn, scan_fn, device="cpu"):
    """Exclusive scan of the concatenation of all ranks' `local_in`.
    reduce_fn(local_in) -> int32 total; scan_fn(local_in, seed) -> local exclusive scan + seed.
    Returns (local_out, grand_total)."""
    total = reduce_fn(local_in)
    totals = _all_gather_ints(dist, total, world, device)
    seeds, grand = exclusive_prefix_i32(totals)
    return scan_fn(local_in, seeds[rank]), grand


def sharded_compact(local_in, rank, world, dist, compact_fn, device="cpu"):
    """Order-preserving removal of zeros over all ranks' shards.
    compact_fn(local_in) -> kept elements of the shard.  Returns (local_kept, global_offset, global_count):
    local_kept belongs at [global_offset, global_offset + len) of the global output."""
    kept = compact_fn(local_in)
    counts = _all_gather_ints(dist, len(kept), world, device)
    offsets, total = exclusive_prefix_i32(counts)
    return kept, offsets[rank], total


def render_split(render_fn, allreduce_fn, first_iter, n_iters, rank, world):
    """render_fn(first, count, stride) accumulates into the rank's private image and returns it;
    allreduce_fn(image) sums it over ranks in place.  Returns the global SUM image."""
    first, count, stride = iteration_shard(first_iter, n_iters, rank, world)
    img = render_fn(first, count, stride)
    if world > 1:
        allreduce_fn(img)
    return img


def as_numpy_i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)
