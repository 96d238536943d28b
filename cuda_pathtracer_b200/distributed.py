"""Multi-GPU host logic (one process per GPU, torch.distributed): only where the paths shard.

* Path tracer -- iterations are independent (the seed depends on (iter, index, depth) only,
  path_tracer/src/pathtrace.cu:57-62), so rank g of N traces iterations first+g, first+g+N, ...
  on a replicated scene into a private accumulation image, and ONE all-reduce(sum) of the
  3*W*H floats closes the job.  No collective inside the loop.
* Scan -- contiguous shards; every rank reduces its shard to one int, ONE all-gather of those
  totals, and each rank scans its shard seeded with the exclusive prefix of the totals before it
  (int32 wrap-around adds are associative, so the result is bit-exact).
* Compact -- same shards; ONE all-gather of the per-shard kept counts gives every shard's offset
  in the global output.  Optionally (rebalance_compacted) the survivors are then moved into evenly
  sized destination shards of the global output: a source shard overlaps at most two of them when the
  input was sharded evenly, so the exchange is a handful of point-to-point copies over NVLink
  (NCCL send / recv), not an all-to-all.

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


def overlap(b0, e0, b1, e1):
    """[begin, end) of the intersection of two half-open ranges (empty: begin == end)"""
    b, e = max(b0, b1), min(e0, e1)
    return (b, e) if b < e else (b, b)


def rebalance_plan(counts, rank, world, align=4):
    """What rank `rank` sends and receives to turn 'survivors of shard r at global offset off[r]' into
    'rank d owns [scan_shard(total, d, world))' of the global output.
    Returns (sends, recvs, my_begin, my_end, total): sends = [(dst, local_begin, local_end)] into my kept
    array, recvs = [(src, out_begin, out_end)] into my destination slice (both in rank order, empty ones
    dropped, the local piece included as dst == src == rank)."""
    offsets, total = [], 0
    for c in counts:
        offsets.append(total)
        total += int(c)
    my_b, my_e = scan_shard(total, rank, world, align)
    src_b, src_e = offsets[rank], offsets[rank] + int(counts[rank])
    sends, recvs = [], []
    for d in range(world):
        db, de = scan_shard(total, d, world, align)
        b, e = overlap(src_b, src_e, db, de)
        if b < e:
            sends.append((d, b - src_b, e - src_b))
    for r in range(world):
        b, e = overlap(offsets[r], offsets[r] + int(counts[r]), my_b, my_e)
        if b < e:
            recvs.append((r, b - my_b, e - my_b))
    return sends, recvs, my_b, my_e, total


def rebalance_compacted(kept, counts, rank, world, dist, empty_like):
    """Moves the survivors into evenly sized destination shards (see rebalance_plan).  kept: this rank's
    survivors (torch tensor); counts: every rank's kept count (from sharded_compact's all-gather);
    empty_like(n) -> uninitialised tensor of n elements on kept's device.  Returns (my_slice, my_begin, total)."""
    sends, recvs, my_b, my_e, total = rebalance_plan(counts, rank, world)
    out = empty_like(my_e - my_b)
    ops = []
    for d, b, e in sends:
        if d == rank:
            continue
        ops.append(dist.P2POp(dist.isend, kept[b:e], d))
    for r, b, e in recvs:
        if r == rank:
            continue
        ops.append(dist.P2POp(dist.irecv, out[b:e], r))
    reqs = dist.batch_isend_irecv(ops) if ops else []
    for d, b, e in sends:                                  # the piece that stays on this rank
        if d == rank:
            (rb, re) = next((x[1], x[2]) for x in recvs if x[0] == rank)
            out[rb:re].copy_(kept[b:e])
    for q in reqs:
        q.wait()
    return out, my_b, total


def sharded_compact_counts(local_in, rank, world, dist, compact_fn, device="cpu"):
    """sharded_compact that also returns every rank's kept count (the input of rebalance_compacted)"""
    kept = compact_fn(local_in)
    counts = _all_gather_ints(dist, len(kept), world, device)
    offsets, total = exclusive_prefix_i32(counts)
    return kept, offsets[rank], total, counts


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
