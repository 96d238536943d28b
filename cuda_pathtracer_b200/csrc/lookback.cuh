// lookback.cuh -- decoupled look-back machinery shared by the int32 scan/compact kernels and the
// path-pool compaction of the wavefront renderer (sm_100a).
//
// One 64-bit descriptor per tile:  [ epoch:30 | status:2 | value:32 ].  A single 64-bit relaxed
// store publishes value and status together, so no fence is needed between them.  The epoch is
// bumped by the host on every launch that reuses the descriptor array, which makes stale words
// from an earlier launch read as "not ready" and removes the per-launch memset.
//
// Tile ids are handed out by an atomic ticket, so a tile's predecessors have always started:
// the spin in the look-back cannot deadlock whatever order the hardware schedules CTAs in.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace b200 {

constexpr unsigned LB_STATUS_X = 0u;  // nothing published yet
constexpr unsigned LB_STATUS_A = 1u;  // value = this tile's aggregate
constexpr unsigned LB_STATUS_P = 2u;  // value = inclusive prefix up to and including this tile
constexpr unsigned FULL_MASK = 0xffffffffu;

__device__ __forceinline__ unsigned long long lb_pack(unsigned epoch, unsigned status, int value)
{
    return ((unsigned long long)((epoch << 2) | status) << 32) | (unsigned)value;
}

__device__ __forceinline__ void lb_store(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned long long lb_load(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL_MASK, v, o);
    return v;
}

// inclusive scan across the 32 lanes of a warp
__device__ __forceinline__ int warp_inclusive_scan(int v, int lane)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(FULL_MASK, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// Called by ONE full warp of the CTA that owns `tile`.  Publishes the tile's aggregate, walks
// back over the predecessors' descriptors 32 at a time, publishes the inclusive prefix and
// returns the tile's exclusive prefix (seed included) in every lane.
__device__ __forceinline__ int lb_exclusive_prefix(unsigned long long* desc, unsigned tile,
                                                   unsigned epoch, int aggregate, int seed,
                                                   int lane)
{
    if (tile == 0) {
        if (lane == 0) lb_store(desc, lb_pack(epoch, LB_STATUS_P, seed + aggregate));
        return seed;
    }
    if (lane == 0) lb_store(desc + tile, lb_pack(epoch, LB_STATUS_A, aggregate));

    int exclusive = 0;
    int look = (int)tile - 1;
    while (true) {
        const int idx = look - lane;
        unsigned status = LB_STATUS_P;  // lanes that fall off the front: prefix 0 (never the
        int value = 0;                  // nearest P, because tile 0 always publishes P)
        bool waiting;
        do {
            waiting = false;
            if (idx >= 0) {
                const unsigned long long d = lb_load(desc + idx);
                const unsigned hi = (unsigned)(d >> 32);
                status = ((hi >> 2) == epoch) ? (hi & 3u) : LB_STATUS_X;
                value = (int)(unsigned)d;
                waiting = (status == LB_STATUS_X);
            }
        } while (__any_sync(FULL_MASK, waiting));

        const unsigned p_mask = __ballot_sync(FULL_MASK, status == LB_STATUS_P);
        if (p_mask) {
            const int first = __ffs(p_mask) - 1;
            exclusive += warp_sum(lane <= first ? value : 0);
            break;
        }
        exclusive += warp_sum(value);
        look -= 32;
    }
    if (lane == 0) lb_store(desc + tile, lb_pack(epoch, LB_STATUS_P, exclusive + aggregate));
    return exclusive;
}

}  // namespace b200
