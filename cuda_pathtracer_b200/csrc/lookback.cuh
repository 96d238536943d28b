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
// back over the predecessors' descriptors 32 * K at a time, publishes the inclusive prefix and
// returns the tile's exclusive prefix (seed included) in every lane.
//
// Polling discipline (measured on a B200 with benchmarks/sc_tune.cu's phase trace, profiles/):
// a descriptor that is not published yet ("X") is re-polled ONLY by the lane that owns it, and
// only if it is nearer than the nearest published prefix, with a __nanosleep back-off between
// polls.  Re-loading the whole window from every waiting warp turns the few L2 lines around the
// prefix frontier into a hot spot: hundreds of warps hammer one or two L2 slices, the polls --
// and the stores that would publish the awaited descriptors -- queue behind each other, and a
// single poll round trip grows from ~0.6 us to several us (the scan then runs at 45 % of HBM).
// STRIDE: distance between descriptors in 64-bit words (1 = packed; 4 = one 32-byte sector each).
template <int K, int BACKOFF_NS, int STRIDE>
__device__ __forceinline__ int lb_exclusive_prefix_cfg(unsigned long long* desc, unsigned tile,
                                                       unsigned epoch, int aggregate, int seed,
                                                       int lane)
{
    if (tile == 0) {
        if (lane == 0) lb_store(desc, lb_pack(epoch, LB_STATUS_P, seed + aggregate));
        return seed;
    }
    if (lane == 0) lb_store(desc + (size_t)tile * STRIDE, lb_pack(epoch, LB_STATUS_A, aggregate));

    int exclusive = 0;
    int look = (int)tile - 1;
    while (true) {
        // window slot k*32 + lane is tile look - (k*32 + lane); nearer tiles have lower k
        unsigned status[K];
        int value[K];
#pragma unroll
        for (int k = 0; k < K; ++k) {
            const int idx = look - (k * 32 + lane);
            const unsigned long long d = idx >= 0 ? lb_load(desc + (size_t)idx * STRIDE) : lb_pack(epoch, LB_STATUS_P, 0);
            const unsigned hi = (unsigned)(d >> 32);
            status[k] = ((hi >> 2) == epoch) ? (hi & 3u) : LB_STATUS_X;
            value[k] = (int)(unsigned)d;
        }
        int done_k = -1;                 // first k whose 32 descriptors hold a P with no X before it
        unsigned done_mask = 0;
#pragma unroll
        for (int k = 0; k < K; ++k) {
            if (done_k < 0) {            // uniform across the warp
                while (true) {
                    const unsigned x_mask = __ballot_sync(FULL_MASK, status[k] == LB_STATUS_X);
                    const unsigned p_mask = __ballot_sync(FULL_MASK, status[k] == LB_STATUS_P);
                    const unsigned below_p = p_mask ? ((p_mask & (0u - p_mask)) - 1u) : FULL_MASK;
                    const unsigned pending = x_mask & below_p;   // unpublished and nearer than any P
                    if (!pending) {
                        if (p_mask) { done_k = k; done_mask = p_mask; }
                        break;
                    }
                    if ((pending >> lane) & 1u) {
                        if (BACKOFF_NS > 0) __nanosleep(BACKOFF_NS);
                        const unsigned long long d = lb_load(desc + (size_t)(look - (k * 32 + lane)) * STRIDE);
                        const unsigned hi = (unsigned)(d >> 32);
                        status[k] = ((hi >> 2) == epoch) ? (hi & 3u) : LB_STATUS_X;
                        value[k] = (int)(unsigned)d;
                    }
                }
            }
        }
        int part = 0;
        if (done_k >= 0) {
            const int first = __ffs(done_mask) - 1;
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (k < done_k || (k == done_k && lane <= first)) part += value[k];
            exclusive += warp_sum(part);
            break;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) part += value[k];
        exclusive += warp_sum(part);
        look -= 32 * K;
    }
    if (lane == 0) lb_store(desc + (size_t)tile * STRIDE, lb_pack(epoch, LB_STATUS_P, exclusive + aggregate));
    return exclusive;
}

constexpr int LB_DEFAULT_BACKOFF_NS = 0;     // measured: back-off does not help once descriptors own a line

template <int K>
__device__ __forceinline__ int lb_exclusive_prefix_wide(unsigned long long* desc, unsigned tile,
                                                        unsigned epoch, int aggregate, int seed,
                                                        int lane)
{
    return lb_exclusive_prefix_cfg<K, LB_DEFAULT_BACKOFF_NS, 1>(desc, tile, epoch, aggregate, seed, lane);
}

__device__ __forceinline__ int lb_exclusive_prefix(unsigned long long* desc, unsigned tile,
                                                   unsigned epoch, int aggregate, int seed,
                                                   int lane)
{
    return lb_exclusive_prefix_cfg<1, LB_DEFAULT_BACKOFF_NS, 1>(desc, tile, epoch, aggregate, seed, lane);
}

}  // namespace b200
