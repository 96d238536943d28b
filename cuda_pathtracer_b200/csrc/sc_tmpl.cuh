// sc_tmpl.cuh -- kernel templates of the int32 scan / stream-compaction library (sm_100a).
//
// Shared by sc_kernels.cu (the library instantiates ONE tuned configuration per kernel) and by
// benchmarks/sc_tune.cu (which instantiates a grid of configurations to pick that one on a B200).
//
// Data layout inside a tile (THREADS * VEC * 4 ints): warp w owns 32 * VEC * 4 consecutive ints;
// its j-th 128-bit vector for lane l is vector index j*32 + l, so every warp-wide load/store
// instruction covers 512 contiguous bytes.  The order of elements is therefore (warp, j, lane,
// component); the scans below are carried out in exactly that order.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "lookback.cuh"

namespace b200 {
namespace sct {

__device__ __forceinline__ int4 ld_stream(const int4* p)
{
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ void st_stream(int4* p, const int4& v)
{
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
                 "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

template <int THREADS, int VEC>
struct Cfg {
    static constexpr int threads = THREADS;
    static constexpr int vec = VEC;
    static constexpr int items = VEC * 4;
    static constexpr int tile = THREADS * VEC * 4;
    static constexpr int warps = THREADS / 32;
};

// Loads this thread's VEC vectors of tile `tile`.  Out-of-range elements read as 0.
// VEC_OK: the pointer is 16-byte aligned (checked on the host).
template <class C, bool VEC_OK>
__device__ __forceinline__ void load_tile(const int* __restrict__ in, unsigned n, unsigned tile,
                                          int warp, int lane, int4 (&v)[C::vec])
{
    const unsigned long long warp_base = (unsigned long long)tile * C::tile + warp * (32 * C::items);
    const bool full = (tile + 1ull) * C::tile <= n;
    if (VEC_OK && full) {
        const int4* p = reinterpret_cast<const int4*>(in + warp_base) + lane;
#pragma unroll
        for (int j = 0; j < C::vec; ++j) v[j] = ld_stream(p + j * 32);
    } else {
#pragma unroll
        for (int j = 0; j < C::vec; ++j) {
            const unsigned long long e = warp_base + (j * 32 + lane) * 4;
            v[j].x = e + 0 < n ? __ldg(in + e + 0) : 0;
            v[j].y = e + 1 < n ? __ldg(in + e + 1) : 0;
            v[j].z = e + 2 < n ? __ldg(in + e + 2) : 0;
            v[j].w = e + 3 < n ? __ldg(in + e + 3) : 0;
        }
    }
}

template <class C, bool VEC_OK>
__device__ __forceinline__ void store_tile(int* __restrict__ out, unsigned n, unsigned tile,
                                           int warp, int lane, const int4 (&v)[C::vec])
{
    const unsigned long long warp_base = (unsigned long long)tile * C::tile + warp * (32 * C::items);
    const bool full = (tile + 1ull) * C::tile <= n;
    if (VEC_OK && full) {
        int4* p = reinterpret_cast<int4*>(out + warp_base) + lane;
#pragma unroll
        for (int j = 0; j < C::vec; ++j) st_stream(p + j * 32, v[j]);
    } else {
#pragma unroll
        for (int j = 0; j < C::vec; ++j) {
            const unsigned long long e = warp_base + (j * 32 + lane) * 4;
            if (e + 0 < n) out[e + 0] = v[j].x;
            if (e + 1 < n) out[e + 1] = v[j].y;
            if (e + 2 < n) out[e + 2] = v[j].z;
            if (e + 3 < n) out[e + 3] = v[j].w;
        }
    }
}

// Phase 1 of the tile scan: this warp's total goes to shared memory (the tile aggregate is known
// to everybody after ONE barrier, before any intra-warp scan has run, so the look-back warp can
// publish it as early as possible).
template <class C>
__device__ __forceinline__ void tile_warp_totals(const int (&s)[C::vec], int* smem_warp, int warp, int lane)
{
    int t = 0;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) t += s[j];
    t = warp_sum(t);
    if (lane == 0) smem_warp[warp] = t;
}

// Phase 2 (after the barrier): excl[j] = sum of everything in the tile that precedes vector j of
// this thread; returns the tile aggregate.
template <class C>
__device__ __forceinline__ int tile_exclusive_from_totals(const int (&s)[C::vec], int (&excl)[C::vec],
                                                          const int* smem_warp, int warp, int lane)
{
    int warp_prefix = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < C::warps; ++w) {
        const int t = smem_warp[w];
        if (w < warp) warp_prefix += t;
        aggregate += t;
    }
    int carry = warp_prefix;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) {
        const int inc = warp_inclusive_scan(s[j], lane);
        excl[j] = carry + inc - s[j];
        carry += __shfl_sync(FULL_MASK, inc, 31);
    }
    return aggregate;
}

__device__ __forceinline__ int4 scan_vec(const int4& v, int base)
{
    int4 o;
    o.x = base;
    o.y = o.x + v.x;
    o.z = o.y + v.y;
    o.w = o.z + v.z;
    return o;
}

template <class C>
struct ScanSmem {
    int warp_sums[C::warps];
    int tile_exclusive;
    unsigned tile;
};

// Tile ids.  TICKET: handed out by an atomic counter in start order (tickets[epoch&1] is ours, the
// other one is re-armed for the next launch; launches sharing a workspace are stream-ordered), so
// a tile's predecessors have always started whatever order the hardware dispatches CTAs in.
// !TICKET: blockIdx.x (CTAs of a 1-D grid are dispatched in index order on every CUDA GPU to
// date -- the assumption CUB's single-pass scan makes as well); saves one L2 round trip per tile.
template <bool TICKET>
__device__ __forceinline__ unsigned tile_id(unsigned* tickets, unsigned epoch, unsigned* smem_tile)
{
    if (!TICKET) return blockIdx.x;
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(&tickets[epoch & 1u], 1u);
        if (t == 0) tickets[(epoch + 1u) & 1u] = 0u;
        *smem_tile = t;
    }
    __syncthreads();
    return *smem_tile;
}

// ------------------------------------------------------------------------------------------------
// k_scan_lookback: one tile per CTA, single pass, 8 B/element.
// ------------------------------------------------------------------------------------------------
template <class C, bool VEC_OK, int LBW, bool TICKET, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
__global__ void __launch_bounds__(C::threads)
k_scan_lookback(const int* __restrict__ in, int* __restrict__ out, unsigned n,
                unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets,
                unsigned epoch, const int* __restrict__ d_seed, int* __restrict__ d_total)
{
    __shared__ ScanSmem<C> sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = tile_id<TICKET>(tickets, epoch, &sm.tile);

    int4 v[C::vec];
    load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);
    int s[C::vec], excl[C::vec];
#pragma unroll
    for (int j = 0; j < C::vec; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
    tile_warp_totals<C>(s, sm.warp_sums, warp, lane);
    __syncthreads();
    if (warp == 0) {                 // publish + look back first; the other warps scan meanwhile
        int aggregate = 0;
#pragma unroll
        for (int w = 0; w < C::warps; ++w) aggregate += sm.warp_sums[w];
        const int seed = d_seed ? __ldg(d_seed) : 0;
        const int e = lb_exclusive_prefix_cfg<LBW, BACKOFF, STRIDE>(desc, tile, epoch, aggregate, seed, lane);
        if (lane == 0) {
            sm.tile_exclusive = e;
            if (d_total && tile == gridDim.x - 1) *d_total = e + aggregate;
        }
    }
    tile_exclusive_from_totals<C>(s, excl, sm.warp_sums, warp, lane);
    __syncthreads();
    const int tile_exclusive = sm.tile_exclusive;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) v[j] = scan_vec(v[j], tile_exclusive + excl[j]);
    store_tile<C, VEC_OK>(out, n, tile, warp, lane, v);
}

// ------------------------------------------------------------------------------------------------
// k_compact: flag + scan + scatter in one pass.  Kept elements of a tile are first packed in
// shared memory in order, then written as one contiguous run; the run is written with 128-bit
// stores from the first 16-byte-aligned destination element on.
// ------------------------------------------------------------------------------------------------
// dynamic shared memory: C::tile ints (the packed survivors of the tile)
template <class C>
constexpr size_t compact_smem_bytes() { return (size_t)C::tile * sizeof(int); }

template <class C, bool VEC_OK, int LBW, bool TICKET, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
__global__ void __launch_bounds__(C::threads)
k_compact(const int* __restrict__ in, int* __restrict__ out, unsigned n,
          unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets, unsigned epoch,
          int* __restrict__ d_count)
{
    extern __shared__ __align__(16) int packed[];
    __shared__ ScanSmem<C> sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = tile_id<TICKET>(tickets, epoch, &sm.tile);

    int4 v[C::vec];
    load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);   // out-of-range reads as 0 == dropped
    int s[C::vec], excl[C::vec];
#pragma unroll
    for (int j = 0; j < C::vec; ++j)
        s[j] = (v[j].x != 0) + (v[j].y != 0) + (v[j].z != 0) + (v[j].w != 0);
    tile_warp_totals<C>(s, sm.warp_sums, warp, lane);
    __syncthreads();
    int kept = 0;
#pragma unroll
    for (int w = 0; w < C::warps; ++w) kept += sm.warp_sums[w];
    if (warp == 0) {
        const int e = lb_exclusive_prefix_cfg<LBW, BACKOFF, STRIDE>(desc, tile, epoch, kept, 0, lane);
        if (lane == 0) {
            sm.tile_exclusive = e;
            if (tile == gridDim.x - 1) *d_count = e + kept;
        }
    }
    tile_exclusive_from_totals<C>(s, excl, sm.warp_sums, warp, lane);
#pragma unroll
    for (int j = 0; j < C::vec; ++j) {
        int p = excl[j];
        if (v[j].x != 0) packed[p++] = v[j].x;
        if (v[j].y != 0) packed[p++] = v[j].y;
        if (v[j].z != 0) packed[p++] = v[j].z;
        if (v[j].w != 0) packed[p++] = v[j].w;
    }
    __syncthreads();
    const int base = sm.tile_exclusive;
    int* dst = out + base;
    // head: up to 3 elements until dst + head is 16-byte aligned; body: int4; tail: scalars.
    // Shared-memory reads of the body are misaligned by `head` ints, so they are done as scalars.
    const int mis = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
    int head = mis ? 4 - mis : 0;
    if (head > kept) head = kept;
    if ((int)threadIdx.x < head) dst[threadIdx.x] = packed[threadIdx.x];
    const int body4 = (kept - head) >> 2;
    int4* dst4 = reinterpret_cast<int4*>(dst + head);
    for (int q = threadIdx.x; q < body4; q += C::threads) {
        const int* src = packed + head + 4 * q;
        st_stream(dst4 + q, make_int4(src[0], src[1], src[2], src[3]));
    }
    const int done = head + 4 * body4;
    if ((int)threadIdx.x < kept - done) dst[done + threadIdx.x] = packed[done + threadIdx.x];
}

// ------------------------------------------------------------------------------------------------
// k_compact_ws: the single-pass compaction with the look-back OFF THE CRITICAL PATH (warp-specialised, persistent) --
// the scheme of the renderer's bounce kernel (pt_kernels.cu: k_bounce).  In k_compact every tile's warps stand at a barrier
// while warp 0 looks back over its predecessors (ncu: that barrier is the largest stall, 8 warps per issue slot).  Here
//   * a CTA is C::warps compute warps + 1 look-back warp and keeps drawing tiles from an atomic ticket (so a tile's
//     predecessors are always held by running CTAs: no deadlock whatever the residency);
//   * the compute warps of tile k count, hand {tile, kept} to the look-back warp through a two-entry mbarrier mailbox, pack
//     the survivors into one of TWO shared-memory buffers and go straight on;
//   * one tile later they pick the exclusive prefix up and write the packed run of tile k out (128-bit stores from the first
//     aligned destination element on, as in k_compact) -- the look-back has had a whole tile's load / count / pack to finish.
// One named barrier per tile (the warp counts; it also orders the two buffers' reuse).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned ws_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ws_mbar_init(unsigned long long* bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(ws_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void ws_mbar_arrive(unsigned long long* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(ws_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void ws_mbar_wait(unsigned long long* bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WS_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WS_DONE_%=;\n"
        "bra WS_WAIT_%=;\n"
        "WS_DONE_%=:\n"
        "}\n" ::"r"(ws_smem_u32(bar)),
        "r"(parity)
        : "memory");
}

template <class C>
struct WsShared {
    unsigned long long req_bar[2], resp_bar[2];
    int req[2][2];                 // {tile (-1: stop), kept}
    int resp[2];                   // exclusive prefix of the tile with this parity
    unsigned tile[2];              // the CTA's tile of this parity (drawn by thread 0 one iteration ahead)
    int warp_sums[2][C::warps];
};

template <class C>
constexpr size_t compact_ws_smem_bytes() { return (size_t)2 * C::tile * sizeof(int); }

// the packed survivors `packed[0 .. kept)` of one tile -> out[base ..): head scalars, aligned 128-bit body, tail scalars
template <int THREADS>
__device__ __forceinline__ void write_run(int* __restrict__ out, const int* packed, int base, int kept)
{
    int* dst = out + base;
    const int mis = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
    int head = mis ? 4 - mis : 0;
    if (head > kept) head = kept;
    if ((int)threadIdx.x < head) dst[threadIdx.x] = packed[threadIdx.x];
    const int body4 = (kept - head) >> 2;
    int4* dst4 = reinterpret_cast<int4*>(dst + head);
    for (int q = threadIdx.x; q < body4; q += THREADS) {
        const int* src = packed + head + 4 * q;
        st_stream(dst4 + q, make_int4(src[0], src[1], src[2], src[3]));
    }
    const int done = head + 4 * body4;
    if ((int)threadIdx.x < kept - done) dst[done + threadIdx.x] = packed[done + threadIdx.x];
}

template <class C, bool VEC_OK, int LBW, int STRIDE>
__global__ void __launch_bounds__(C::threads + 32)
k_compact_ws(const int* __restrict__ in, int* __restrict__ out, unsigned n, unsigned long long* __restrict__ desc,
             unsigned* __restrict__ tickets, unsigned epoch, int* __restrict__ d_count)
{
    extern __shared__ __align__(16) int packed2[];            // [2][C::tile]
    __shared__ WsShared<C> sh;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    if (threadIdx.x == 0) {
        ws_mbar_init(&sh.req_bar[0], 1); ws_mbar_init(&sh.req_bar[1], 1);
        ws_mbar_init(&sh.resp_bar[0], 1); ws_mbar_init(&sh.resp_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sh.tile[0] = atomicAdd(&tickets[2], 1u);              // tickets[2]: next tile, tickets[3]: CTAs done (both zero between launches)
    }
    __syncthreads();

    if (warp == C::warps) {                                    // ---- the look-back warp ----
        for (int it = 0;; ++it) {
            ws_mbar_wait(&sh.req_bar[it & 1], (unsigned)((it >> 1) & 1));
            const int tile = sh.req[it & 1][0], kept = sh.req[it & 1][1];
            if (tile < 0) break;
            const int e = lb_exclusive_prefix_cfg<LBW, LB_DEFAULT_BACKOFF_NS, STRIDE>(desc, (unsigned)tile, epoch, kept, 0, lane);
            __syncwarp();                                      // every lane has read the request before it can be answered
            if (lane == 0) {
                if ((unsigned)tile == tiles - 1) *d_count = e + kept;
                sh.resp[it & 1] = e;
                ws_mbar_arrive(&sh.resp_bar[it & 1]);
            }
        }
        return;
    }

    // ---- the compute warps ----
    int prev_kept = 0, it = 0;
    for (;; ++it) {
        const unsigned tile = sh.tile[it & 1];                 // written before the previous iteration's barrier
        if (tile >= tiles) break;                              // uniform over the CTA
        int4 v[C::vec];
        load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);      // out-of-range reads as 0 == dropped
        int s[C::vec], excl[C::vec];
#pragma unroll
        for (int j = 0; j < C::vec; ++j)
            s[j] = (v[j].x != 0) + (v[j].y != 0) + (v[j].z != 0) + (v[j].w != 0);
        tile_warp_totals<C>(s, sh.warp_sums[it & 1], warp, lane);
        if (threadIdx.x == 0) sh.tile[(it + 1) & 1] = atomicAdd(&tickets[2], 1u);
        asm volatile("bar.sync 1, %0;" ::"n"(C::threads) : "memory");
        const int kept = tile_exclusive_from_totals<C>(s, excl, sh.warp_sums[it & 1], warp, lane);
        if (threadIdx.x == 0) {
            sh.req[it & 1][0] = (int)tile;
            sh.req[it & 1][1] = kept;
            ws_mbar_arrive(&sh.req_bar[it & 1]);
        }
        if (it > 0) {                                          // the previous tile's run, now that its prefix is known
            ws_mbar_wait(&sh.resp_bar[(it - 1) & 1], (unsigned)(((it - 1) >> 1) & 1));
            write_run<C::threads>(out, packed2 + ((it - 1) & 1) * C::tile, sh.resp[(it - 1) & 1], prev_kept);
        }
        int* packed = packed2 + (it & 1) * C::tile;
#pragma unroll
        for (int j = 0; j < C::vec; ++j) {
            int q = excl[j];
            if (v[j].x != 0) packed[q++] = v[j].x;
            if (v[j].y != 0) packed[q++] = v[j].y;
            if (v[j].z != 0) packed[q++] = v[j].z;
            if (v[j].w != 0) packed[q++] = v[j].w;
        }
        prev_kept = kept;
    }
    asm volatile("bar.sync 1, %0;" ::"n"(C::threads) : "memory");      // the last tile's buffer is complete
    if (it > 0) {
        ws_mbar_wait(&sh.resp_bar[(it - 1) & 1], (unsigned)(((it - 1) >> 1) & 1));
        write_run<C::threads>(out, packed2 + ((it - 1) & 1) * C::tile, sh.resp[(it - 1) & 1], prev_kept);
    }
    if (threadIdx.x == 0) {                                    // tell the look-back warp to leave
        sh.req[it & 1][0] = -1;
        ws_mbar_arrive(&sh.req_bar[it & 1]);
        // the last CTA to get here (every CTA has drawn its last ticket by now) re-arms the counters for the next launch
        if (atomicAdd(&tickets[3], 1u) == gridDim.x - 1) { tickets[2] = 0u; tickets[3] = 0u; }
    }
}

// ------------------------------------------------------------------------------------------------
// Work-efficient three-phase scan: k_reduce_tiles (up-sweep), recursion on the tile sums,
// k_downsweep (per-tile scan seeded with the scanned tile sum).  12 B/element.
// ------------------------------------------------------------------------------------------------
template <class C, bool VEC_OK>
__global__ void __launch_bounds__(C::threads)
k_reduce_tiles(const int* __restrict__ in, unsigned n, int* __restrict__ tile_sums)
{
    __shared__ int warp_sums[C::warps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int4 v[C::vec];
    load_tile<C, VEC_OK>(in, n, blockIdx.x, warp, lane, v);
    int s = 0;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) s += v[j].x + v[j].y + v[j].z + v[j].w;
    s = warp_sum(s);
    if (lane == 0) warp_sums[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < C::warps; ++w) t += warp_sums[w];
        tile_sums[blockIdx.x] = t;
    }
}

template <class C, bool VEC_OK>
__global__ void __launch_bounds__(C::threads)
k_downsweep(const int* __restrict__ in, int* __restrict__ out, unsigned n,
            const int* __restrict__ tile_offsets, const int* __restrict__ d_seed,
            int* __restrict__ d_total)
{
    __shared__ int warp_sums[C::warps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = blockIdx.x;
    int4 v[C::vec];
    load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);
    int s[C::vec], excl[C::vec];
#pragma unroll
    for (int j = 0; j < C::vec; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
    tile_warp_totals<C>(s, warp_sums, warp, lane);
    __syncthreads();
    const int aggregate = tile_exclusive_from_totals<C>(s, excl, warp_sums, warp, lane);
    int offset = tile_offsets ? __ldg(tile_offsets + tile) : 0;
    if (d_seed) offset += __ldg(d_seed);
    if (d_total && threadIdx.x == 0 && tile == gridDim.x - 1) *d_total = offset + aggregate;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) v[j] = scan_vec(v[j], offset + excl[j]);
    store_tile<C, VEC_OK>(out, n, tile, warp, lane, v);
}

// ------------------------------------------------------------------------------------------------
// Whole-array sum (the per-shard total a multi-GPU scan exchanges): persistent grid-stride loop.
// ------------------------------------------------------------------------------------------------
template <class C, bool VEC_OK>
__global__ void __launch_bounds__(C::threads)
k_reduce_total(const int* __restrict__ in, unsigned n, int* __restrict__ d_total)
{
    __shared__ int warp_sums[C::warps];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int acc = 0;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    for (unsigned tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int4 v[C::vec];
        load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);
#pragma unroll
        for (int j = 0; j < C::vec; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    acc = warp_sum(acc);
    if (lane == 0) warp_sums[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < C::warps; ++w) t += warp_sums[w];
        atomicAdd(d_total, t);
    }
}

}  // namespace sct
}  // namespace b200
