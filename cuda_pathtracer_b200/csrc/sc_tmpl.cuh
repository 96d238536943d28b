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
// k_scan_wave: the single-pass scan for arrays small enough that ONE wave of CTAs (one per SM) holds all of them in
// registers (n <= #SMs * 28672 ints: up to 2^22 on a B200).  With at most ~148 tiles there is nothing to walk back over:
// every CTA publishes its aggregate, and thread t of CTA b reads the aggregate of CTA t < b -- all predecessors in ONE
// round trip, however many there are -- a block sum gives the exclusive prefix.  (k_scan_lookback at 2^22 is 1024 tiles whose
// 32-descriptor windows and the prefix frontier meet after ~16 dependent round trips: the ramp, not the bandwidth, is
// what a small scan costs.)  CTAs are dispatched in blockIdx order and wait only on lower ones.
// ------------------------------------------------------------------------------------------------
template <class C, bool VEC_OK, int STRIDE>
__global__ void __launch_bounds__(C::threads, 1)
k_scan_wave(const int* __restrict__ in, int* __restrict__ out, unsigned n, unsigned long long* __restrict__ desc,
            unsigned epoch, const int* __restrict__ d_seed, int* __restrict__ d_total)
{
    __shared__ int warp_sums[C::warps];
    __shared__ int pred_sums[C::warps];
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
    if (threadIdx.x == 0) lb_store(desc + (size_t)tile * STRIDE, lb_pack(epoch, LB_STATUS_A, aggregate));
    int part = 0;
    if ((unsigned)(warp * 32) < tile) {                          // (uniform per warp)
        for (unsigned t = threadIdx.x; t < tile; t += C::threads) {
            unsigned long long d;
            do { d = lb_load(desc + (size_t)t * STRIDE); } while ((unsigned)(d >> 34) != epoch);
            part += (int)(unsigned)d;
        }
        part = warp_sum(part);
    }
    if (lane == 0) pred_sums[warp] = part;
    __syncthreads();
    int tile_exclusive = d_seed ? __ldg(d_seed) : 0;
    const int pred_warps = (int)((tile + 31u) / 32u) < C::warps ? (int)((tile + 31u) / 32u) : C::warps;
    for (int w = 0; w < pred_warps; ++w) tile_exclusive += pred_sums[w];
    if (d_total && threadIdx.x == 0 && tile == gridDim.x - 1) *d_total = tile_exclusive + aggregate;
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
// k_scan_tma: the single-pass scan as a persistent TMA ring (one CTA per SM).
//
// In k_scan_lookback a tile's loads are issued when its CTA starts and the CTA then stands in the look-back before it can
// store: the bytes in flight per SM follow the look-back latency.  Here the data never waits in registers:
//   * ONE thread (the DMA thread) draws tiles from an atomic ticket and moves them with 1-D bulk copies
//     (cp.async.bulk, SASS UBLKCP): global -> a ring of STAGES shared-memory buffers, completion on an mbarrier, and -- once
//     a buffer holds the finished scan -- shared -> global as a bulk store.  Loads run STAGES - DELAY - 2 tiles ahead of
//     the arithmetic whatever the look-back does;
//   * CWARPS compute warps make two passes over a buffer: A (when the tile lands) sums it -- the per-warp totals go to the
//     look-back warp through an mbarrier -- and B (DELAY tiles later, when the tile's exclusive prefix has come back)
//     scans it in place; then fence.proxy.async + an mbarrier hand the buffer to the DMA thread;
//   * NLB look-back warps (local tile i belongs to warp i % NLB) publish the aggregate, look back, publish the prefix.
// No CTA-wide barrier after set-up: four mbarriers per stage (full -> req -> resp -> store) in a strict cycle, so the phase
// parity of use u of a stage is u & 1 for all four.  Tiles are in ticket order and a CTA's tiles ascend, so the smallest
// tile without a prefix only ever waits on published tiles: no deadlock whatever the residency.
// Needs 16-byte aligned pointers (bulk copies); the ragged end of the array (n % 4 ints) is moved by the DMA thread itself.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ws_mbar_expect_tx(unsigned long long* bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(ws_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ws_bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(ws_smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(ws_smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void ws_bulk_s2g(void* dst, const void* src, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(ws_smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ws_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void ws_bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void ws_bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void ws_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr unsigned TMA_NO_TILE = 0xffffffffu;

template <int STAGES, int CWARPS>
struct TmaScanShared {
    unsigned long long full_bar[STAGES], req_bar[STAGES], resp_bar[STAGES], store_bar[STAGES];
    unsigned tile[STAGES];         // the tile in this stage (TMA_NO_TILE: no more tiles)
    unsigned stop_at;              // local index of the first tile that does not exist (written before its full_bar arrive)
    int resp[STAGES];              // exclusive prefix of that tile
    int warp_sums[STAGES][CWARPS];
};

template <int STAGE_INTS, int STAGES>
constexpr size_t scan_tma_smem_bytes() { return (size_t)STAGES * STAGE_INTS * sizeof(int); }

template <int STAGE_INTS, int CWARPS, int NLB>
constexpr int scan_tma_threads() { return (CWARPS + 1 + NLB) * 32; }

template <int STAGE_INTS, int STAGES, int CWARPS, int DELAY, int NLB, int LBW, int STRIDE>
__global__ void __launch_bounds__((CWARPS + 1 + NLB) * 32, 1)
k_scan_tma(const int* in, int* out, unsigned n, unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets,
           unsigned epoch, const int* __restrict__ d_seed, int* __restrict__ d_total)
{
    static_assert(STAGE_INTS % (CWARPS * 128) == 0, "a stage is a whole number of 128-bit vectors per lane");
    static_assert(DELAY + 2 <= STAGES && NLB <= STAGES, "ring too short");
    constexpr int VPL = STAGE_INTS / (CWARPS * 128);           // 128-bit vectors per lane per tile
    constexpr int WARP_INTS = 32 * VPL * 4;
    extern __shared__ __align__(128) int ring[];                // [STAGES][STAGE_INTS]
    __shared__ TmaScanShared<STAGES, CWARPS> sh;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(((unsigned long long)n + STAGE_INTS - 1) / STAGE_INTS);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ws_mbar_init(&sh.full_bar[s], 1); ws_mbar_init(&sh.req_bar[s], CWARPS);
            ws_mbar_init(&sh.resp_bar[s], 1); ws_mbar_init(&sh.store_bar[s], CWARPS);
        }
        sh.stop_at = 0xffffffffu;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == CWARPS) {                                       // ---- the DMA thread ----
        if (lane != 0) return;
        unsigned issued = 0, stored = 0;
        bool more = true;
        unsigned next_t = atomicAdd(&tickets[2], 1u);           // tickets[2]: next tile, tickets[3]: CTAs done (zero between launches)
        for (;;) {
            if (more && issued - stored < (unsigned)STAGES) {   // a free stage: the next tile's load
                const unsigned t = next_t;
                const int st = (int)(issued % STAGES);
                if (t >= tiles) {
                    sh.tile[st] = TMA_NO_TILE;
                    sh.stop_at = issued;
                    ws_mbar_arrive(&sh.full_bar[st]);
                    more = false;
                    continue;
                }
                next_t = atomicAdd(&tickets[2], 1u);            // (one ahead: its round trip hides behind the waits below)
                sh.tile[st] = t;
                const unsigned long long base = (unsigned long long)t * STAGE_INTS;
                const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
                int* buf = ring + (size_t)st * STAGE_INTS;
                const unsigned whole = valid & ~3u;
                for (unsigned k = whole; k < valid; ++k) buf[k] = in[base + k];
                if (whole) {
                    ws_mbar_expect_tx(&sh.full_bar[st], whole * 4u);
                    ws_bulk_g2s(buf, in + base, whole * 4u, &sh.full_bar[st]);
                } else {
                    ws_mbar_arrive(&sh.full_bar[st]);
                }
                ++issued;
                continue;
            }
            if (stored < issued) {                              // the oldest tile: store it once its scan is in the buffer
                const int st = (int)(stored % STAGES);
                ws_mbar_wait(&sh.store_bar[st], (stored / STAGES) & 1u);
                const unsigned t = sh.tile[st];
                const unsigned long long base = (unsigned long long)t * STAGE_INTS;
                const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
                const int* buf = ring + (size_t)st * STAGE_INTS;
                const unsigned whole = valid & ~3u;
                if (whole) ws_bulk_s2g(out + base, buf, whole * 4u);
                ws_bulk_commit();
                for (unsigned k = whole; k < valid; ++k) out[base + k] = buf[k];
                ws_bulk_wait_read0();                           // the buffer may be overwritten
                ++stored;
                continue;
            }
            break;
        }
        ws_bulk_wait0();
        // the last CTA to get here (every CTA has drawn its last ticket by now) re-arms the counters for the next launch
        if (atomicAdd(&tickets[3], 1u) == gridDim.x - 1) { tickets[2] = 0u; tickets[3] = 0u; }
        return;
    }

    if (warp > CWARPS) {                                        // ---- the look-back warps ----
        const int seed = d_seed ? __ldg(d_seed) : 0;
        for (unsigned i = (unsigned)(warp - CWARPS - 1);; i += NLB) {
            const int st = (int)(i % STAGES);
            ws_mbar_wait(&sh.req_bar[st], (i / STAGES) & 1u);
            if (i >= *(volatile unsigned*)&sh.stop_at) break;
            const unsigned t = sh.tile[st];
            int aggregate = 0;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) aggregate += sh.warp_sums[st][w];
            const int e = lb_exclusive_prefix_cfg<LBW, LB_DEFAULT_BACKOFF_NS, STRIDE>(desc, t, epoch, aggregate, seed, lane);
            __syncwarp();
            if (lane == 0) {
                if (d_total && t == tiles - 1) *d_total = e + aggregate;
                sh.resp[st] = e;
                ws_mbar_arrive(&sh.resp_bar[st]);
            }
        }
        return;
    }

    // ---- the compute warps ----
    unsigned done_at = 0xffffffffu;
    for (unsigned i = 0;; ++i) {
        if (i < done_at) {                                      // pass A over local tile i: its sum
            const int st = (int)(i % STAGES);
            ws_mbar_wait(&sh.full_bar[st], (i / STAGES) & 1u);
            const unsigned t = sh.tile[st];
            if (t == TMA_NO_TILE) {
                done_at = i;
                if (lane == 0) ws_mbar_arrive(&sh.req_bar[st]);  // the look-back warp of local tile i sees stop_at and leaves
            } else {
                const int4* p = reinterpret_cast<const int4*>(ring + (size_t)st * STAGE_INTS + warp * WARP_INTS) + lane;
                const unsigned long long base = (unsigned long long)t * STAGE_INTS;
                const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
                int s = 0;
                if (valid == (unsigned)STAGE_INTS) {
#pragma unroll
                    for (int j = 0; j < VPL; ++j) { const int4 v = p[j * 32]; s += (v.x + v.y) + (v.z + v.w); }
                } else {
#pragma unroll
                    for (int j = 0; j < VPL; ++j) {
                        const int4 v = p[j * 32];
                        const unsigned e = (unsigned)(warp * WARP_INTS + (j * 32 + lane) * 4);
                        s += (e + 0 < valid ? v.x : 0) + (e + 1 < valid ? v.y : 0) + (e + 2 < valid ? v.z : 0) + (e + 3 < valid ? v.w : 0);
                    }
                }
                s = warp_sum(s);
                if (lane == 0) {
                    sh.warp_sums[st][warp] = s;
                    ws_mbar_arrive(&sh.req_bar[st]);
                }
            }
        }
        if (i >= (unsigned)DELAY) {                             // pass B over local tile i - DELAY: the scan, in place
            const unsigned k = i - DELAY;
            if (k >= done_at) break;
            const int st = (int)(k % STAGES);
            ws_mbar_wait(&sh.resp_bar[st], (k / STAGES) & 1u);
            int carry = sh.resp[st];
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) if (w < warp) carry += sh.warp_sums[st][w];
            int4* p = reinterpret_cast<int4*>(ring + (size_t)st * STAGE_INTS + warp * WARP_INTS) + lane;
            const unsigned t = sh.tile[st];
            const unsigned long long base = (unsigned long long)t * STAGE_INTS;
            const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
            int4 v[VPL];
            int s[VPL];
#pragma unroll
            for (int j = 0; j < VPL; ++j) v[j] = p[j * 32];
            if (valid != (unsigned)STAGE_INTS) {
#pragma unroll
                for (int j = 0; j < VPL; ++j) {
                    const unsigned e = (unsigned)(warp * WARP_INTS + (j * 32 + lane) * 4);
                    if (e + 0 >= valid) v[j].x = 0;
                    if (e + 1 >= valid) v[j].y = 0;
                    if (e + 2 >= valid) v[j].z = 0;
                    if (e + 3 >= valid) v[j].w = 0;
                }
            }
#pragma unroll
            for (int j = 0; j < VPL; ++j) s[j] = (v[j].x + v[j].y) + (v[j].z + v[j].w);
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int inc = warp_inclusive_scan(s[j], lane);
                p[j * 32] = scan_vec(v[j], carry + inc - s[j]);
                carry += __shfl_sync(FULL_MASK, inc, 31);
            }
            ws_fence_async_smem();                              // generic-proxy writes -> visible to the bulk store
            __syncwarp();
            if (lane == 0) ws_mbar_arrive(&sh.store_bar[st]);
        }
    }
    if (NLB > 1) {      // the other look-back warps wait for local tiles done_at + 1 ..: wake them too.  Every compute warp must
                        // have made its pass-A arrivals first (an early one would fall into a phase that is still open)
        asm volatile("bar.sync 1, %0;" ::"n"(CWARPS * 32) : "memory");
        if (lane == 0) {
#pragma unroll
            for (int q = 1; q < NLB; ++q) ws_mbar_arrive(&sh.req_bar[(done_at + q) % STAGES]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_compact_tma: stream compaction on the same ring (see k_scan_tma).  Pass A counts the non-zero elements of a stage, pass B
// -- once the tile's exclusive prefix `base` is back -- packs the survivors IN PLACE: every compute warp first takes its part of
// the tile into registers, one named barrier, then the survivors go to packed[mis + rank], mis = (out + base) mod 16 bytes, so
// that the 16-byte aligned part of the run out[base .. base + kept) is also 16-byte aligned in shared memory and leaves as ONE
// bulk store; the up to 3 + 3 ints before and after it are written by the DMA thread.  A stage is STAGE_INTS + 4 ints.
// `in` must be 16-byte aligned (bulk loads); `out` may be anything.
// ------------------------------------------------------------------------------------------------
template <int STAGE_INTS, int STAGES>
constexpr size_t compact_tma_smem_bytes() { return (size_t)STAGES * (STAGE_INTS + 4) * sizeof(int); }

template <int STAGE_INTS, int STAGES, int CWARPS, int DELAY, int NLB, int LBW, int STRIDE>
__global__ void __launch_bounds__((CWARPS + 1 + NLB) * 32, 1)
k_compact_tma(const int* in, int* out, unsigned n, unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets,
              unsigned epoch, int* __restrict__ d_count)
{
    static_assert(STAGE_INTS % (CWARPS * 128) == 0, "a stage is a whole number of 128-bit vectors per lane");
    static_assert(DELAY + 2 <= STAGES && NLB <= STAGES, "ring too short");
    constexpr int VPL = STAGE_INTS / (CWARPS * 128);
    constexpr int WARP_INTS = 32 * VPL * 4;
    constexpr int PITCH = STAGE_INTS + 4;
    extern __shared__ __align__(128) int ring[];                // [STAGES][PITCH]
    __shared__ TmaScanShared<STAGES, CWARPS> sh;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(((unsigned long long)n + STAGE_INTS - 1) / STAGE_INTS);
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ws_mbar_init(&sh.full_bar[s], 1); ws_mbar_init(&sh.req_bar[s], CWARPS);
            ws_mbar_init(&sh.resp_bar[s], 1); ws_mbar_init(&sh.store_bar[s], CWARPS);
        }
        sh.stop_at = 0xffffffffu;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (warp == CWARPS) {                                       // ---- the DMA thread ----
        if (lane != 0) return;
        unsigned issued = 0, stored = 0;
        bool more = true;
        unsigned next_t = atomicAdd(&tickets[2], 1u);
        for (;;) {
            if (more && issued - stored < (unsigned)STAGES) {
                const unsigned t = next_t;
                const int st = (int)(issued % STAGES);
                if (t >= tiles) {
                    sh.tile[st] = TMA_NO_TILE;
                    sh.stop_at = issued;
                    ws_mbar_arrive(&sh.full_bar[st]);
                    more = false;
                    continue;
                }
                next_t = atomicAdd(&tickets[2], 1u);
                sh.tile[st] = t;
                const unsigned long long base = (unsigned long long)t * STAGE_INTS;
                const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
                int* buf = ring + (size_t)st * PITCH;
                const unsigned whole = valid & ~3u;
                for (unsigned k = whole; k < valid; ++k) buf[k] = in[base + k];
                if (whole) {
                    ws_mbar_expect_tx(&sh.full_bar[st], whole * 4u);
                    ws_bulk_g2s(buf, in + base, whole * 4u, &sh.full_bar[st]);
                } else {
                    ws_mbar_arrive(&sh.full_bar[st]);
                }
                ++issued;
                continue;
            }
            if (stored < issued) {                              // the oldest tile: its packed run goes out
                const int st = (int)(stored % STAGES);
                ws_mbar_wait(&sh.store_bar[st], (stored / STAGES) & 1u);
                int kept = 0;
#pragma unroll
                for (int w = 0; w < CWARPS; ++w) kept += sh.warp_sums[st][w];
                int* dst = out + sh.resp[st];
                const int mis = (int)((reinterpret_cast<uintptr_t>(dst) >> 2) & 3u);
                int head = mis ? 4 - mis : 0;
                if (head > kept) head = kept;
                const int* packed = ring + (size_t)st * PITCH + mis;
                const int body4 = (kept - head) >> 2;
                if (body4) ws_bulk_s2g(dst + head, packed + head, (unsigned)body4 * 16u);
                ws_bulk_commit();
                for (int k = 0; k < head; ++k) dst[k] = packed[k];
                for (int k = head + 4 * body4; k < kept; ++k) dst[k] = packed[k];
                ws_bulk_wait_read0();
                ++stored;
                continue;
            }
            break;
        }
        ws_bulk_wait0();
        if (atomicAdd(&tickets[3], 1u) == gridDim.x - 1) { tickets[2] = 0u; tickets[3] = 0u; }
        return;
    }

    if (warp > CWARPS) {                                        // ---- the look-back warps ----
        for (unsigned i = (unsigned)(warp - CWARPS - 1);; i += NLB) {
            const int st = (int)(i % STAGES);
            ws_mbar_wait(&sh.req_bar[st], (i / STAGES) & 1u);
            if (i >= *(volatile unsigned*)&sh.stop_at) break;
            const unsigned t = sh.tile[st];
            int kept = 0;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) kept += sh.warp_sums[st][w];
            const int e = lb_exclusive_prefix_cfg<LBW, LB_DEFAULT_BACKOFF_NS, STRIDE>(desc, t, epoch, kept, 0, lane);
            __syncwarp();
            if (lane == 0) {
                if (t == tiles - 1) *d_count = e + kept;
                sh.resp[st] = e;
                ws_mbar_arrive(&sh.resp_bar[st]);
            }
        }
        return;
    }

    // ---- the compute warps ----
    unsigned done_at = 0xffffffffu;
    for (unsigned i = 0;; ++i) {
        if (i < done_at) {                                      // pass A over local tile i: how many it keeps
            const int st = (int)(i % STAGES);
            ws_mbar_wait(&sh.full_bar[st], (i / STAGES) & 1u);
            const unsigned t = sh.tile[st];
            if (t == TMA_NO_TILE) {
                done_at = i;
                if (lane == 0) ws_mbar_arrive(&sh.req_bar[st]);
            } else {
                const int4* p = reinterpret_cast<const int4*>(ring + (size_t)st * PITCH + warp * WARP_INTS) + lane;
                const unsigned long long base = (unsigned long long)t * STAGE_INTS;
                const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
                int s = 0;
                if (valid == (unsigned)STAGE_INTS) {
#pragma unroll
                    for (int j = 0; j < VPL; ++j) { const int4 v = p[j * 32]; s += (v.x != 0) + (v.y != 0) + (v.z != 0) + (v.w != 0); }
                } else {
#pragma unroll
                    for (int j = 0; j < VPL; ++j) {
                        const int4 v = p[j * 32];
                        const unsigned e = (unsigned)(warp * WARP_INTS + (j * 32 + lane) * 4);
                        s += (e + 0 < valid && v.x != 0) + (e + 1 < valid && v.y != 0) + (e + 2 < valid && v.z != 0) + (e + 3 < valid && v.w != 0);
                    }
                }
                s = warp_sum(s);
                if (lane == 0) {
                    sh.warp_sums[st][warp] = s;
                    ws_mbar_arrive(&sh.req_bar[st]);
                }
            }
        }
        if (i >= (unsigned)DELAY) {                             // pass B over local tile i - DELAY: pack in place
            const unsigned k = i - DELAY;
            if (k >= done_at) break;
            const int st = (int)(k % STAGES);
            ws_mbar_wait(&sh.resp_bar[st], (k / STAGES) & 1u);
            const int mis = (int)((reinterpret_cast<uintptr_t>(out + sh.resp[st]) >> 2) & 3u);
            int carry = mis;
#pragma unroll
            for (int w = 0; w < CWARPS; ++w) if (w < warp) carry += sh.warp_sums[st][w];
            int* stage = ring + (size_t)st * PITCH;
            const int4* p = reinterpret_cast<const int4*>(stage + warp * WARP_INTS) + lane;
            const unsigned t = sh.tile[st];
            const unsigned long long base = (unsigned long long)t * STAGE_INTS;
            const unsigned valid = (unsigned long long)n - base < (unsigned long long)STAGE_INTS ? (unsigned)(n - base) : (unsigned)STAGE_INTS;
            int4 v[VPL];
#pragma unroll
            for (int j = 0; j < VPL; ++j) v[j] = p[j * 32];
            if (valid != (unsigned)STAGE_INTS) {
#pragma unroll
                for (int j = 0; j < VPL; ++j) {
                    const unsigned e = (unsigned)(warp * WARP_INTS + (j * 32 + lane) * 4);
                    if (e + 0 >= valid) v[j].x = 0;
                    if (e + 1 >= valid) v[j].y = 0;
                    if (e + 2 >= valid) v[j].z = 0;
                    if (e + 3 >= valid) v[j].w = 0;
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(CWARPS * 32) : "memory");   // the whole tile is in registers: overwrite it
#pragma unroll
            for (int j = 0; j < VPL; ++j) {
                const int c = (v[j].x != 0) + (v[j].y != 0) + (v[j].z != 0) + (v[j].w != 0);
                const int inc = warp_inclusive_scan(c, lane);
                int q = carry + inc - c;
                if (v[j].x != 0) stage[q++] = v[j].x;
                if (v[j].y != 0) stage[q++] = v[j].y;
                if (v[j].z != 0) stage[q++] = v[j].z;
                if (v[j].w != 0) stage[q++] = v[j].w;
                carry += __shfl_sync(FULL_MASK, inc, 31);
            }
            ws_fence_async_smem();
            __syncwarp();
            if (lane == 0) ws_mbar_arrive(&sh.store_bar[st]);
        }
    }
    if (NLB > 1) {
        asm volatile("bar.sync 1, %0;" ::"n"(CWARPS * 32) : "memory");
        if (lane == 0) {
#pragma unroll
            for (int q = 1; q < NLB; ++q) ws_mbar_arrive(&sh.req_bar[(done_at + q) % STAGES]);
        }
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
