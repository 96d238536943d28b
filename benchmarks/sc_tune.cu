// sc_tune.cu -- configuration sweep of the single-pass scan / compact kernel templates on a B200.
//
//   nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -o benchmarks/sc_tune benchmarks/sc_tune.cu
//   benchmarks/sc_tune [lg_n = 28] [iters = 10]  > profiles/rN_sc_tune.jsonl
//
// For every (threads, vectors/thread, look-back window, ticket) it checks the result against a
// host scan (bit-exact) and prints one JSON line with the median CUDA-event time and the
// algorithmic GB/s (scan: 8 B/element; compact: 4 n + 4 count).  Inputs (1 GiB at 2^28) are far
// larger than the 126 MB L2, so successive launches never hit in L2.  The chosen configuration is
// recorded in cuda_pathtracer_b200/csrc/sc_kernels.cuh.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

#include "../cuda_pathtracer_b200/csrc/sc_tmpl.cuh"

using namespace b200;

#define CK(x)                                                                         \
    do {                                                                              \
        cudaError_t e__ = (x);                                                        \
        if (e__ != cudaSuccess) {                                                     \
            fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e__)); \
            exit(2);                                                                  \
        }                                                                             \
    } while (0)

struct Env {
    unsigned n;
    int *d_in, *d_flags, *d_out, *d_want_scan, *d_want_compact;
    int want_count;
    unsigned long long* desc;
    unsigned* tickets;
    unsigned epoch;
    int* d_small;
    int* d_diff;
    int iters;
};

__global__ void k_diff(const int* a, const int* b, unsigned n, int* diff)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned stride = gridDim.x * blockDim.x;
    int bad = 0;
    for (; i < n; i += stride) bad |= (a[i] != b[i]);
    if (bad) atomicOr(diff, 1);
}

__global__ void k_copy(const int4* __restrict__ in, int4* __restrict__ out, unsigned n4)
{
    unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) sct::st_stream(out + i, sct::ld_stream(in + i));
}

static bool same(Env& e, const int* a, const int* b, unsigned n)
{
    CK(cudaMemset(e.d_diff, 0, 4));
    if (n) k_diff<<<1184, 256>>>(a, b, n, e.d_diff);
    int h = 0;
    CK(cudaMemcpy(&h, e.d_diff, 4, cudaMemcpyDeviceToHost));
    return h == 0;
}

static float median_ms(Env& e, const std::function<void()>& launch)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    for (int i = 0; i < 3; ++i) launch();
    std::vector<float> t;
    for (int i = 0; i < e.iters; ++i) {
        CK(cudaEventRecord(a));
        launch();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        t.push_back(ms);
    }
    CK(cudaGetLastError());
    std::sort(t.begin(), t.end());
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    return t[t.size() / 2];
}

template <int TH, int VEC, int LBW, bool TICKET, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
static void run_scan(Env& e)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    auto k = sct::k_scan_lookback<C, true, LBW, TICKET, BACKOFF, STRIDE>;
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, TH, 0));
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    CK(cudaMemset(e.tickets, 0, 8));
    auto launch = [&] {
        e.epoch += 1;
        k<<<tiles, TH>>>(e.d_in, e.d_out, e.n, e.desc, e.tickets, e.epoch, nullptr, e.d_small);
    };
    const float ms = median_ms(e, launch);
    const bool ok = same(e, e.d_out, e.d_want_scan, e.n);
    printf("{\"op\": \"scan\", \"threads\": %d, \"vec\": %d, \"tile_ints\": %d, \"window\": %d, \"ticket\": %d, "
           "\"backoff_ns\": %d, \"desc_stride\": %d, \"regs\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           TH, VEC, C::tile, 32 * LBW, (int)TICKET, BACKOFF, STRIDE, fa.numRegs, occ, ms, 8.0 * e.n / ms / 1e6,
           ok ? "true" : "false");
    fflush(stdout);
}

template <int TH, int VEC, int LBW, bool TICKET, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
static void run_compact(Env& e)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    auto k = sct::k_compact<C, true, LBW, TICKET, BACKOFF, STRIDE>;
    const size_t smem = sct::compact_smem_bytes<C>();
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, TH, smem));
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    CK(cudaMemset(e.tickets, 0, 8));
    auto launch = [&] {
        e.epoch += 1;
        k<<<tiles, TH, smem>>>(e.d_flags, e.d_out, e.n, e.desc, e.tickets, e.epoch, e.d_small);
    };
    const float ms = median_ms(e, launch);
    int count = -1;
    CK(cudaMemcpy(&count, e.d_small, 4, cudaMemcpyDeviceToHost));
    const bool ok = count == e.want_count && same(e, e.d_out, e.d_want_compact, (unsigned)e.want_count);
    printf("{\"op\": \"compact\", \"threads\": %d, \"vec\": %d, \"tile_ints\": %d, \"window\": %d, \"ticket\": %d, "
           "\"backoff_ns\": %d, \"desc_stride\": %d, \"regs\": %d, \"ctas_per_sm\": %d, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           TH, VEC, C::tile, 32 * LBW, (int)TICKET, BACKOFF, STRIDE, fa.numRegs, occ, ms,
           (4.0 * e.n + 4.0 * e.want_count) / ms / 1e6, ok ? "true" : "false");
    fflush(stdout);
}


namespace b200 {
namespace sct {
// ------------------------------------------------------------------------------------------------
// EXPERIMENT (kept in the tuner only, not in the library): persistent, software-pipelined forms of the two
// single-pass kernels.  Result on a B200 at 2^28 (profiles/r8_sc_tune_persist.jsonl): bit-exact but SLOWER than
// one tile per CTA -- scan 0.42 ms vs 0.39, compact 0.53 vs 0.40: the static round-robin keeps all CTAs in
// lockstep (everybody loads, then everybody stores), and a CTA still has at most one tile in flight.
//
// In the one-tile-per-CTA kernels above a CTA's life is load -> local scan -> look-back -> store, and
// nothing of that CTA is in flight from HBM while its warp 0 waits for the predecessors' descriptors
// (the predecessors started at about the same time, so that wait is a sizeable part of the life:
// the same scan WITHOUT the look-back streams at 6.8 TB/s, with it at 5.4, profiles/r1_sc_tune_v3).
// Here the grid is as many CTAs as are resident at once and CTA c owns tiles c, c + G, c + 2G, ...
// The input of the CTA's NEXT tile is fetched with cp.async (LDGSTS: no registers held while in
// flight) into one of two shared-memory landing buffers at the top of every iteration, so the whole
// iteration -- look-back wait included -- overlaps that fetch.  Every thread fetches, and later
// reads, only its own vectors of the landing buffer: no barrier is involved in the hand-over.
// Forward progress: a tile's predecessors are owned by CTAs that are all resident and work their own
// tiles off in ascending order, and a CTA publishes the aggregate of a tile before it looks back.
// The last (partial) tile, and unaligned pointers, take the guarded synchronous load instead.
// dynamic shared memory: 2 * C::tile ints.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

template <class C>
constexpr size_t persist_smem_bytes() { return (size_t)2 * C::tile * sizeof(int); }

// fetch of `tile` (known to be full, pointers 16-byte aligned) into landing buffer `buf`
template <class C>
__device__ __forceinline__ void prefetch_tile(const int* __restrict__ in, unsigned tile, int warp, int lane, int4* buf)
{
    const int4* p = reinterpret_cast<const int4*>(in + (unsigned long long)tile * C::tile + warp * (32 * C::items)) + lane;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) cp_async16(buf + j * C::threads + threadIdx.x, p + j * 32);
}

template <class C, bool VEC_OK, int LBW, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
__global__ void __launch_bounds__(C::threads)
k_scan_persist(const int* __restrict__ in, int* __restrict__ out, unsigned n,
               unsigned long long* __restrict__ desc, unsigned epoch, const int* __restrict__ d_seed,
               int* __restrict__ d_total)
{
    extern __shared__ __align__(16) int4 landing[];          // [2][C::vec][C::threads]
    __shared__ ScanSmem<C> sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    const int seed = d_seed ? __ldg(d_seed) : 0;
    bool fetched = false;
    int cur = 0;
    for (unsigned tile = blockIdx.x; tile < tiles; tile += gridDim.x, cur ^= 1) {
        int4 v[C::vec];
        if (fetched) {
            cp_async_wait_all();
            const int4* b = landing + cur * (C::vec * C::threads);
#pragma unroll
            for (int j = 0; j < C::vec; ++j) v[j] = b[j * C::threads + threadIdx.x];
        } else {
            load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);
        }
        const unsigned long long next = (unsigned long long)tile + gridDim.x;
        fetched = VEC_OK && (next + 1ull) * C::tile <= n;
        if (fetched) prefetch_tile<C>(in, (unsigned)next, warp, lane, landing + (cur ^ 1) * (C::vec * C::threads));

        int s[C::vec], excl[C::vec];
#pragma unroll
        for (int j = 0; j < C::vec; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
        tile_warp_totals<C>(s, sm.warp_sums, warp, lane);
        __syncthreads();
        if (warp == 0) {                 // publish + look back first; the other warps scan meanwhile
            int aggregate = 0;
#pragma unroll
            for (int w = 0; w < C::warps; ++w) aggregate += sm.warp_sums[w];
            const int e = lb_exclusive_prefix_cfg<LBW, BACKOFF, STRIDE>(desc, tile, epoch, aggregate, seed, lane);
            if (lane == 0) {
                sm.tile_exclusive = e;
                if (d_total && tile == tiles - 1) *d_total = e + aggregate;
            }
        }
        tile_exclusive_from_totals<C>(s, excl, sm.warp_sums, warp, lane);
        __syncthreads();
        const int tile_exclusive = sm.tile_exclusive;
#pragma unroll
        for (int j = 0; j < C::vec; ++j) v[j] = scan_vec(v[j], tile_exclusive + excl[j]);
        store_tile<C, VEC_OK>(out, n, tile, warp, lane, v);
    }
}

// dynamic shared memory: 2 * C::tile ints (landing buffers) + C::tile ints (the packed survivors)
template <class C>
constexpr size_t compact_persist_smem_bytes() { return (size_t)3 * C::tile * sizeof(int); }

template <class C, bool VEC_OK, int LBW, int BACKOFF = LB_DEFAULT_BACKOFF_NS, int STRIDE = 1>
__global__ void __launch_bounds__(C::threads)
k_compact_persist(const int* __restrict__ in, int* __restrict__ out, unsigned n,
                  unsigned long long* __restrict__ desc, unsigned epoch, int* __restrict__ d_count)
{
    extern __shared__ __align__(16) int4 landing[];          // [2][C::vec][C::threads], then packed[C::tile]
    int* packed = reinterpret_cast<int*>(landing + 2 * C::vec * C::threads);
    __shared__ ScanSmem<C> sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    bool fetched = false;
    int cur = 0;
    for (unsigned tile = blockIdx.x; tile < tiles; tile += gridDim.x, cur ^= 1) {
        int4 v[C::vec];
        if (fetched) {
            cp_async_wait_all();
            const int4* b = landing + cur * (C::vec * C::threads);
#pragma unroll
            for (int j = 0; j < C::vec; ++j) v[j] = b[j * C::threads + threadIdx.x];
        } else {
            load_tile<C, VEC_OK>(in, n, tile, warp, lane, v);   // out-of-range reads as 0 == dropped
        }
        const unsigned long long next = (unsigned long long)tile + gridDim.x;
        fetched = VEC_OK && (next + 1ull) * C::tile <= n;
        if (fetched) prefetch_tile<C>(in, (unsigned)next, warp, lane, landing + (cur ^ 1) * (C::vec * C::threads));

        int s[C::vec], excl[C::vec];
#pragma unroll
        for (int j = 0; j < C::vec; ++j)
            s[j] = (v[j].x != 0) + (v[j].y != 0) + (v[j].z != 0) + (v[j].w != 0);
        tile_warp_totals<C>(s, sm.warp_sums, warp, lane);
        __syncthreads();                                     // (also: everybody is done with the previous tile's `packed`)
        int kept = 0;
#pragma unroll
        for (int w = 0; w < C::warps; ++w) kept += sm.warp_sums[w];
        if (warp == 0) {
            const int e = lb_exclusive_prefix_cfg<LBW, BACKOFF, STRIDE>(desc, tile, epoch, kept, 0, lane);
            if (lane == 0) {
                sm.tile_exclusive = e;
                if (tile == tiles - 1) *d_count = e + kept;
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
        // head: up to 3 elements until dst + head is 16-byte aligned; body: int4; tail: scalars
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
}

}  // namespace sct
}  // namespace b200

// ---- persistent, software-pipelined kernels (k_scan_persist / k_compact_persist) ---------------------
// ctas_per_sm_cap: 0 = as many as fit, else at most that many (fewer, fatter streams)
template <int TH, int VEC, int STRIDE>
static void run_scan_persist(Env& e, int ctas_per_sm_cap)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    auto k = sct::k_scan_persist<C, true, 1, 0, STRIDE>;
    const size_t smem = sct::persist_smem_bytes<C>();
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0, sms = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, TH, smem));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int per_sm = occ;
    if (ctas_per_sm_cap > 0 && per_sm > ctas_per_sm_cap) per_sm = ctas_per_sm_cap;
    unsigned grid = (unsigned)(sms * per_sm);
    if (grid > tiles) grid = tiles;
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    auto launch = [&] {
        e.epoch += 1;
        k<<<grid, TH, smem>>>(e.d_in, e.d_out, e.n, e.desc, e.epoch, nullptr, e.d_small);
    };
    const float ms = median_ms(e, launch);
    const bool ok = same(e, e.d_out, e.d_want_scan, e.n);
    printf("{\"op\": \"scan_persist\", \"threads\": %d, \"vec\": %d, \"tile_ints\": %d, \"desc_stride\": %d, \"regs\": %d, "
           "\"ctas_per_sm\": %d, \"occ\": %d, \"grid\": %u, \"smem\": %zu, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           TH, VEC, C::tile, STRIDE, fa.numRegs, per_sm, occ, grid, smem, ms, 8.0 * e.n / ms / 1e6, ok ? "true" : "false");
    fflush(stdout);
}

template <int TH, int VEC, int STRIDE>
static void run_compact_persist(Env& e, int ctas_per_sm_cap)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    auto k = sct::k_compact_persist<C, true, 1, 0, STRIDE>;
    const size_t smem = sct::compact_persist_smem_bytes<C>();
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0, sms = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, TH, smem));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    int per_sm = occ;
    if (ctas_per_sm_cap > 0 && per_sm > ctas_per_sm_cap) per_sm = ctas_per_sm_cap;
    unsigned grid = (unsigned)(sms * per_sm);
    if (grid > tiles) grid = tiles;
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    auto launch = [&] {
        e.epoch += 1;
        k<<<grid, TH, smem>>>(e.d_flags, e.d_out, e.n, e.desc, e.epoch, e.d_small);
    };
    const float ms = median_ms(e, launch);
    int count = -1;
    CK(cudaMemcpy(&count, e.d_small, 4, cudaMemcpyDeviceToHost));
    const bool ok = count == e.want_count && same(e, e.d_out, e.d_want_compact, (unsigned)e.want_count);
    printf("{\"op\": \"compact_persist\", \"threads\": %d, \"vec\": %d, \"tile_ints\": %d, \"desc_stride\": %d, \"regs\": %d, "
           "\"ctas_per_sm\": %d, \"occ\": %d, \"grid\": %u, \"smem\": %zu, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           TH, VEC, C::tile, STRIDE, fa.numRegs, per_sm, occ, grid, smem, ms, (4.0 * e.n + 4.0 * e.want_count) / ms / 1e6,
           ok ? "true" : "false");
    fflush(stdout);
}

template <int TH, int VEC>
static void run_persist_cfg(Env& e)
{
    for (int cap : {0, 4, 2}) {
        run_scan_persist<TH, VEC, 16>(e, cap);
        run_compact_persist<TH, VEC, 16>(e, cap);
    }
}

// ---- phase trace of the single-pass scan: where does a tile's lifetime go? ---------------------
__device__ __forceinline__ unsigned long long gtime()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

struct TileTrace {
    unsigned long long t_start, t_loaded, t_published, t_prefix, t_end;
    int steps, spins, smid, pad;
};

template <class C, int LBW, int BACKOFF>
__global__ void __launch_bounds__(C::threads)
k_scan_trace(const int* __restrict__ in, int* __restrict__ out, unsigned n, unsigned long long* __restrict__ desc,
             unsigned epoch, TileTrace* __restrict__ tr)
{
    __shared__ sct::ScanSmem<C> sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = blockIdx.x;
    TileTrace t;
    t.t_start = gtime();
    int4 v[C::vec];
    sct::load_tile<C, true>(in, n, tile, warp, lane, v);
    int s[C::vec], excl[C::vec];
#pragma unroll
    for (int j = 0; j < C::vec; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
    sct::tile_warp_totals<C>(s, sm.warp_sums, warp, lane);
    __syncthreads();
    t.t_loaded = gtime();
    if (warp == 0) {
        int aggregate = 0;
#pragma unroll
        for (int w = 0; w < C::warps; ++w) aggregate += sm.warp_sums[w];
        const int steps = 0, spins = 0;      // (the polling loop itself lives in lookback.cuh)
        t.t_published = gtime();
        const int exclusive = lb_exclusive_prefix_cfg<LBW, BACKOFF, 1>(desc, tile, epoch, aggregate, 0, lane);
        t.t_prefix = gtime();
        t.steps = steps;
        t.spins = spins;
        if (lane == 0) sm.tile_exclusive = exclusive;
    }
    sct::tile_exclusive_from_totals<C>(s, excl, sm.warp_sums, warp, lane);
    __syncthreads();
    const int tile_exclusive = sm.tile_exclusive;
#pragma unroll
    for (int j = 0; j < C::vec; ++j) v[j] = sct::scan_vec(v[j], tile_exclusive + excl[j]);
    sct::store_tile<C, true>(out, n, tile, warp, lane, v);
    if (threadIdx.x == 0) {
        t.t_end = gtime();
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        t.smid = (int)smid;
        t.pad = 0;
        tr[tile] = t;
    }
}

template <int TH, int VEC, int LBW, int BACKOFF>
static void run_trace(Env& e, const char* path)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    TileTrace* d_tr;
    CK(cudaMalloc(&d_tr, sizeof(TileTrace) * (size_t)tiles));
    for (int i = 0; i < 3; ++i) {
        e.epoch += 1;
        k_scan_trace<C, LBW, BACKOFF><<<tiles, TH>>>(e.d_in, e.d_out, e.n, e.desc, e.epoch, d_tr);
    }
    CK(cudaDeviceSynchronize());
    const bool ok = same(e, e.d_out, e.d_want_scan, e.n);
    std::vector<TileTrace> tr(tiles);
    CK(cudaMemcpy(tr.data(), d_tr, sizeof(TileTrace) * (size_t)tiles, cudaMemcpyDeviceToHost));
    CK(cudaFree(d_tr));
    FILE* f = fopen(path, "wb");
    if (f) { fwrite(tr.data(), sizeof(TileTrace), tiles, f); fclose(f); }
    unsigned long long t0 = ~0ull, t1 = 0;
    double load = 0, pub = 0, look = 0, store = 0, steps = 0, spins = 0;
    for (const TileTrace& t : tr) {
        if (t.t_start < t0) t0 = t.t_start;
        if (t.t_end > t1) t1 = t.t_end;
        load += (double)(t.t_loaded - t.t_start);
        pub += (double)(t.t_published - t.t_loaded);
        look += (double)(t.t_prefix - t.t_published);
        store += (double)(t.t_end - t.t_prefix);
        steps += t.steps;
        spins += t.spins;
    }
    printf("{\"op\": \"trace\", \"threads\": %d, \"vec\": %d, \"window\": %d, \"backoff_ns\": %d, \"ok\": %s, \"tiles\": %u, \"span_us\": %.1f, "
           "\"mean_ns\": {\"load\": %.0f, \"publish\": %.0f, \"lookback\": %.0f, \"store\": %.0f}, \"mean_steps\": %.2f, "
           "\"mean_spins\": %.2f}\n",
           TH, VEC, 32 * LBW, BACKOFF, ok ? "true" : "false", tiles, (double)(t1 - t0) / 1e3, load / tiles, pub / tiles, look / tiles,
           store / tiles, steps / tiles, spins / tiles);
    fflush(stdout);
}

template <int TH, int VEC>
static void run_downsweep(Env& e)
{
    using C = sct::Cfg<TH, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + C::tile - 1) / C::tile);
    auto launch = [&] { sct::k_downsweep<C, true><<<tiles, TH>>>(e.d_in, e.d_out, e.n, nullptr, nullptr, nullptr); };
    const float ms = median_ms(e, launch);
    printf("{\"op\": \"tile_scan_only\", \"threads\": %d, \"vec\": %d, \"tile_ints\": %d, \"ms\": %.4f, \"gbs\": %.1f}\n", TH,
           VEC, C::tile, ms, 8.0 * e.n / ms / 1e6);
    fflush(stdout);
}

template <int TH, int VEC, int LBW, int BACKOFF, int STRIDE>
static void run_both(Env& e, bool compact)
{
    run_scan<TH, VEC, LBW, false, BACKOFF, STRIDE>(e);
    if (compact) run_compact<TH, VEC, LBW, false, BACKOFF, STRIDE>(e);
}

template <int TH, int VEC>
static void run_cfg(Env& e)
{
    run_downsweep<TH, VEC>(e);
    const bool compact = sct::Cfg<TH, VEC>::tile * 4 <= 96 * 1024;
    run_both<TH, VEC, 1, 100, 1>(e, compact);
    run_both<TH, VEC, 1, 100, 4>(e, false);
    run_both<TH, VEC, 1, 0, 16>(e, compact);
    run_both<TH, VEC, 1, 50, 16>(e, false);
    run_both<TH, VEC, 1, 100, 16>(e, compact);
    run_both<TH, VEC, 1, 200, 16>(e, false);
    run_both<TH, VEC, 1, 0, 32>(e, false);
    run_both<TH, VEC, 1, 100, 32>(e, compact);
    run_both<TH, VEC, 1, 100, 64>(e, false);
    run_both<TH, VEC, 2, 100, 16>(e, false);
    run_scan<TH, VEC, 1, true, 100, 16>(e);
}

// ---- persistent TMA-ring scan (sct::k_scan_tma) ------------------------------------------------------
template <int STAGE_INTS, int STAGES, int CWARPS, int DELAY, int NLB, int LBW>
static void run_scan_tma(Env& e)
{
    auto k = sct::k_scan_tma<STAGE_INTS, STAGES, CWARPS, DELAY, NLB, LBW, 16>;
    const size_t smem = sct::scan_tma_smem_bytes<STAGE_INTS, STAGES>();
    const int threads = sct::scan_tma_threads<STAGE_INTS, CWARPS, NLB>();
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + STAGE_INTS - 1) / STAGE_INTS);
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0, sms = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, threads, smem));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    unsigned grid = (unsigned)(sms * occ);
    if (grid > tiles) grid = tiles;
    if (grid == 0) { printf("{\"op\": \"scan_tma\", \"error\": \"does not fit\"}\n"); return; }
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    unsigned* tk = nullptr;                                     // [2], [3]: the kernel's tile / done counters
    CK(cudaMalloc(&tk, 16));
    CK(cudaMemset(tk, 0, 16));
    auto launch = [&] {
        e.epoch += 1;
        k<<<grid, threads, smem>>>(e.d_in, e.d_out, e.n, e.desc, tk, e.epoch, nullptr, e.d_small);
    };
    const float ms = median_ms(e, launch);
    int total = 0, want_total = 0, last_in = 0;
    CK(cudaMemcpy(&total, e.d_small, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&want_total, e.d_want_scan + (e.n - 1), 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&last_in, e.d_in + (e.n - 1), 4, cudaMemcpyDeviceToHost));
    const bool ok = same(e, e.d_out, e.d_want_scan, e.n) && total == want_total + last_in;
    printf("{\"op\": \"scan_tma\", \"stage_ints\": %d, \"stages\": %d, \"cwarps\": %d, \"delay\": %d, \"nlb\": %d, \"window\": %d, "
           "\"regs\": %d, \"occ\": %d, \"grid\": %u, \"smem\": %zu, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           STAGE_INTS, STAGES, CWARPS, DELAY, NLB, 32 * LBW, fa.numRegs, occ, grid, smem, ms, 8.0 * e.n / ms / 1e6,
           ok ? "true" : "false");
    fflush(stdout);
    CK(cudaFree(tk));
}

template <int STAGE_INTS, int STAGES, int CWARPS, int DELAY, int NLB, int LBW>
static void run_compact_tma(Env& e)
{
    auto k = sct::k_compact_tma<STAGE_INTS, STAGES, CWARPS, DELAY, NLB, LBW, 16>;
    const size_t smem = sct::compact_tma_smem_bytes<STAGE_INTS, STAGES>();
    const int threads = sct::scan_tma_threads<STAGE_INTS, CWARPS, NLB>();
    const unsigned tiles = (unsigned)(((unsigned long long)e.n + STAGE_INTS - 1) / STAGE_INTS);
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaFuncAttributes fa;
    CK(cudaFuncGetAttributes(&fa, k));
    int occ = 0, sms = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k, threads, smem));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    unsigned grid = (unsigned)(sms * occ);
    if (grid > tiles) grid = tiles;
    if (grid == 0) { printf("{\"op\": \"compact_tma\", \"error\": \"does not fit\"}\n"); return; }
    CK(cudaMemset(e.d_out, 0xff, (size_t)e.n * 4));
    unsigned* tk = nullptr;
    CK(cudaMalloc(&tk, 16));
    CK(cudaMemset(tk, 0, 16));
    auto launch = [&] {
        e.epoch += 1;
        k<<<grid, threads, smem>>>(e.d_flags, e.d_out, e.n, e.desc, tk, e.epoch, e.d_small);
    };
    const float ms = median_ms(e, launch);
    int count = -1;
    CK(cudaMemcpy(&count, e.d_small, 4, cudaMemcpyDeviceToHost));
    const bool ok = count == e.want_count && same(e, e.d_out, e.d_want_compact, (unsigned)e.want_count);
    printf("{\"op\": \"compact_tma\", \"stage_ints\": %d, \"stages\": %d, \"cwarps\": %d, \"delay\": %d, \"nlb\": %d, \"window\": %d, "
           "\"regs\": %d, \"occ\": %d, \"grid\": %u, \"smem\": %zu, \"ms\": %.4f, \"gbs\": %.1f, \"ok\": %s}\n",
           STAGE_INTS, STAGES, CWARPS, DELAY, NLB, 32 * LBW, fa.numRegs, occ, grid, smem, ms,
           (4.0 * e.n + 4.0 * e.want_count) / ms / 1e6, ok ? "true" : "false");
    fflush(stdout);
    CK(cudaFree(tk));
}

int main(int argc, char** argv)
{
    const int lg = argc > 1 ? atoi(argv[1]) : 28;
    Env e;
    e.iters = argc > 2 ? atoi(argv[2]) : 10;
    e.n = 1u << lg;
    const size_t bytes = (size_t)e.n * 4;
    std::vector<int> h(e.n), f(e.n), ws(e.n), wc;
    srand(565);
    for (unsigned i = 0; i < e.n; ++i) { const int r = rand(); h[i] = r % 50; f[i] = (r >> 8) % 4; }
    int run = 0;
    wc.reserve(e.n);
    for (unsigned i = 0; i < e.n; ++i) {
        ws[i] = run;
        run = (int)((unsigned)run + (unsigned)h[i]);
        if (f[i]) wc.push_back(f[i]);
    }
    e.want_count = (int)wc.size();
    CK(cudaMalloc(&e.d_in, bytes));
    CK(cudaMalloc(&e.d_flags, bytes));
    CK(cudaMalloc(&e.d_out, bytes));
    CK(cudaMalloc(&e.d_want_scan, bytes));
    CK(cudaMalloc(&e.d_want_compact, bytes));
    CK(cudaMemcpy(e.d_in, h.data(), bytes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e.d_flags, f.data(), bytes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e.d_want_scan, ws.data(), bytes, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(e.d_want_compact, wc.data(), wc.size() * 4, cudaMemcpyHostToDevice));
    const size_t max_tiles = (e.n + 2047) / 2048;
    CK(cudaMalloc(&e.desc, max_tiles * 8 * 64));
    CK(cudaMemset(e.desc, 0, max_tiles * 8 * 64));
    CK(cudaMalloc(&e.tickets, 8));
    CK(cudaMemset(e.tickets, 0, 8));
    CK(cudaMalloc(&e.d_small, 16));
    CK(cudaMalloc(&e.d_diff, 4));
    e.epoch = 0;

    {
        const unsigned n4 = e.n / 4;
        auto launch = [&] { k_copy<<<(n4 + 255) / 256, 256>>>((const int4*)e.d_in, (int4*)e.d_out, n4); };
        const float ms = median_ms(e, launch);
        printf("{\"op\": \"copy\", \"ms\": %.4f, \"gbs\": %.1f}\n", ms, 8.0 * e.n / ms / 1e6);
        auto launch2 = [&] { CK(cudaMemcpyAsync(e.d_out, e.d_in, bytes, cudaMemcpyDeviceToDevice)); };
        const float ms2 = median_ms(e, launch2);
        printf("{\"op\": \"cudaMemcpyD2D\", \"ms\": %.4f, \"gbs\": %.1f}\n", ms2, 8.0 * e.n / ms2 / 1e6);
    }
    if (argc > 3 && std::string(argv[3]) == "tma") {          // benchmarks/sc_tune 28 10 tma
        run_scan<128, 8, 1, false, 0, 16>(e);                 // the one-tile-per-CTA kernel the library uses
        //           ints  stages warps delay nlb window/32
        run_scan_tma<8192, 7, 8, 3, 4, 1>(e);
        run_scan_tma<7168, 8, 8, 3, 4, 1>(e);                 // finer stages: more of them for the same shared memory
        run_scan_tma<7168, 8, 8, 4, 4, 1>(e);
        run_scan_tma<6144, 9, 8, 3, 4, 1>(e);
        run_scan_tma<6144, 9, 8, 4, 4, 1>(e);
        run_scan_tma<6144, 9, 8, 5, 6, 1>(e);
        run_scan_tma<5120, 11, 8, 5, 6, 1>(e);
        run_compact<256, 8, 1, false, 0, 16>(e);
        run_compact_tma<8192, 7, 8, 3, 4, 1>(e);
        run_compact_tma<7168, 8, 8, 3, 4, 1>(e);
        run_compact_tma<7168, 8, 8, 4, 4, 1>(e);
        run_compact_tma<6144, 9, 8, 3, 4, 1>(e);
        run_compact_tma<6144, 9, 8, 4, 4, 1>(e);
        run_compact_tma<6144, 9, 8, 5, 6, 1>(e);
        run_compact_tma<5120, 11, 8, 5, 6, 1>(e);
        return 0;
    }
    if (argc > 3 && std::string(argv[3]) == "persist") {      // benchmarks/sc_tune 28 10 persist
        run_scan<128, 8, 1, false, 0, 16>(e);                 // the one-tile-per-CTA kernels the library used so far
        run_compact<256, 8, 1, false, 0, 16>(e);
        run_persist_cfg<128, 8>(e);
        run_persist_cfg<256, 4>(e);
        run_persist_cfg<256, 8>(e);
        run_persist_cfg<128, 4>(e);
        run_persist_cfg<512, 4>(e);
        run_persist_cfg<64, 8>(e);
        run_persist_cfg<128, 16>(e);
        return 0;
    }
    if (argc > 3) {                      // trace mode: benchmarks/sc_tune 28 10 out_prefix
        const std::string pre = argv[3];
        run_trace<256, 4, 1, 0>(e, (pre + "_256x4_b0.bin").c_str());
        run_trace<256, 4, 1, 100>(e, (pre + "_256x4_b100.bin").c_str());
        run_trace<512, 4, 1, 100>(e, (pre + "_512x4_b100.bin").c_str());
        return 0;
    }
    run_cfg<64, 8>(e);
    run_cfg<64, 16>(e);
    run_cfg<128, 4>(e);
    run_cfg<128, 8>(e);
    run_cfg<128, 16>(e);
    run_cfg<256, 4>(e);
    run_cfg<256, 8>(e);
    run_cfg<512, 4>(e);
    return 0;
}
