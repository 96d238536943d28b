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
