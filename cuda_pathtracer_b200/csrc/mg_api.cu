// mg_api.cu -- the C ABI of include/b200_mg.h: the two hot paths over the K GPUs of one box, in ONE process.
//
// Plumbing only -- the kernels are libb200pt.so's and libb200sc.so's, reached through their own C ABIs:
//   * one host thread per device enqueues that device's share (b200pt_render / b200_*_dev are asynchronous),
//   * the one collective of each path is issued for all devices from the calling thread inside
//     ncclGroupStart / ncclGroupEnd on the devices' own streams, so everything stays stream-ordered,
//   * nothing synchronises the host between the phases except the final copy to host memory.
// The reference has no multi-GPU code (path_tracer/src/main.cpp:34-47); the partitioning follows SURVEY.md 8e.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <functional>
#include <new>
#include <thread>
#include <vector>

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>            // types and prototypes only: the library is bound at run time, see NcclApi below

#include "../../include/b200_mg.h"
#include "../../include/b200_sc.h"

struct b200mg {
    int K = 0;
    std::vector<int> dev;
    std::vector<ncclComm_t> comm;            // empty when K == 1: no collective is ever issued
    std::vector<cudaStream_t> stream;
    std::vector<b200_sc_ctx*> sc;
    std::vector<int*> d_small;               // per device: [0] total / count, [4 .. 4+K) gathered, [4+K .. 4+2K) seeds
    std::vector<int*> d_in, d_out;           // per device staging of the host-pointer scan / compact (grow-only)
    std::vector<size_t> cap;                 // ... capacity in ints
    std::vector<cudaEvent_t> ev;             // per device: 4 events (kernels begin / end, exchange begin / end)
};

struct b200mg_pt {
    b200mg* g = nullptr;
    int W = 0, H = 0;
    std::vector<b200pt*> h;
    std::vector<float*> d_red;               // the all-reduced SUM image on every device (K > 1)
    void* d_pbo = nullptr;                   // device 0: uchar4 preview
    bool reduced = false;
};

namespace {

// NCCL is bound with dlopen when the first multi-device group is created, not at link time: a process that already
// carries an NCCL (a Python process that imported torch brings its own libnccl.so.2) keeps using THAT one -- dlopen finds
// the loaded library by its SONAME -- and a process that never asks for more than one device never loads NCCL at all.
// (A link-time dependency made `import torch` fail when this library happened to be loaded first: the system's older
// libnccl.so.2 then shadowed the one torch is built against.)
struct NcclApi {
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    bool ok = false;
};

NcclApi& nccl()
{
    static NcclApi api = [] {
        NcclApi a;
        void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
        if (!lib) return a;
        a.CommInitAll = (decltype(a.CommInitAll))dlsym(lib, "ncclCommInitAll");
        a.CommDestroy = (decltype(a.CommDestroy))dlsym(lib, "ncclCommDestroy");
        a.GroupStart = (decltype(a.GroupStart))dlsym(lib, "ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))dlsym(lib, "ncclGroupEnd");
        a.AllReduce = (decltype(a.AllReduce))dlsym(lib, "ncclAllReduce");
        a.AllGather = (decltype(a.AllGather))dlsym(lib, "ncclAllGather");
        a.GetErrorString = (decltype(a.GetErrorString))dlsym(lib, "ncclGetErrorString");
        a.ok = a.CommInitAll && a.CommDestroy && a.GroupStart && a.GroupEnd && a.AllReduce && a.AllGather && a.GetErrorString;
        return a;
    }();
    return api;
}

struct OnDevice {
    int prev = -1;
    explicit OnDevice(int d) { cudaGetDevice(&prev); if (prev != d) cudaSetDevice(d); else prev = -1; }
    ~OnDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

inline int nccl_code(ncclResult_t r) { return r == ncclSuccess ? 0 : B200_E_NCCL + (int)r; }

// fn(g) on one host thread per device (the thread has the device current); first non-zero code wins
int for_each_device(const b200mg* G, const std::function<int(int)>& fn)
{
    std::vector<int> rc((size_t)G->K, 0);
    if (G->K == 1) {
        OnDevice on(G->dev[0]);
        return fn(0);
    }
    std::vector<std::thread> th;
    th.reserve((size_t)G->K);
    for (int g = 0; g < G->K; ++g)
        th.emplace_back([&, g] {
            if (cudaSetDevice(G->dev[g]) != cudaSuccess) { rc[g] = (int)cudaGetLastError(); return; }
            rc[g] = fn(g);
        });
    for (auto& t : th) t.join();
    for (int r : rc) if (r) return r;
    return 0;
}

#define CKC(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) return (int)e__; } while (0)

// [begin, end) of device g's contiguous shard of n ints; boundaries are multiples of 4 ints (16 bytes) so that every
// shard keeps the 128-bit vector path (cuda_pathtracer_b200/distributed.py::scan_shard is the same rule)
void shard_of(int n, int g, int K, int& b, int& e)
{
    long long per = ((long long)n + K - 1) / K;
    per = (per + 3) / 4 * 4;
    const long long bb = std::min<long long>(n, (long long)g * per);
    b = (int)bb;
    e = (int)std::min<long long>(n, bb + per);
}

int ensure_staging(b200mg* G, int g, size_t ints)
{
    if (G->cap[g] >= ints && G->d_in[g]) return 0;
    cudaFree(G->d_in[g]); cudaFree(G->d_out[g]);
    G->d_in[g] = G->d_out[g] = nullptr; G->cap[g] = 0;
    const size_t want = std::max<size_t>((ints + 3) / 4 * 4, 1024);
    cudaError_t e = cudaMalloc((void**)&G->d_in[g], want * sizeof(int));
    if (e == cudaSuccess) e = cudaMalloc((void**)&G->d_out[g], want * sizeof(int));
    if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e; }
    G->cap[g] = want;
    return 0;
}

// ONE all-gather of one int per device: small[0] of every device -> small[4 .. 4+K) of every device
int gather_ints(b200mg* G)
{
    // (the events go around the group: NCCL enqueues a group's kernels at ncclGroupEnd, not at the call inside it)
    for (int g = 0; g < G->K; ++g) { OnDevice on(G->dev[g]); cudaEventRecord(G->ev[4 * g + 2], G->stream[g]); }
    ncclResult_t r = nccl().GroupStart();
    for (int g = 0; g < G->K && r == ncclSuccess; ++g) {
        OnDevice on(G->dev[g]);
        r = nccl().AllGather(G->d_small[g], G->d_small[g] + 4, 1, ncclInt32, G->comm[g], G->stream[g]);
    }
    const ncclResult_t r2 = nccl().GroupEnd();
    for (int g = 0; g < G->K; ++g) { OnDevice on(G->dev[g]); cudaEventRecord(G->ev[4 * g + 3], G->stream[g]); }
    return nccl_code(r != ncclSuccess ? r : r2);
}

float max_elapsed(b200mg* G, int first, int second)
{
    float worst = 0.f;
    for (int g = 0; g < G->K; ++g) {
        OnDevice on(G->dev[g]);
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, G->ev[4 * g + first], G->ev[4 * g + second]) == cudaSuccess) worst = std::max(worst, ms);
        else cudaGetLastError();
    }
    return worst;
}

}  // namespace

extern "C" {

int b200mg_create(int n_devices, const int* devices, b200mg** out)
{
    if (!out) return B200_E_BADARG;
    int visible = 0;
    if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) { cudaGetLastError(); return B200_E_NODEVICE; }
    if (n_devices <= 0) n_devices = visible;
    if (n_devices > visible && !devices) return B200_E_BADARG;
    b200mg* G = new (std::nothrow) b200mg();
    if (!G) return B200_E_NOMEM;
    G->K = n_devices;
    for (int g = 0; g < n_devices; ++g) {
        const int d = devices ? devices[g] : g;
        if (d < 0 || d >= visible || std::find(G->dev.begin(), G->dev.end(), d) != G->dev.end()) { delete G; return B200_E_BADARG; }
        G->dev.push_back(d);
    }
    const size_t K = (size_t)G->K;
    G->stream.assign(K, nullptr); G->sc.assign(K, nullptr); G->d_small.assign(K, nullptr);
    G->d_in.assign(K, nullptr); G->d_out.assign(K, nullptr); G->cap.assign(K, 0); G->ev.assign(4 * K, nullptr);
    int rc = 0;
    for (int g = 0; g < G->K && !rc; ++g) {
        OnDevice on(G->dev[g]);
        cudaError_t e = cudaStreamCreateWithFlags(&G->stream[g], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc((void**)&G->d_small[g], sizeof(int) * (16 + 2 * K));
        if (e == cudaSuccess) e = cudaMemset(G->d_small[g], 0, sizeof(int) * (16 + 2 * K));
        for (int k = 0; k < 4 && e == cudaSuccess; ++k) e = cudaEventCreate(&G->ev[4 * g + k]);
        rc = (int)e;
        if (!rc) rc = b200_sc_ctx_create(&G->sc[g]);
    }
    if (!rc && G->K > 1) {
        G->comm.assign(K, nullptr);
        rc = nccl().ok ? nccl_code(nccl().CommInitAll(G->comm.data(), G->K, G->dev.data())) : B200_E_NCCL + 99;
        if (rc) G->comm.clear();
    }
    for (int g = 0; g < G->K && !rc; ++g) {      // load the scan library's kernels on every device (CUDA loads modules lazily)
        OnDevice on(G->dev[g]);
        int* s4 = G->d_small[g];
        rc = b200_reduce_i32_dev(G->sc[g], 4, s4 + 4, s4, G->stream[g]);
        if (!rc) rc = b200_scan_i32_dev(G->sc[g], 4, s4 + 4, s4, B200_SCAN_LOOKBACK, nullptr, nullptr, G->stream[g]);
        if (!rc) rc = b200_compact_i32_dev(G->sc[g], 4, s4 + 4, s4, s4 + 8, G->stream[g]);
        if (!rc) rc = (int)cudaMemsetAsync(s4, 0, sizeof(int) * 12, G->stream[g]);
        if (!rc) rc = (int)cudaStreamSynchronize(G->stream[g]);
    }
    if (!rc && G->K > 1) {            // NCCL sets its channels up lazily, at the first collective (tens of ms): pay that here
        rc = gather_ints(G);
        for (int g = 0; g < G->K && !rc; ++g) {
            OnDevice on(G->dev[g]);
            rc = (int)cudaStreamSynchronize(G->stream[g]);
        }
    }
    if (rc) { b200mg_destroy(G); return rc; }
    *out = G;
    return 0;
}

int b200mg_destroy(b200mg* G)
{
    if (!G) return 0;
    for (ncclComm_t c : G->comm) if (c) nccl().CommDestroy(c);
    for (int g = 0; g < G->K; ++g) {
        OnDevice on(G->dev[g]);
        if ((size_t)g < G->sc.size()) b200_sc_ctx_destroy(G->sc[g]);
        if ((size_t)g < G->d_small.size()) { cudaFree(G->d_small[g]); cudaFree(G->d_in[g]); cudaFree(G->d_out[g]); }
        for (int k = 0; k < 4; ++k) if ((size_t)(4 * g + k) < G->ev.size() && G->ev[4 * g + k]) cudaEventDestroy(G->ev[4 * g + k]);
        if ((size_t)g < G->stream.size() && G->stream[g]) cudaStreamDestroy(G->stream[g]);
    }
    delete G;
    return 0;
}

int b200mg_device_count(const b200mg* G) { return G ? G->K : 0; }

// ---- path tracer ---------------------------------------------------------------------------------------------
int b200mg_pt_create(b200mg* G, const b200pt_scene* scene, const b200pt_scene_ext* ext, b200mg_pt** out)
{
    if (!G || !out || !scene) return B200_E_BADARG;
    b200mg_pt* R = new (std::nothrow) b200mg_pt();
    if (!R) return B200_E_NOMEM;
    R->g = G;
    R->W = scene->camera.resolution[0];
    R->H = scene->camera.resolution[1];
    R->h.assign((size_t)G->K, nullptr);
    R->d_red.assign((size_t)G->K, nullptr);
    int rc = for_each_device(G, [&](int g) {          // the scene is replicated: every device uploads its own copy
        int r = b200pt_create_ext(scene, ext, &R->h[g]);
        if (!r && G->K > 1) {
            cudaError_t e = cudaMalloc((void**)&R->d_red[g], sizeof(float) * 3 * (size_t)R->W * R->H);
            if (e != cudaSuccess) { cudaGetLastError(); r = e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e; }
        }
        return r;
    });
    if (!rc) {
        OnDevice on(G->dev[0]);
        cudaError_t e = cudaMalloc(&R->d_pbo, (size_t)R->W * R->H * 4);
        if (e != cudaSuccess) { cudaGetLastError(); rc = e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e; }
    }
    if (rc) { b200mg_pt_destroy(R); return rc; }
    *out = R;
    return 0;
}

int b200mg_pt_destroy(b200mg_pt* R)
{
    if (!R) return 0;
    for (int g = 0; g < R->g->K; ++g) {
        OnDevice on(R->g->dev[g]);
        cudaStreamSynchronize(R->g->stream[g]);
        b200pt_destroy(R->h[g]);
        cudaFree(R->d_red[g]);
        if (g == 0) cudaFree(R->d_pbo);
    }
    delete R;
    return 0;
}

int b200mg_pt_set_options(b200mg_pt* R, const b200pt_options* o)
{
    if (!R || !o) return B200_E_BADARG;
    for (b200pt* h : R->h) { const int r = b200pt_set_options(h, o); if (r) return r; }
    return 0;
}

int b200mg_pt_reset(b200mg_pt* R)
{
    if (!R) return B200_E_BADARG;
    R->reduced = false;
    return for_each_device(R->g, [&](int g) {
        CKC(cudaStreamSynchronize(R->g->stream[g]));
        return b200pt_reset(R->h[g]);
    });
}

int b200mg_pt_write_sum(b200mg_pt* R, const float* host_rgb)
{
    if (!R || !host_rgb) return B200_E_BADARG;
    int r = b200mg_pt_reset(R);                          // the other devices start from zero: the sum over devices is host_rgb
    if (!r) r = b200pt_write_sum(R->h[0], host_rgb);
    return r;
}

int b200mg_pt_render(b200mg_pt* R, int first_iter, int n_iters)
{
    if (!R || n_iters < 0 || first_iter < 0) return B200_E_BADARG;
    const int K = R->g->K;
    R->reduced = false;
    return for_each_device(R->g, [&](int g) {
        const int count = n_iters > g ? (n_iters - g + K - 1) / K : 0;
        return b200pt_render(R->h[g], first_iter + g, count, K, R->g->stream[g], 0);
    });
}

int b200mg_pt_reduce_read_sum(b200mg_pt* R, float* host_rgb, float* allreduce_ms)
{
    if (!R || !host_rgb) return B200_E_BADARG;
    b200mg* G = R->g;
    const size_t n = 3 * (size_t)R->W * R->H;
    if (allreduce_ms) *allreduce_ms = 0.f;
    if (G->K == 1) {
        OnDevice on(G->dev[0]);
        CKC(cudaStreamSynchronize(G->stream[0]));
        R->reduced = true;
        return b200pt_read_sum(R->h[0], host_rgb);
    }
    for (int g = 0; g < G->K; ++g) { OnDevice on(G->dev[g]); cudaEventRecord(G->ev[4 * g + 2], G->stream[g]); }
    ncclResult_t r = nccl().GroupStart();
    for (int g = 0; g < G->K && r == ncclSuccess; ++g) {
        OnDevice on(G->dev[g]);
        void* img = nullptr;
        const int rc = b200pt_sum_device_ptr(R->h[g], &img, nullptr);
        if (rc) { nccl().GroupEnd(); return rc; }
        r = nccl().AllReduce(img, R->d_red[g], n, ncclFloat, ncclSum, G->comm[g], G->stream[g]);
    }
    const ncclResult_t r2 = nccl().GroupEnd();
    if (r != ncclSuccess || r2 != ncclSuccess) return nccl_code(r != ncclSuccess ? r : r2);
    for (int g = 0; g < G->K; ++g) { OnDevice on(G->dev[g]); cudaEventRecord(G->ev[4 * g + 3], G->stream[g]); }
    {
        OnDevice on(G->dev[0]);
        CKC(cudaMemcpyAsync(host_rgb, R->d_red[0], n * sizeof(float), cudaMemcpyDeviceToHost, G->stream[0]));
    }
    for (int g = 0; g < G->K; ++g) {
        OnDevice on(G->dev[g]);
        CKC(cudaStreamSynchronize(G->stream[g]));
    }
    if (allreduce_ms) *allreduce_ms = max_elapsed(G, 2, 3);
    R->reduced = true;
    return 0;
}

int b200mg_pt_preview_u8(b200mg_pt* R, int iter_divisor, void* host_rgba8)
{
    if (!R || !host_rgba8 || iter_divisor <= 0) return B200_E_BADARG;
    if (!R->reduced) return B200_E_BADARG;               // b200mg_pt_reduce_read_sum first
    b200mg* G = R->g;
    OnDevice on(G->dev[0]);
    int rc = G->K == 1 ? b200pt_tonemap_u8(R->h[0], iter_divisor, R->d_pbo, G->stream[0])
                       : b200pt_tonemap_buffer_u8(R->h[0], R->d_red[0], iter_divisor, R->d_pbo, G->stream[0]);
    if (rc) return rc;
    CKC(cudaMemcpyAsync(host_rgba8, R->d_pbo, (size_t)R->W * R->H * 4, cudaMemcpyDeviceToHost, G->stream[0]));
    CKC(cudaStreamSynchronize(G->stream[0]));
    return 0;
}

int b200mg_pt_get_stats(b200mg_pt* R, b200pt_stats* total)
{
    if (!R || !total) return B200_E_BADARG;
    std::memset(total, 0, sizeof(*total));
    for (int g = 0; g < R->g->K; ++g) {
        b200pt_stats s;
        const int rc = b200pt_get_stats(R->h[g], &s);
        if (rc) return rc;
        total->path_segments += s.path_segments;
        total->iterations += s.iterations;
        total->kernel_launches += s.kernel_launches;
        total->last_render_ms = std::max(total->last_render_ms, s.last_render_ms);
        total->traced_depth = std::max(total->traced_depth, s.traced_depth);
        total->bounce_kernel_ms += s.bounce_kernel_ms;
        total->bounce_kernel_launches += s.bounce_kernel_launches;
    }
    return 0;
}

// ---- scan / compact --------------------------------------------------------------------------------------------
int b200mg_scan_i32(b200mg* G, int n, int* odata, const int* idata, float* kernel_ms, float* exchange_ms)
{
    if (!G || n < 0 || (n > 0 && (!odata || !idata))) return B200_E_BADARG;
    if (kernel_ms) *kernel_ms = 0.f;
    if (exchange_ms) *exchange_ms = 0.f;
    if (n == 0) return 0;
    const int K = G->K;
    // phase A: shard to the device, per-shard total
    int rc = for_each_device(G, [&](int g) {
        int b, e;
        shard_of(n, g, K, b, e);
        int r = ensure_staging(G, g, (size_t)(e - b));
        if (!r) r = b200_sc_ctx_reserve(G->sc[g], e - b);      // the look-back workspace too: no allocation inside the timed region
        if (r) return r;
        cudaStream_t s = G->stream[g];
        if (e > b) CKC(cudaMemcpyAsync(G->d_in[g], idata + b, sizeof(int) * (size_t)(e - b), cudaMemcpyHostToDevice, s));
        CKC(cudaStreamSynchronize(s));       // every shard is resident before any device's kernel timer starts (copies are
        return 0;                            // excluded from kernel_ms, the reference's convention; the devices' copies end at different times)
    });
    if (rc) return rc;
    rc = for_each_device(G, [&](int g) {
        int b, e;
        shard_of(n, g, K, b, e);
        cudaStream_t s = G->stream[g];
        CKC(cudaEventRecord(G->ev[4 * g + 0], s));
        return K > 1 ? b200_reduce_i32_dev(G->sc[g], e - b, G->d_in[g], G->d_small[g], s) : 0;
    });
    if (rc) return rc;
    // the one exchange: every device learns every shard's total
    if (K > 1) { rc = gather_ints(G); if (rc) return rc; }
    // phase B: seed = exclusive prefix of the totals before this shard (on the device), seeded single-pass scan
    rc = for_each_device(G, [&](int g) {
        int b, e;
        shard_of(n, g, K, b, e);
        cudaStream_t s = G->stream[g];
        int r = 0;
        const int* seed = nullptr;
        if (K > 1) {
            r = b200_scan_i32_dev(G->sc[g], K, G->d_small[g] + 4 + K, G->d_small[g] + 4, B200_SCAN_LOOKBACK, nullptr, nullptr, s);
            if (r) return r;
            seed = G->d_small[g] + 4 + K + g;
        }
        r = b200_scan_i32_dev(G->sc[g], e - b, G->d_out[g], G->d_in[g], B200_SCAN_LOOKBACK, seed, nullptr, s);
        if (r) return r;
        CKC(cudaEventRecord(G->ev[4 * g + 1], s));
        if (e > b) CKC(cudaMemcpyAsync(odata + b, G->d_out[g], sizeof(int) * (size_t)(e - b), cudaMemcpyDeviceToHost, s));
        CKC(cudaStreamSynchronize(s));
        return 0;
    });
    if (rc) return rc;
    if (kernel_ms) *kernel_ms = max_elapsed(G, 0, 1);
    if (exchange_ms && K > 1) *exchange_ms = max_elapsed(G, 2, 3);
    return 0;
}

int b200mg_compact_i32(b200mg* G, int n, int* odata, const int* idata, int* count, float* kernel_ms, float* exchange_ms)
{
    if (!G || n < 0 || !count || (n > 0 && (!odata || !idata))) return B200_E_BADARG;
    if (kernel_ms) *kernel_ms = 0.f;
    if (exchange_ms) *exchange_ms = 0.f;
    *count = 0;
    if (n == 0) return 0;
    const int K = G->K;
    int rc = for_each_device(G, [&](int g) {
        int b, e;
        shard_of(n, g, K, b, e);
        int r = ensure_staging(G, g, (size_t)(e - b));
        if (!r) r = b200_sc_ctx_reserve(G->sc[g], e - b);
        if (r) return r;
        cudaStream_t s = G->stream[g];
        if (e > b) CKC(cudaMemcpyAsync(G->d_in[g], idata + b, sizeof(int) * (size_t)(e - b), cudaMemcpyHostToDevice, s));
        CKC(cudaStreamSynchronize(s));       // (as in the scan: all shards resident before the kernel timers start)
        return 0;
    });
    if (rc) return rc;
    rc = for_each_device(G, [&](int g) {
        int b, e;
        shard_of(n, g, K, b, e);
        cudaStream_t s = G->stream[g];
        int r = 0;
        CKC(cudaEventRecord(G->ev[4 * g + 0], s));
        r = b200_compact_i32_dev(G->sc[g], e - b, G->d_out[g], G->d_in[g], G->d_small[g], s);
        if (r) return r;
        if (K == 1) CKC(cudaEventRecord(G->ev[4 * g + 1], s));
        return 0;
    });
    if (rc) return rc;
    if (K > 1) {
        rc = gather_ints(G);                              // the one exchange: every shard's kept count
        if (rc) return rc;
        for (int g = 0; g < K; ++g) { OnDevice on(G->dev[g]); cudaEventRecord(G->ev[4 * g + 1], G->stream[g]); }
    }
    std::vector<int> counts((size_t)K, 0);
    {
        OnDevice on(G->dev[0]);
        CKC(cudaMemcpyAsync(counts.data(), K > 1 ? G->d_small[0] + 4 : G->d_small[0], sizeof(int) * (size_t)K, cudaMemcpyDeviceToHost, G->stream[0]));
        CKC(cudaStreamSynchronize(G->stream[0]));
    }
    std::vector<long long> off((size_t)K + 1, 0);
    for (int g = 0; g < K; ++g) off[g + 1] = off[g] + counts[g];
    rc = for_each_device(G, [&](int g) {                  // survivors of shard g belong at [off[g], off[g+1]) of the output
        cudaStream_t s = G->stream[g];
        if (counts[g] > 0) CKC(cudaMemcpyAsync(odata + off[g], G->d_out[g], sizeof(int) * (size_t)counts[g], cudaMemcpyDeviceToHost, s));
        CKC(cudaStreamSynchronize(s));
        return 0;
    });
    if (rc) return rc;
    *count = (int)off[K];
    if (kernel_ms) *kernel_ms = max_elapsed(G, 0, 1);
    if (exchange_ms && K > 1) *exchange_ms = max_elapsed(G, 2, 3);
    return 0;
}

const char* b200mg_error_string(int code)
{
    if (code == B200_E_NCCL + 99) return "b200mg: libnccl.so.2 could not be loaded (needed for more than one device)";
    if (code >= B200_E_NCCL && code < B200_E_NCCL + 99)
        return nccl().ok ? nccl().GetErrorString((ncclResult_t)(code - B200_E_NCCL)) : "b200mg: NCCL error";
    return b200pt_error_string(code);
}

}  // extern "C"
