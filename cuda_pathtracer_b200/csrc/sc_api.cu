// sc_api.cu -- the C ABI of include/b200_sc.h on top of the kernels in sc_kernels.cu.
//
// Host-side mirror of what the reference does around its kernels (allocation, H2D/D2H copies,
// CUDA-event timing of the kernels only: stream_compaction/stream_compaction/efficient.cu:152-176,
// naive.cu:36-66, thrust.cu:20-29), minus its per-call cudaMalloc/cudaFree/cudaGetDeviceProperties:
// the workspace is grow-only and lives in a context.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "../../include/b200_sc.h"
#include "sc_kernels.cuh"

using namespace b200;

struct b200_sc_ctx {
    int device = 0;
    int num_sms = 148;
    unsigned epoch = 0;
    unsigned long long* desc = nullptr;   // tile descriptors (look-back)
    size_t desc_cap = 0;                  // in words
    unsigned* tickets = nullptr;          // 4 counters: [0], [1] tile tickets by epoch parity (SC_LB_TICKET builds), [2], [3] k_compact_ws
    int* level_sums = nullptr;            // tile sums of every level of the work-efficient scan
    size_t level_cap = 0;                 // in ints
    int* scratch[2] = {nullptr, nullptr}; // naive ping-pong
    size_t scratch_cap = 0;               // in ints, each
    int* stage[2] = {nullptr, nullptr};   // staging for host-pointer calls (in, out)
    size_t stage_cap = 0;
    int* small = nullptr;                 // 4 ints: count / total slots for host-pointer calls
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int last_launches = 0;
    unsigned ring_min_n = SC_RING_MIN_N;  // from how many elements on the TMA-ring kernels run (b200_sc_ctx_set_ring_min_n)
};

namespace {

#define CK(x)                                   \
    do {                                        \
        cudaError_t e__ = (x);                  \
        if (e__ != cudaSuccess) return (int)e__; \
    } while (0)

int ensure(void** p, size_t* cap, size_t want_elems, size_t elem_bytes, bool zero)
{
    if (*cap >= want_elems && *p) return 0;
    if (*p) { cudaFree(*p); *p = nullptr; *cap = 0; }
    size_t bytes = want_elems * elem_bytes;
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); return e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e; }
    if (zero) {      // (null-stream memset: make it visible to launches on non-blocking caller streams)
        e = cudaMemset(*p, 0, bytes);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) return (int)e;
    }
    *cap = want_elems;
    return 0;
}

int ensure_lookback(b200_sc_ctx* c, unsigned n)
{
    const size_t tiles = sc_desc_words(n);
    if (c->desc_cap < tiles) {
        // new array is zero => every word reads as "epoch 0, not ready"; epochs start at 1
        size_t want = tiles < 16384 ? 16384 : tiles;
        int r = ensure((void**)&c->desc, &c->desc_cap, want, sizeof(unsigned long long), true);
        if (r) return r;
    }
    return 0;
}

// A context is used from one stream at a time (launches sharing its workspace must be stream-ordered, b200_sc.h);
// the wrap-around wipe is therefore enqueued on the launch stream, behind the launches that still use the old epochs.
// ping-pong buffers of the naive scan
int ensure_naive(b200_sc_ctx* c, unsigned n)
{
    const size_t want = ((size_t)n + 3) & ~(size_t)3;
    if (c->scratch_cap >= want) return 0;
    size_t cap0 = c->scratch_cap, cap1 = c->scratch_cap;
    int r = ensure((void**)&c->scratch[0], &cap0, want, sizeof(int), false);
    if (r) return r;
    r = ensure((void**)&c->scratch[1], &cap1, want, sizeof(int), false);
    if (r) return r;
    c->scratch_cap = want;
    return 0;
}

unsigned next_epoch(b200_sc_ctx* c, cudaStream_t s)
{
    c->epoch += 1;
    if (c->epoch >= (1u << 30)) {   // wrap: wipe descriptors once every 2^30 launches
        cudaMemsetAsync(c->desc, 0, c->desc_cap * sizeof(unsigned long long), s);
        // ... and BOTH ticket counters: 2^30 - 1 and 1 are both odd, so the first launch after the wrap would
        // otherwise reuse the counter the last launch before it has just run up (SC_LB_TICKET builds only)
        cudaMemsetAsync(c->tickets, 0, 2 * sizeof(unsigned), s);
        c->epoch = 1;
    }
    return c->epoch;
}

// work-efficient scan, recursive over levels; sums of level L live at level_sums + off
int efficient_rec(b200_sc_ctx* c, const int* in, int* out, unsigned n, int* sums, const int* d_seed,
                  int* d_total, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (tiles == 1) {
        CK(launch_downsweep(in, out, n, nullptr, d_seed, d_total, s));
        c->last_launches += 1;
        return 0;
    }
    CK(launch_reduce_tiles(in, n, sums, s));
    c->last_launches += 1;
    // scan the tile sums in place (each tile reads its inputs before it writes them)
    const size_t next = ((size_t)tiles + 3) & ~(size_t)3;   // keep every level 16-B aligned
    int r = efficient_rec(c, sums, sums, tiles, sums + next, nullptr, nullptr, s);
    if (r) return r;
    CK(launch_downsweep(in, out, n, sums, d_seed, d_total, s));
    c->last_launches += 1;
    return 0;
}

size_t efficient_level_ints(unsigned n)
{
    size_t total = 0;
    unsigned t = sc_num_tiles(n);
    while (t > 1) { total += ((size_t)t + 3) & ~(size_t)3; t = sc_num_tiles(t); }
    return total + 4;
}

int ilog2ceil_u(unsigned x)
{
    int lg = 0;
    while ((1ull << lg) < x) ++lg;
    return lg;
}

b200_sc_ctx* g_default = nullptr;
std::mutex g_mutex;

int default_ctx(b200_sc_ctx** out)
{
    if (!g_default) {
        int r = b200_sc_ctx_create(&g_default);
        if (r) return r;
    }
    *out = g_default;
    return 0;
}

bool is_device_ptr(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

}  // namespace

extern "C" {

int b200_sc_ctx_create(b200_sc_ctx** out)
{
    if (!out) return B200_E_BADARG;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); return B200_E_NODEVICE; }
    b200_sc_ctx* c = new (std::nothrow) b200_sc_ctx();
    if (!c) return B200_E_NOMEM;
    cudaError_t e = cudaGetDevice(&c->device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, c->device);
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->tickets, 4 * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMemset(c->tickets, 0, 4 * sizeof(unsigned));
    if (e == cudaSuccess) e = cudaMalloc((void**)&c->small, 4 * sizeof(int));
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&c->ev1);
    if (e != cudaSuccess) {          // nothing allocated so far is leaked
        cudaGetLastError();
        b200_sc_ctx_destroy(c);
        return e == cudaErrorMemoryAllocation ? B200_E_NOMEM : (int)e;
    }
    *out = c;
    return 0;
}

int b200_sc_ctx_destroy(b200_sc_ctx* c)
{
    if (!c) return 0;
    cudaFree(c->desc); cudaFree(c->tickets); cudaFree(c->level_sums);
    cudaFree(c->scratch[0]); cudaFree(c->scratch[1]);
    cudaFree(c->stage[0]); cudaFree(c->stage[1]); cudaFree(c->small);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    delete c;
    return 0;
}

int b200_sc_ctx_reserve(b200_sc_ctx* c, int n)
{
    if (!c || n < 0) return B200_E_BADARG;
    if (n == 0) return 0;
    int r = ensure_lookback(c, (unsigned)n);
    if (!r) r = ensure((void**)&c->level_sums, &c->level_cap, efficient_level_ints((unsigned)n), sizeof(int), false);
    return r;
}

int b200_scan_i32_dev(b200_sc_ctx* c, int n, int* d_out, const int* d_in, int variant,
                      const int* d_seed, int* d_total, void* stream)
{
    if (!c || n < 0 || (n > 0 && (!d_out || !d_in))) return B200_E_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    c->last_launches = 0;
    if (n == 0) {
        if (d_total) {
            if (d_seed) CK(cudaMemcpyAsync(d_total, d_seed, sizeof(int), cudaMemcpyDeviceToDevice, s));
            else CK(cudaMemsetAsync(d_total, 0, sizeof(int), s));
        }
        return 0;
    }
    const unsigned un = (unsigned)n;
    switch (variant) {
    case B200_SCAN_LOOKBACK:
    case B200_SCAN_LOOKBACK_TILE:
    case B200_SCAN_LOOKBACK_RING:
    case B200_SCAN_LOOKBACK_WAVE: {
        const bool ring_ok = scan_ring_usable(d_in, d_out);
        if (variant == B200_SCAN_LOOKBACK_RING && !ring_ok) return B200_E_BADARG;
        const bool wave_ok = scan_wave_usable(un, c->num_sms);
        if (variant == B200_SCAN_LOOKBACK_WAVE && !wave_ok) return B200_E_BADARG;
        bool ring = variant == B200_SCAN_LOOKBACK_RING || (variant == B200_SCAN_LOOKBACK && ring_ok && un >= c->ring_min_n);
        const bool wave = variant == B200_SCAN_LOOKBACK_WAVE || (variant == B200_SCAN_LOOKBACK && !ring && wave_ok && un >= SC_WAVE_MIN_N);
        int r = ensure_lookback(c, un);
        if (r) return r;
        const unsigned epoch = next_epoch(c, s);
        if (wave) {
            CK(launch_scan_wave(d_in, d_out, un, c->desc, epoch, d_seed, d_total, c->num_sms, s));
        } else {
            cudaError_t e = cudaErrorLaunchOutOfResources;
            if (ring) e = launch_scan_ring(d_in, d_out, un, c->desc, c->tickets, epoch, d_seed, d_total, c->num_sms, s);
            // a device whose shared memory cannot hold the ring (nothing was launched): the register-tile kernel, unless the ring
            // was asked for by name
            if (ring && e == cudaErrorLaunchOutOfResources && variant == B200_SCAN_LOOKBACK) { cudaGetLastError(); ring = false; }
            if (!ring) e = launch_scan_lookback(d_in, d_out, un, c->desc, c->tickets, epoch, d_seed, d_total, s);
            CK(e);
        }
        c->last_launches = 1;
        return 0;
    }
    case B200_SCAN_EFFICIENT: {
        int r = ensure((void**)&c->level_sums, &c->level_cap, efficient_level_ints(un), sizeof(int), false);
        if (r) return r;
        return efficient_rec(c, d_in, d_out, un, c->level_sums, d_seed, d_total, s);
    }
    case B200_SCAN_NAIVE: {
        const int passes = ilog2ceil_u(un);
        const int* src = d_in;
        if (passes > 0) {
            const int rr = ensure_naive(c, un);
            if (rr) return rr;
            for (int d = 0; d < passes; ++d) {
                int* dst = c->scratch[d & 1];
                CK(launch_naive_pass(un, 1u << d, dst, src, s));
                c->last_launches += ((1u << d) % 4u == 0u && un % 4u != 0u && un >= 4u) ? 2 : 1;
                src = dst;
            }
        }
        CK(launch_naive_shift(un, d_out, src, d_seed, d_total, s));
        c->last_launches += 1;
        return 0;
    }
    default:
        return B200_E_BADARG;
    }
}

int b200_compact_i32_dev(b200_sc_ctx* c, int n, int* d_out, const int* d_in, int* d_count, void* stream)
{
    if (!c || n < 0 || !d_count || (n > 0 && (!d_out || !d_in))) return B200_E_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    c->last_launches = 0;
    if (n == 0) { CK(cudaMemsetAsync(d_count, 0, sizeof(int), s)); return 0; }
    int r = ensure_lookback(c, (unsigned)n);
    if (r) return r;
    const unsigned epoch = next_epoch(c, s);
    cudaError_t e = cudaErrorLaunchOutOfResources;
    const bool ring = compact_ring_usable(d_in) && (unsigned)n >= c->ring_min_n;
    if (ring) e = launch_compact_ring(d_in, d_out, (unsigned)n, c->desc, c->tickets, epoch, d_count, c->num_sms, s);
    if (!ring || e == cudaErrorLaunchOutOfResources) {     // (no ring on this device: nothing was launched)
        if (ring) cudaGetLastError();
        e = launch_compact_lookback(d_in, d_out, (unsigned)n, c->desc, c->tickets, epoch, d_count, c->num_sms, s);
    }
    CK(e);
    c->last_launches = 1;
    return 0;
}

int b200_sc_ctx_set_ring_min_n(b200_sc_ctx* c, int min_n)
{
    if (!c || min_n < 0) return B200_E_BADARG;
    c->ring_min_n = min_n == 0 ? SC_RING_MIN_N : (unsigned)min_n;
    return 0;
}

int b200_reduce_i32_dev(b200_sc_ctx* c, int n, const int* d_in, int* d_total, void* stream)
{
    if (!c || n < 0 || !d_total || (n > 0 && !d_in)) return B200_E_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    c->last_launches = 0;
    CK(cudaMemsetAsync(d_total, 0, sizeof(int), s));
    if (n == 0) return 0;
    CK(launch_reduce_total(d_in, (unsigned)n, d_total, c->num_sms, s));
    c->last_launches = 1;
    return 0;
}

int b200_sc_last_launches(const b200_sc_ctx* c) { return c ? c->last_launches : 0; }

// ---- host-or-device pointer entry points ------------------------------------------------------

static int run_hostapi(int n, int* odata, const int* idata, int variant, bool compact, int* count,
                       float* kernel_ms)
{
    if (n < 0 || (n > 0 && (!odata || !idata)) || (compact && !count)) return B200_E_BADARG;
    std::lock_guard<std::mutex> lock(g_mutex);
    b200_sc_ctx* c = nullptr;
    int r = default_ctx(&c);
    if (r) return r;
    if (kernel_ms) *kernel_ms = 0.f;
    if (n == 0) { if (count) *count = 0; return 0; }

    const bool in_dev = is_device_ptr(idata), out_dev = is_device_ptr(odata);
    const int* d_in = idata;
    int* d_out = odata;
    if (!in_dev || !out_dev) {
        const size_t want = ((size_t)n + 3) & ~(size_t)3;
        if (c->stage_cap < want) {
            size_t cap0 = c->stage_cap, cap1 = c->stage_cap;
            r = ensure((void**)&c->stage[0], &cap0, want, sizeof(int), false);
            if (r) return r;
            r = ensure((void**)&c->stage[1], &cap1, want, sizeof(int), false);
            if (r) return r;
            c->stage_cap = want;
        }
    }
    if (!in_dev) {
        CK(cudaMemcpy(c->stage[0], idata, (size_t)n * sizeof(int), cudaMemcpyHostToDevice));
        d_in = c->stage[0];
    }
    if (!out_dev) d_out = c->stage[1];

    // the timer spans the kernels only (the reference's convention, efficient.cu:166-169): the workspace a first call of this
    // size needs is allocated before it starts
    r = b200_sc_ctx_reserve(c, n);
    if (!r && !compact && variant == B200_SCAN_NAIVE) r = ensure_naive(c, (unsigned)n);
    if (r) return r;
    CK(cudaEventRecord(c->ev0, 0));
    if (compact) r = b200_compact_i32_dev(c, n, d_out, d_in, c->small, nullptr);
    else r = b200_scan_i32_dev(c, n, d_out, d_in, variant, nullptr, nullptr, nullptr);
    if (r) return r;
    CK(cudaEventRecord(c->ev1, 0));
    CK(cudaEventSynchronize(c->ev1));
    if (kernel_ms) CK(cudaEventElapsedTime(kernel_ms, c->ev0, c->ev1));

    int kept = n;
    if (compact) {
        CK(cudaMemcpy(&kept, c->small, sizeof(int), cudaMemcpyDeviceToHost));
        *count = kept;
    }
    if (!out_dev && kept > 0)
        CK(cudaMemcpy(odata, d_out, (size_t)kept * sizeof(int), cudaMemcpyDeviceToHost));
    CK(cudaDeviceSynchronize());
    return 0;
}

int b200_scan_i32(int n, int* odata, const int* idata, int variant, float* kernel_ms)
{
    if (variant < B200_SCAN_LOOKBACK || variant > B200_SCAN_LOOKBACK_WAVE) return B200_E_BADARG;
    return run_hostapi(n, odata, idata, variant, false, nullptr, kernel_ms);
}

int b200_compact_i32(int n, int* odata, const int* idata, int* count, float* kernel_ms)
{
    return run_hostapi(n, odata, idata, 0, true, count, kernel_ms);
}

// ---- allocation-free scan (include/b200_sc.h: no counterpart in the reference) ----
// The look-back kernel's scratch in a file-scope device array instead of a cudaMalloc'd workspace.  In HOST code the
// name `g_noalloc_desc` is only a stand-in for the symbol, not a device address: the address is fetched once per
// device, with that device current, through cudaGetSymbolAddress, and THAT pointer goes to the kernel.
static_assert(SC_LB_TILE == 4096 && SC_LB_STRIDE == 16, "b200_sc.h states one 128-byte descriptor per 4096 ints");
constexpr size_t NOALLOC_DESC_WORDS = (size_t)(B200_SCAN_NOALLOC_MAX_N / SC_LB_TILE) * SC_LB_STRIDE;
static_assert(NOALLOC_DESC_WORDS * sizeof(unsigned long long) == B200_SCAN_NOALLOC_SCRATCH_BYTES, "header and array must agree");
static __device__ __align__(256) unsigned long long g_noalloc_desc[NOALLOC_DESC_WORDS];
static __device__ __align__(256) unsigned g_noalloc_tickets[64];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
struct NoallocDevice { unsigned long long* desc; unsigned* tickets; bool ready; };
NoallocDevice g_noalloc[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mutex;
}  // namespace

int b200_scan_i32_noalloc(int n, int* d_out, const int* d_in, void* stream)
{
    if (n < 0 || (n > 0 && (!d_out || !d_in))) return B200_E_BADARG;
    if (n > B200_SCAN_NOALLOC_MAX_N) return B200_E_TOOLARGE;
    if (n == 0) return 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return B200_E_NODEVICE; }
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return B200_E_BADARG;
    NoallocDevice nd;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc[dev].ready) {                  // set-up, once per device, with that device current
            void *p = nullptr, *t = nullptr;
            CK(cudaGetSymbolAddress(&p, g_noalloc_desc));
            CK(cudaGetSymbolAddress(&t, g_noalloc_tickets));
            g_noalloc[dev].desc = (unsigned long long*)p;
            g_noalloc[dev].tickets = (unsigned*)t;
            g_noalloc[dev].ready = true;
        }
        nd = g_noalloc[dev];
    }
    cudaStream_t s = (cudaStream_t)stream;
    // The array keeps the previous call's descriptors: the ones this launch reads must start as "nothing published"
    // (epoch 0).  The epoch is then a constant, so a captured graph replays correctly.
    CK(cudaMemsetAsync(nd.desc, 0, sc_desc_words((unsigned)n) * sizeof(unsigned long long), s));
    CK(cudaMemsetAsync(nd.tickets, 0, 2 * sizeof(unsigned), s));
    CK(launch_scan_lookback(d_in, d_out, (unsigned)n, nd.desc, nd.tickets, 1u, nullptr, nullptr, s));
    return 0;
}

const char* b200_error_string(int code)
{
    switch (code) {
    case B200_OK: return "ok";
    case B200_E_BADARG: return "b200: bad argument";
    case B200_E_NODEVICE: return "b200: no CUDA device (this library has no CPU fallback)";
    case B200_E_NOMEM: return "b200: out of device memory";
    case B200_E_TOOLARGE: return "b200: n exceeds B200_SCAN_NOALLOC_MAX_N (2^26 ints): the allocation-free scan has a fixed scratch array";
    default: return cudaGetErrorString((cudaError_t)code);
    }
}

}  // extern "C"
