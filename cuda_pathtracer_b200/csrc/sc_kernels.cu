// sc_kernels.cu -- hand-written sm_100a kernels of the int32 scan / stream-compaction library.
//
// Replaces: Efficient::gpu_scan_block + gpu_prefix_add (+ recursion)
// (stream_compaction/stream_compaction/efficient.cu:45-106,138-149), Common::kernMapToBoolean +
// kernScatter (common.cu:25-46), Naive::gpu_naive_scan + gpu_incl2excl_pfxsum (naive.cu:15-32)
// and the thrust::exclusive_scan call (thrust.cu:25).  No Thrust/CUB anywhere.
//
// Roofline: HBM.  Algorithmic bytes: look-back scan 8 B/element (read once, write once);
// compact 4 B read + 4 B per kept element; work-efficient scan 12 B/element; naive
// 8 B * ilog2ceil(n) + 8 B per element.
//
// Data layout inside a tile (SC_TILE = 4096 ints, 8 warps): warp w owns 512 consecutive ints;
// its j-th 128-bit vector for lane l is vector index j*32 + l, so every warp-wide load/store
// instruction covers 512 contiguous bytes.  The order of elements is therefore (warp, j, lane,
// component); the scan below is carried out in exactly that order.
#include "sc_kernels.cuh"
#include "sc_tmpl.cuh"

namespace b200 {

namespace {

// The tuned configurations (benchmarks/sc_tune.cu on a B200, profiles/): threads x 128-bit vectors
// per thread, look-back window = 32 * LBW descriptors.
using CfgLB = sct::Cfg<SC_LB_THREADS, SC_LB_VEC>;      // single-pass scan
using CfgCP = sct::Cfg<SC_CP_THREADS, SC_CP_VEC>;      // single-pass compact
using CfgWE = sct::Cfg<SC_THREADS, SC_VEC>;            // work-efficient reduce / down-sweep, reduce
constexpr int LBW = SC_LB_WINDOW;
constexpr bool TICKET = SC_LB_TICKET;
constexpr int BACKOFF = SC_LB_BACKOFF_NS;
constexpr int STRIDE = SC_LB_STRIDE;

// ------------------------------------------------------------------------------------------------
// Naive scan: one Hillis-Steele pass  o[i] = in[i] + (i >= off ? in[i-off] : 0).
// Buffers are library-owned (16-B aligned); passes with off % 4 == 0 move 128-bit vectors.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_naive_pass_vec(unsigned n4, unsigned off4, int4* __restrict__ out, const int4* __restrict__ in)
{
    const unsigned q = blockIdx.x * 256u + threadIdx.x;
    if (q >= n4) return;
    int4 a = in[q];
    if (q >= off4) {
        const int4 b = in[q - off4];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    out[q] = a;
}

__global__ void __launch_bounds__(256)
k_naive_pass_scalar(unsigned begin, unsigned n, unsigned off, int* __restrict__ out,
                    const int* __restrict__ in)
{
    const unsigned i = begin + blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    int a = in[i];
    if (i >= off) a += in[i - off];
    out[i] = a;
}

__global__ void __launch_bounds__(256)
k_naive_shift(unsigned n, int* __restrict__ out, const int* __restrict__ incl,
              const int* __restrict__ d_seed, int* __restrict__ d_total)
{
    const unsigned i = blockIdx.x * 256u + threadIdx.x;
    if (i >= n) return;
    const int seed = d_seed ? __ldg(d_seed) : 0;
    out[i] = seed + (i == 0 ? 0 : incl[i - 1]);
    if (d_total && i == n - 1) *d_total = seed + incl[i];
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
cudaError_t launch_scan_lookback(const int* d_in, int* d_out, unsigned n, unsigned long long* desc,
                                 unsigned* tickets, unsigned epoch, const int* d_seed, int* d_total,
                                 cudaStream_t s)
{
    const unsigned tiles = sc_lb_num_tiles(n);
    if (aligned16(d_in) && aligned16(d_out))
        sct::k_scan_lookback<CfgLB, true, LBW, TICKET, BACKOFF, STRIDE><<<tiles, CfgLB::threads, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_seed, d_total);
    else
        sct::k_scan_lookback<CfgLB, false, LBW, TICKET, BACKOFF, STRIDE><<<tiles, CfgLB::threads, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_seed, d_total);
    return cudaGetLastError();
}

// The single-wave scan (sct::k_scan_wave): 896 threads x {1, 2, 4, 8} vectors, the smallest tile that covers n with at most
// one CTA per SM.  Returns cudaErrorInvalidValue when n is too large for one wave (the caller then takes another kernel).
template <int VEC>
static cudaError_t launch_wave_cfg(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned epoch,
                                   const int* d_seed, int* d_total, cudaStream_t s)
{
    using C = sct::Cfg<SC_WAVE_THREADS, VEC>;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    if (aligned16(d_in) && aligned16(d_out))
        sct::k_scan_wave<C, true, STRIDE><<<tiles, C::threads, 0, s>>>(d_in, d_out, n, desc, epoch, d_seed, d_total);
    else
        sct::k_scan_wave<C, false, STRIDE><<<tiles, C::threads, 0, s>>>(d_in, d_out, n, desc, epoch, d_seed, d_total);
    return cudaGetLastError();
}

bool scan_wave_usable(unsigned n, int num_sms)
{
    return num_sms > 0 && (unsigned long long)n <= (unsigned long long)num_sms * (SC_WAVE_THREADS * 8 * 4);
}

cudaError_t launch_scan_wave(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned epoch,
                             const int* d_seed, int* d_total, int num_sms, cudaStream_t s)
{
    if (!scan_wave_usable(n, num_sms)) return cudaErrorInvalidValue;
    const unsigned long long per_sm = ((unsigned long long)n + num_sms - 1) / num_sms;
    if (per_sm <= SC_WAVE_THREADS * 1 * 4) return launch_wave_cfg<1>(d_in, d_out, n, desc, epoch, d_seed, d_total, s);
    if (per_sm <= SC_WAVE_THREADS * 2 * 4) return launch_wave_cfg<2>(d_in, d_out, n, desc, epoch, d_seed, d_total, s);
    if (per_sm <= SC_WAVE_THREADS * 4 * 4) return launch_wave_cfg<4>(d_in, d_out, n, desc, epoch, d_seed, d_total, s);
    return launch_wave_cfg<8>(d_in, d_out, n, desc, epoch, d_seed, d_total, s);
}

// The persistent TMA-ring scan (sct::k_scan_tma): one CTA per SM, tuned on a B200 with benchmarks/sc_tune.cu
// (profiles/r2z_sc_tune_tma_*.jsonl): 8192-int stages x 7 (224 KB of shared memory), 8 compute warps, pass B three tiles
// behind pass A, 4 look-back warps.
cudaError_t launch_scan_ring(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned* tickets,
                             unsigned epoch, const int* d_seed, int* d_total, int num_sms, cudaStream_t s)
{
    constexpr int STAGE = SC_RING_STAGE_INTS, STAGES = SC_RING_STAGES, CW = SC_RING_WARPS, DELAY = SC_RING_DELAY, NLB = SC_RING_LB_WARPS;
    auto k = sct::k_scan_tma<STAGE, STAGES, CW, DELAY, NLB, LBW, STRIDE>;
    constexpr size_t smem = sct::scan_tma_smem_bytes<STAGE, STAGES>();
    constexpr int threads = sct::scan_tma_threads<STAGE, CW, NLB>();
    if (!aligned16(d_in) || !aligned16(d_out)) return cudaErrorMisalignedAddress;
    static int per_sm_of[64] = {0};          // function attributes are per DEVICE (a process may drive several)
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (per_sm_of[dev] == 0) {
        int a = 0;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, k, threads, smem);
        if (e == cudaErrorInvalidValue) { cudaGetLastError(); a = 0; e = cudaSuccess; }   // more shared memory than the device has
        if (e != cudaSuccess) return e;
        per_sm_of[dev] = a > 0 ? a : -1;
    }
    if (per_sm_of[dev] <= 0) return cudaErrorLaunchOutOfResources;
    const unsigned tiles = (unsigned)(((unsigned long long)n + STAGE - 1) / STAGE);
    unsigned grid = (unsigned)(per_sm_of[dev] * num_sms);
    if (grid > tiles) grid = tiles;
    k<<<grid, threads, smem, s>>>(d_in, d_out, n, desc, tickets, epoch, d_seed, d_total);
    return cudaGetLastError();
}

bool scan_ring_usable(const int* d_in, const int* d_out) { return aligned16(d_in) && aligned16(d_out); }

// The compaction on the TMA ring (sct::k_compact_tma), same shape as the scan's: 0.871 of the HBM peak at 2^28 against 0.756
// for k_compact_ws (profiles/r2z_sc_tune_tma_compact_2p28.jsonl).
cudaError_t launch_compact_ring(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned* tickets,
                                unsigned epoch, int* d_count, int num_sms, cudaStream_t s)
{
    constexpr int STAGE = SC_RING_STAGE_INTS, STAGES = SC_RING_STAGES, CW = SC_RING_WARPS, DELAY = SC_RING_DELAY, NLB = SC_RING_LB_WARPS;
    auto k = sct::k_compact_tma<STAGE, STAGES, CW, DELAY, NLB, LBW, STRIDE>;
    constexpr size_t smem = sct::compact_tma_smem_bytes<STAGE, STAGES>();
    constexpr int threads = sct::scan_tma_threads<STAGE, CW, NLB>();
    if (!aligned16(d_in)) return cudaErrorMisalignedAddress;
    static int per_sm_of[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (per_sm_of[dev] == 0) {
        int a = 0;
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, k, threads, smem);
        if (e == cudaErrorInvalidValue) { cudaGetLastError(); a = 0; e = cudaSuccess; }   // more shared memory than the device has
        if (e != cudaSuccess) return e;
        per_sm_of[dev] = a > 0 ? a : -1;
    }
    if (per_sm_of[dev] <= 0) return cudaErrorLaunchOutOfResources;
    const unsigned tiles = (unsigned)(((unsigned long long)n + STAGE - 1) / STAGE);
    unsigned grid = (unsigned)(per_sm_of[dev] * num_sms);
    if (grid > tiles) grid = tiles;
    k<<<grid, threads, smem, s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    return cudaGetLastError();
}

bool compact_ring_usable(const int* d_in) { return aligned16(d_in); }

cudaError_t launch_compact_lookback(const int* d_in, int* d_out, unsigned n, unsigned long long* desc,
                                    unsigned* tickets, unsigned epoch, int* d_count, int num_sms, cudaStream_t s)
{
#if SC_COMPACT_WS
    using C = sct::Cfg<SC_WS_THREADS, SC_WS_VEC>;
    constexpr size_t smem = sct::compact_ws_smem_bytes<C>();
    // as many CTAs as are resident at once.  Function attributes are per DEVICE: set up once for each device that launches
    // (a process may drive several, include/b200_mg.h)
    static int per_sm_of[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (per_sm_of[dev] == 0) {
        int a = 0, b = 0;
        cudaError_t e = cudaFuncSetAttribute(sct::k_compact_ws<C, true, LBW, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(sct::k_compact_ws<C, false, LBW, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, sct::k_compact_ws<C, true, LBW, STRIDE>, C::threads + 32, smem);
        if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sct::k_compact_ws<C, false, LBW, STRIDE>, C::threads + 32, smem);
        if (e != cudaSuccess) return e;
        per_sm_of[dev] = (a < b ? a : b) > 0 ? (a < b ? a : b) : -1;
    }
    const int per_sm = per_sm_of[dev];
    if (per_sm <= 0) return cudaErrorLaunchOutOfResources;
    const unsigned tiles = (unsigned)(((unsigned long long)n + C::tile - 1) / C::tile);
    unsigned grid = (unsigned)(per_sm * num_sms);
    if (grid > tiles) grid = tiles;
    // (the descriptor array is sized for the smaller scan tiles: sc_desc_words)
    if (aligned16(d_in))
        sct::k_compact_ws<C, true, LBW, STRIDE><<<grid, C::threads + 32, smem, s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    else
        sct::k_compact_ws<C, false, LBW, STRIDE><<<grid, C::threads + 32, smem, s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    return cudaGetLastError();
#else
    (void)num_sms;
    const unsigned tiles = sc_cp_num_tiles(n);
    static const cudaError_t attr = [] {
        cudaError_t e = cudaFuncSetAttribute(sct::k_compact<CfgCP, true, LBW, TICKET, BACKOFF, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)sct::compact_smem_bytes<CfgCP>());
        if (e != cudaSuccess) return e;
        return cudaFuncSetAttribute(sct::k_compact<CfgCP, false, LBW, TICKET, BACKOFF, STRIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)sct::compact_smem_bytes<CfgCP>());
    }();
    if (attr != cudaSuccess) return attr;
    if (aligned16(d_in))
        sct::k_compact<CfgCP, true, LBW, TICKET, BACKOFF, STRIDE><<<tiles, CfgCP::threads, sct::compact_smem_bytes<CfgCP>(), s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    else
        sct::k_compact<CfgCP, false, LBW, TICKET, BACKOFF, STRIDE><<<tiles, CfgCP::threads, sct::compact_smem_bytes<CfgCP>(), s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    return cudaGetLastError();
#endif
}

cudaError_t launch_reduce_tiles(const int* d_in, unsigned n, int* d_tile_sums, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in))
        sct::k_reduce_tiles<CfgWE, true><<<tiles, SC_THREADS, 0, s>>>(d_in, n, d_tile_sums);
    else
        sct::k_reduce_tiles<CfgWE, false><<<tiles, SC_THREADS, 0, s>>>(d_in, n, d_tile_sums);
    return cudaGetLastError();
}

cudaError_t launch_downsweep(const int* d_in, int* d_out, unsigned n, const int* d_tile_offsets,
                             const int* d_seed, int* d_total, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in) && aligned16(d_out))
        sct::k_downsweep<CfgWE, true><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, d_tile_offsets, d_seed, d_total);
    else
        sct::k_downsweep<CfgWE, false><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, d_tile_offsets, d_seed, d_total);
    return cudaGetLastError();
}

cudaError_t launch_naive_pass(unsigned n, unsigned offset, int* d_out, const int* d_in, cudaStream_t s)
{
    const bool vec = (offset % 4u == 0u) && aligned16(d_in) && aligned16(d_out);
    unsigned scalar_begin = 0;
    if (vec) {
        const unsigned n4 = n / 4u;
        if (n4) {
            k_naive_pass_vec<<<(n4 + 255u) / 256u, 256, 0, s>>>(
                n4, offset / 4u, reinterpret_cast<int4*>(d_out), reinterpret_cast<const int4*>(d_in));
        }
        scalar_begin = n4 * 4u;
    }
    if (scalar_begin < n) {
        const unsigned m = n - scalar_begin;
        k_naive_pass_scalar<<<(m + 255u) / 256u, 256, 0, s>>>(scalar_begin, n, offset, d_out, d_in);
    }
    return cudaGetLastError();
}

cudaError_t launch_naive_shift(unsigned n, int* d_out, const int* d_incl, const int* d_seed,
                               int* d_total, cudaStream_t s)
{
    k_naive_shift<<<(n + 255u) / 256u, 256, 0, s>>>(n, d_out, d_incl, d_seed, d_total);
    return cudaGetLastError();
}

cudaError_t launch_reduce_total(const int* d_in, unsigned n, int* d_total, int num_sms, cudaStream_t s)
{
    unsigned grid = sc_num_tiles(n);
    const unsigned cap = (unsigned)num_sms * 8u;
    if (grid > cap) grid = cap;
    if (aligned16(d_in))
        sct::k_reduce_total<CfgWE, true><<<grid, SC_THREADS, 0, s>>>(d_in, n, d_total);
    else
        sct::k_reduce_total<CfgWE, false><<<grid, SC_THREADS, 0, s>>>(d_in, n, d_total);
    return cudaGetLastError();
}

}  // namespace b200
