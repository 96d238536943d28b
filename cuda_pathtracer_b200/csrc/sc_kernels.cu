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
#include "lookback.cuh"

namespace b200 {

namespace {

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

// Loads this thread's SC_VEC vectors of tile `tile`.  Out-of-range elements read as 0.
// `vec_ok`: the input pointer is 16-byte aligned (checked on the host).
template <bool VEC_OK>
__device__ __forceinline__ void load_tile(const int* __restrict__ in, unsigned n, unsigned tile,
                                          int warp, int lane, int4 (&v)[SC_VEC])
{
    const unsigned warp_base = tile * SC_TILE + warp * (32 * SC_ITEMS);
    const bool full = (tile + 1u) * (unsigned long long)SC_TILE <= n;
    if (VEC_OK && full) {
        const int4* p = reinterpret_cast<const int4*>(in + warp_base) + lane;
#pragma unroll
        for (int j = 0; j < SC_VEC; ++j) v[j] = ld_stream(p + j * 32);
    } else {
#pragma unroll
        for (int j = 0; j < SC_VEC; ++j) {
            const unsigned e = warp_base + (j * 32 + lane) * 4;
            v[j].x = e + 0 < n ? __ldg(in + e + 0) : 0;
            v[j].y = e + 1 < n ? __ldg(in + e + 1) : 0;
            v[j].z = e + 2 < n ? __ldg(in + e + 2) : 0;
            v[j].w = e + 3 < n ? __ldg(in + e + 3) : 0;
        }
    }
}

template <bool VEC_OK>
__device__ __forceinline__ void store_tile(int* __restrict__ out, unsigned n, unsigned tile,
                                           int warp, int lane, const int4 (&v)[SC_VEC])
{
    const unsigned warp_base = tile * SC_TILE + warp * (32 * SC_ITEMS);
    const bool full = (tile + 1u) * (unsigned long long)SC_TILE <= n;
    if (VEC_OK && full) {
        int4* p = reinterpret_cast<int4*>(out + warp_base) + lane;
#pragma unroll
        for (int j = 0; j < SC_VEC; ++j) st_stream(p + j * 32, v[j]);
    } else {
#pragma unroll
        for (int j = 0; j < SC_VEC; ++j) {
            const unsigned e = warp_base + (j * 32 + lane) * 4;
            if (e + 0 < n) out[e + 0] = v[j].x;
            if (e + 1 < n) out[e + 1] = v[j].y;
            if (e + 2 < n) out[e + 2] = v[j].z;
            if (e + 3 < n) out[e + 3] = v[j].w;
        }
    }
}

// Exclusive scan of one value per vector (s[j]) across the CTA in tile order.
// On return excl[j] = sum of everything in the tile that precedes vector j of this thread;
// returns the tile aggregate.  smem_warp: SC_THREADS/32 ints.  Contains one __syncthreads.
__device__ __forceinline__ int tile_exclusive_scan(const int (&s)[SC_VEC], int (&excl)[SC_VEC],
                                                   int* smem_warp, int warp, int lane)
{
    int inc[SC_VEC];
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) inc[j] = warp_inclusive_scan(s[j], lane);
    int carry = 0;
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) {
        excl[j] = carry + inc[j] - s[j];
        carry += __shfl_sync(FULL_MASK, inc[j], 31);
    }
    if (lane == 0) smem_warp[warp] = carry;
    __syncthreads();
    int warp_prefix = 0, aggregate = 0;
#pragma unroll
    for (int w = 0; w < SC_THREADS / 32; ++w) {
        const int t = smem_warp[w];
        if (w < warp) warp_prefix += t;
        aggregate += t;
    }
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) excl[j] += warp_prefix;
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

struct ScanSmem {
    int warp_sums[SC_THREADS / 32];
    int tile_exclusive;
    unsigned tile;
};

// hand out tile ids in start order; tickets[epoch&1] is ours, the other one is re-armed for the
// next launch (launches that share a workspace are stream-ordered).
__device__ __forceinline__ unsigned take_ticket(unsigned* tickets, unsigned epoch, unsigned* smem_tile)
{
    if (threadIdx.x == 0) {
        const unsigned t = atomicAdd(&tickets[epoch & 1u], 1u);
        if (t == 0) tickets[(epoch + 1u) & 1u] = 0u;
        *smem_tile = t;
    }
    __syncthreads();
    return *smem_tile;
}

// ------------------------------------------------------------------------------------------------
// k_scan_lookback_i32: one tile per CTA, single pass, 8 B/element.
// ------------------------------------------------------------------------------------------------
template <bool VEC_OK>
__global__ void __launch_bounds__(SC_THREADS)
k_scan_lookback_i32(const int* __restrict__ in, int* __restrict__ out, unsigned n,
                    unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets,
                    unsigned epoch, const int* __restrict__ d_seed, int* __restrict__ d_total)
{
    __shared__ ScanSmem sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = take_ticket(tickets, epoch, &sm.tile);

    int4 v[SC_VEC];
    load_tile<VEC_OK>(in, n, tile, warp, lane, v);
    int s[SC_VEC], excl[SC_VEC];
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
    const int aggregate = tile_exclusive_scan(s, excl, sm.warp_sums, warp, lane);

    if (warp == 0) {
        const int seed = d_seed ? __ldg(d_seed) : 0;
        const int e = lb_exclusive_prefix(desc, tile, epoch, aggregate, seed, lane);
        if (lane == 0) {
            sm.tile_exclusive = e;
            if (d_total && tile == gridDim.x - 1) *d_total = e + aggregate;
        }
    }
    __syncthreads();
    const int tile_exclusive = sm.tile_exclusive;
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) v[j] = scan_vec(v[j], tile_exclusive + excl[j]);
    store_tile<VEC_OK>(out, n, tile, warp, lane, v);
}

// ------------------------------------------------------------------------------------------------
// k_compact_i32: flag + scan + scatter in one pass.  Kept elements of a tile are first packed in
// shared memory in order, then written as one contiguous, coalesced run.
// ------------------------------------------------------------------------------------------------
struct CompactSmem {
    int packed[SC_TILE];
    ScanSmem scan;
};

template <bool VEC_OK>
__global__ void __launch_bounds__(SC_THREADS)
k_compact_i32(const int* __restrict__ in, int* __restrict__ out, unsigned n,
              unsigned long long* __restrict__ desc, unsigned* __restrict__ tickets, unsigned epoch,
              int* __restrict__ d_count)
{
    __shared__ CompactSmem sm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = take_ticket(tickets, epoch, &sm.scan.tile);

    int4 v[SC_VEC];
    load_tile<VEC_OK>(in, n, tile, warp, lane, v);   // out-of-range reads as 0 == dropped
    int s[SC_VEC], excl[SC_VEC];
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j)
        s[j] = (v[j].x != 0) + (v[j].y != 0) + (v[j].z != 0) + (v[j].w != 0);
    const int kept = tile_exclusive_scan(s, excl, sm.scan.warp_sums, warp, lane);

    if (warp == 0) {
        const int e = lb_exclusive_prefix(desc, tile, epoch, kept, 0, lane);
        if (lane == 0) {
            sm.scan.tile_exclusive = e;
            if (tile == gridDim.x - 1) *d_count = e + kept;
        }
    }
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) {
        int p = excl[j];
        if (v[j].x != 0) sm.packed[p++] = v[j].x;
        if (v[j].y != 0) sm.packed[p++] = v[j].y;
        if (v[j].z != 0) sm.packed[p++] = v[j].z;
        if (v[j].w != 0) sm.packed[p++] = v[j].w;
    }
    __syncthreads();
    int* dst = out + sm.scan.tile_exclusive;
    for (int i = threadIdx.x; i < kept; i += SC_THREADS) dst[i] = sm.packed[i];
}

// ------------------------------------------------------------------------------------------------
// Work-efficient three-phase scan: k_reduce_tiles (up-sweep), recursion on the tile sums,
// k_downsweep (per-tile scan seeded with the scanned tile sum).  12 B/element.
// ------------------------------------------------------------------------------------------------
template <bool VEC_OK>
__global__ void __launch_bounds__(SC_THREADS)
k_reduce_tiles(const int* __restrict__ in, unsigned n, int* __restrict__ tile_sums)
{
    __shared__ int warp_sums[SC_THREADS / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int4 v[SC_VEC];
    load_tile<VEC_OK>(in, n, blockIdx.x, warp, lane, v);
    int s = 0;
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) s += v[j].x + v[j].y + v[j].z + v[j].w;
    s = warp_sum(s);
    if (lane == 0) warp_sums[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < SC_THREADS / 32; ++w) t += warp_sums[w];
        tile_sums[blockIdx.x] = t;
    }
}

template <bool VEC_OK>
__global__ void __launch_bounds__(SC_THREADS)
k_downsweep(const int* __restrict__ in, int* __restrict__ out, unsigned n,
            const int* __restrict__ tile_offsets, const int* __restrict__ d_seed,
            int* __restrict__ d_total)
{
    __shared__ int warp_sums[SC_THREADS / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned tile = blockIdx.x;
    int4 v[SC_VEC];
    load_tile<VEC_OK>(in, n, tile, warp, lane, v);
    int s[SC_VEC], excl[SC_VEC];
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) s[j] = v[j].x + v[j].y + v[j].z + v[j].w;
    const int aggregate = tile_exclusive_scan(s, excl, warp_sums, warp, lane);
    int offset = tile_offsets ? __ldg(tile_offsets + tile) : 0;
    if (d_seed) offset += __ldg(d_seed);
    if (d_total && threadIdx.x == 0 && tile == gridDim.x - 1) *d_total = offset + aggregate;
#pragma unroll
    for (int j = 0; j < SC_VEC; ++j) v[j] = scan_vec(v[j], offset + excl[j]);
    store_tile<VEC_OK>(out, n, tile, warp, lane, v);
}

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

// ------------------------------------------------------------------------------------------------
// Whole-array sum (the per-shard total a multi-GPU scan exchanges): persistent grid-stride loop.
// ------------------------------------------------------------------------------------------------
template <bool VEC_OK>
__global__ void __launch_bounds__(SC_THREADS)
k_reduce_total(const int* __restrict__ in, unsigned n, int* __restrict__ d_total)
{
    __shared__ int warp_sums[SC_THREADS / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int acc = 0;
    const unsigned tiles = (n + SC_TILE - 1) / SC_TILE;
    for (unsigned tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        int4 v[SC_VEC];
        load_tile<VEC_OK>(in, n, tile, warp, lane, v);
#pragma unroll
        for (int j = 0; j < SC_VEC; ++j) acc += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    acc = warp_sum(acc);
    if (lane == 0) warp_sums[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < SC_THREADS / 32; ++w) t += warp_sums[w];
        atomicAdd(d_total, t);
    }
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
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in) && aligned16(d_out))
        k_scan_lookback_i32<true><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_seed, d_total);
    else
        k_scan_lookback_i32<false><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_seed, d_total);
    return cudaGetLastError();
}

cudaError_t launch_compact_lookback(const int* d_in, int* d_out, unsigned n, unsigned long long* desc,
                                    unsigned* tickets, unsigned epoch, int* d_count, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in))
        k_compact_i32<true><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    else
        k_compact_i32<false><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, desc, tickets, epoch, d_count);
    return cudaGetLastError();
}

cudaError_t launch_reduce_tiles(const int* d_in, unsigned n, int* d_tile_sums, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in))
        k_reduce_tiles<true><<<tiles, SC_THREADS, 0, s>>>(d_in, n, d_tile_sums);
    else
        k_reduce_tiles<false><<<tiles, SC_THREADS, 0, s>>>(d_in, n, d_tile_sums);
    return cudaGetLastError();
}

cudaError_t launch_downsweep(const int* d_in, int* d_out, unsigned n, const int* d_tile_offsets,
                             const int* d_seed, int* d_total, cudaStream_t s)
{
    const unsigned tiles = sc_num_tiles(n);
    if (aligned16(d_in) && aligned16(d_out))
        k_downsweep<true><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, d_tile_offsets, d_seed, d_total);
    else
        k_downsweep<false><<<tiles, SC_THREADS, 0, s>>>(d_in, d_out, n, d_tile_offsets, d_seed, d_total);
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
        k_reduce_total<true><<<grid, SC_THREADS, 0, s>>>(d_in, n, d_total);
    else
        k_reduce_total<false><<<grid, SC_THREADS, 0, s>>>(d_in, n, d_total);
    return cudaGetLastError();
}

}  // namespace b200
