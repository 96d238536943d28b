// sc_kernels.cuh -- internal launcher interface between sc_kernels.cu (device code) and
// sc_api.cu (the C ABI of include/b200_sc.h).  Not part of the public boundary.
#pragma once
#include <cuda_runtime.h>

namespace b200 {

constexpr int SC_THREADS = 256;                       // 8 warps per CTA
constexpr int SC_VEC = 4;                             // int4 loads per thread
constexpr int SC_ITEMS = SC_VEC * 4;                  // 16 ints per thread
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;        // 4096 ints = 16 KB per tile

inline unsigned sc_num_tiles(unsigned n) { return (unsigned)(((unsigned long long)n + SC_TILE - 1) / SC_TILE); }

// single-pass (decoupled look-back) scan and compact: tuned on a B200 with benchmarks/sc_tune.cu
// (profiles/r1_sc_tune_v3.jsonl).  At 2^28 ints: scan 128 x 8 -> 5.5 TB/s, compact 256 x 8 -> 4.7 TB/s;
// the decisive knob is SC_LB_STRIDE (one 128-byte line per descriptor), see lookback.cuh.
constexpr int SC_LB_THREADS = 128;                    // scan: 128 threads x 8 vectors = 4096 ints per tile
constexpr int SC_LB_VEC = 8;
constexpr int SC_LB_TILE = SC_LB_THREADS * SC_LB_VEC * 4;
constexpr int SC_CP_THREADS = 256;                    // compact: 256 threads x 8 vectors = 8192 ints per tile
constexpr int SC_CP_VEC = 8;
constexpr int SC_CP_TILE = SC_CP_THREADS * SC_CP_VEC * 4;
constexpr int SC_LB_WINDOW = 1;                       // look-back window = 32 * this descriptors
constexpr int SC_LB_BACKOFF_NS = 0;                   // sleep between polls of an unpublished descriptor
constexpr int SC_LB_STRIDE = 16;                      // 64-bit words between descriptors (128 B)
constexpr bool SC_LB_TICKET = false;                  // tile id = blockIdx.x (see sc_tmpl.cuh)
inline unsigned sc_lb_num_tiles(unsigned n) { return (unsigned)(((unsigned long long)n + SC_LB_TILE - 1) / SC_LB_TILE); }
inline unsigned sc_cp_num_tiles(unsigned n) { return (unsigned)(((unsigned long long)n + SC_CP_TILE - 1) / SC_CP_TILE); }
// descriptor words a launch over n elements needs (scan tiles are the smaller ones)
inline size_t sc_desc_words(unsigned n) { return (size_t)sc_lb_num_tiles(n) * SC_LB_STRIDE; }

// Single-pass decoupled look-back exclusive scan (B4/B6 replacement).  desc: >= num_tiles words,
// zero at allocation; tickets: 2 unsigned, zero at allocation; epoch: bumped per launch.
cudaError_t launch_scan_lookback(const int* d_in, int* d_out, unsigned n, unsigned long long* desc,
                                 unsigned* tickets, unsigned epoch, const int* d_seed, int* d_total,
                                 cudaStream_t s);

// The scan of a SMALL array in one wave of CTAs (sc_tmpl.cuh: k_scan_wave): every CTA reads all its predecessors' aggregates
// in one round trip.  Usable while n <= num_sms * 28672 ints; the library takes it between SC_WAVE_MIN_N and that limit.
constexpr int SC_WAVE_THREADS = 896;
constexpr unsigned SC_WAVE_MIN_N = (1u << 19) + 1u;   // (up to 2^19 the tile kernel is as fast or faster: profiles/r2z_sc_sweep_small.json)
bool scan_wave_usable(unsigned n, int num_sms);
cudaError_t launch_scan_wave(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned epoch,
                             const int* d_seed, int* d_total, int num_sms, cudaStream_t s);

// The same scan as a persistent TMA ring (sc_tmpl.cuh: k_scan_tma): bulk copies global -> shared -> global, the look-back
// off the critical path.  Needs 16-byte aligned pointers (scan_ring_usable); tickets[2], tickets[3] are its tile / done
// counters (zero between launches, shared with k_compact_ws: launches on one context are stream-ordered).  The library
// takes it from SC_RING_MIN_N elements on (below that the one-tile-per-CTA kernel's shorter ramp wins).
constexpr int SC_RING_STAGE_INTS = 8192;
constexpr int SC_RING_STAGES = 7;
constexpr int SC_RING_WARPS = 8;
constexpr int SC_RING_DELAY = 3;
constexpr int SC_RING_LB_WARPS = 4;
constexpr unsigned SC_RING_MIN_N = 1u << 24;
cudaError_t launch_scan_ring(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned* tickets,
                             unsigned epoch, const int* d_seed, int* d_total, int num_sms, cudaStream_t s);
bool scan_ring_usable(const int* d_in, const int* d_out);
// ... and the compaction on the same ring (k_compact_tma); only d_in must be 16-byte aligned
cudaError_t launch_compact_ring(const int* d_in, int* d_out, unsigned n, unsigned long long* desc, unsigned* tickets,
                                unsigned epoch, int* d_count, int num_sms, cudaStream_t s);
bool compact_ring_usable(const int* d_in);

// Single-pass compaction: keep in[i] != 0, order preserving (B5 replacement).
// (SC_COMPACT_WS: the persistent warp-specialised kernel, look-back off the critical path; tickets[2], tickets[3] are its
// tile / done counters, zero between launches)
#ifndef SC_COMPACT_WS
#define SC_COMPACT_WS 1
#endif
#ifndef SC_WS_THREADS
#define SC_WS_THREADS 256
#endif
#ifndef SC_WS_VEC
#define SC_WS_VEC 8
#endif
cudaError_t launch_compact_lookback(const int* d_in, int* d_out, unsigned n, unsigned long long* desc,
                                    unsigned* tickets, unsigned epoch, int* d_count, int num_sms, cudaStream_t s);

// Work-efficient three-phase scan pieces: per-tile reduce (up-sweep) and per-tile scan with a
// tile offset (down-sweep).  tile_offsets == nullptr => single tile, offset 0.
cudaError_t launch_reduce_tiles(const int* d_in, unsigned n, int* d_tile_sums, cudaStream_t s);
cudaError_t launch_downsweep(const int* d_in, int* d_out, unsigned n, const int* d_tile_offsets,
                             const int* d_seed, int* d_total, cudaStream_t s);

// Naive Hillis-Steele pass and the inclusive->exclusive shift (B3 replacement).
cudaError_t launch_naive_pass(unsigned n, unsigned offset, int* d_out, const int* d_in, cudaStream_t s);
cudaError_t launch_naive_shift(unsigned n, int* d_out, const int* d_incl, const int* d_seed,
                               int* d_total, cudaStream_t s);

// Whole-array sum into *d_total (must be zero on entry).
cudaError_t launch_reduce_total(const int* d_in, unsigned n, int* d_total, int num_sms, cudaStream_t s);

}  // namespace b200
