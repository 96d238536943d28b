/* b200_sc.h -- C ABI of the B200-native int32 scan / stream-compaction library.
 *
 * This is the drop-in boundary for the reference's scan/compact hot path:
 *
 *   void StreamCompaction::{CPU,Naive,Efficient,Thrust}::scan(int n, int* odata, const int* idata)
 *        reference: stream_compaction/stream_compaction/naive.h:7-9, efficient.h:7-11,
 *                   thrust.h:7-9 (and cpu.h:7-13 for the host oracle, which is NOT in this library)
 *   int  StreamCompaction::Efficient::compact(int n, int* odata, const int* idata)
 *        reference: stream_compaction/stream_compaction/efficient.h:10, efficient.cu:187-223
 *
 * Conventions kept from the reference:
 *   - exclusive prefix sum, int32 two's-complement wrap-around (cpu.cu:16-22);
 *   - compact keeps idata[i] != 0, order preserving, returns the count (cpu.cu:40-51);
 *   - odata != idata; both hold >= n ints; the call is synchronous (results valid on return);
 *   - pointers may be HOST pointers (harness, stream_compaction/src/main.cpp:31-138) or DEVICE
 *     pointers (the renderer calls Efficient::scan on device memory, path_tracer/src/pathtrace.cu:397);
 *   - the reported time spans the kernels only, not allocation or copies
 *     (efficient.cu:166-169 startGpuTimer/endGpuTimer placement).
 * Differences: errors are RETURNED (0 = ok, otherwise a cudaError_t value or a B200_E_* code);
 * nothing in here calls exit().  There is no CPU fallback: with no usable GPU every entry
 * point returns an error.
 *
 * Plain C: pointers and sizes only, no CUDA or torch types in the signatures (a stream is
 * passed as void*, which is what cudaStream_t is).
 */
#ifndef B200_SC_H
#define B200_SC_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* scan algorithm selector */
enum {
    B200_SCAN_LOOKBACK = 0,  /* single pass, decoupled look-back: 8 B/element (Thrust::scan name)     */
    B200_SCAN_EFFICIENT = 1, /* work-efficient reduce / scan-sums / down-sweep: 12 B/element
                                (Efficient::scan name; the reference's is efficient.cu:138-176)       */
    B200_SCAN_NAIVE = 2,     /* ilog2ceil(n) Hillis-Steele passes (Naive::scan, naive.cu:36-66)       */
    /* B200_SCAN_LOOKBACK picks between three kernels by size and pointer alignment; these name them (tests, benchmarks): */
    B200_SCAN_LOOKBACK_TILE = 3, /* one tile per CTA, loads and stores from registers                                    */
    B200_SCAN_LOOKBACK_RING = 4, /* persistent, one CTA per SM: TMA bulk copies through a shared-memory ring, look-back
                                    off the critical path; device pointers must be 16-byte aligned (B200_E_BADARG)      */
    B200_SCAN_LOOKBACK_WAVE = 5  /* small arrays: one wave of CTAs, every CTA sums ALL its predecessors' aggregates in one
                                    round trip; n <= #SMs * 28672 ints (B200_E_BADARG beyond)                           */
};

/* error codes outside the cudaError_t range */
enum {
    B200_OK = 0,
    B200_E_BADARG = 10001,
    B200_E_NODEVICE = 10002,
    B200_E_NOMEM = 10003,
    B200_E_TOOLARGE = 10005 /* b200_scan_i32_noalloc: n exceeds B200_SCAN_NOALLOC_MAX_N */
};

/* ---- host-or-device pointer entry points (what the reference's API binds to) ---------------- */

/* Exclusive scan of n ints.  kernel_ms (may be NULL) receives the CUDA-event time of the kernels. */
int b200_scan_i32(int n, int* odata, const int* idata, int variant, float* kernel_ms);

/* Stream compaction of n ints; *count receives the number kept.  Only the first *count
 * elements of odata are written (the reference copies n ints back, efficient.cu:217, whose tail
 * is uninitialised device memory -- callers only ever read the first count). */
int b200_compact_i32(int n, int* odata, const int* idata, int* count, float* kernel_ms);

/* ---- device-resident, asynchronous entry points (renderer, benchmarks, multi-GPU shards) ----- */

typedef struct b200_sc_ctx b200_sc_ctx; /* workspace: tile descriptors, level sums, scratch */

int b200_sc_ctx_create(b200_sc_ctx** ctx);
int b200_sc_ctx_destroy(b200_sc_ctx* ctx);

/* Grow the context's workspace (tile descriptors, level sums) to what a scan / compact of n ints needs, so that later
 * calls with up to n elements allocate nothing (the workspace otherwise grows on first use). */
int b200_sc_ctx_reserve(b200_sc_ctx* ctx, int n);

/* Tuning knob: from how many elements on b200_scan_i32_dev(.., B200_SCAN_LOOKBACK, ..) and b200_compact_i32_dev run their
 * persistent TMA-ring kernels (16-byte aligned device input -- and, for the scan, output -- required; otherwise, and below
 * the limit, the register-tile kernels run).  0 restores the library default (2^24); 1 = whenever the alignment allows;
 * INT_MAX = never.  The result is the same bit for bit whichever kernel runs. */
int b200_sc_ctx_set_ring_min_n(b200_sc_ctx* ctx, int min_n);

/* All pointers are device pointers; work is enqueued on `stream` (a cudaStream_t) and NOT
 * synchronised.  d_seed (may be NULL) is one device int added to every output element: the
 * exclusive prefix of the totals of the shards that precede this one (multi-GPU scan).
 * d_total (may be NULL) receives seed + sum of all n inputs. */
int b200_scan_i32_dev(b200_sc_ctx* ctx, int n, int* d_out, const int* d_in, int variant,
                      const int* d_seed, int* d_total, void* stream);

/* d_count receives the number of elements kept. */
int b200_compact_i32_dev(b200_sc_ctx* ctx, int n, int* d_out, const int* d_in, int* d_count,
                         void* stream);

/* Sum of n ints (int32 wrap) into *d_total: the per-shard total exchanged between GPUs. */
int b200_reduce_i32_dev(b200_sc_ctx* ctx, int n, const int* d_in, int* d_total, void* stream);

/* ---- allocation-free entry point ------------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Exclusive scan of n ints, DEVICE pointers, enqueued on `stream` and not synchronised -- the single-pass look-back
 * kernel of b200_scan_i32_dev(.., B200_SCAN_LOOKBACK, ..), same result bit for bit.  It needs no context: the kernel's
 * scratch (one 128-byte tile descriptor per 4096 ints) lives in a file-scope `static __device__` array of the library,
 * B200_SCAN_NOALLOC_SCRATCH_BYTES (2 MiB) on every GPU that loads the library, held until the process exits.  The
 * call allocates nothing (no cudaMalloc / cudaFree / allocating library routine): it clears the descriptors it is
 * about to use with cudaMemsetAsync on `stream` and launches the kernel.
 *   n > B200_SCAN_NOALLOC_MAX_N (2^26 ints = 256 MiB of input)  ->  B200_E_TOOLARGE, nothing is enqueued;
 *   every call on a device shares that device's array: calls that OVERLAP on the GPU (different streams without an
 *   ordering between them) race on it -- keep them on one stream or order them with events. */
enum {
    B200_SCAN_NOALLOC_MAX_N = 1 << 26,
    B200_SCAN_NOALLOC_SCRATCH_BYTES = (B200_SCAN_NOALLOC_MAX_N / 4096) * 128
};
int b200_scan_i32_noalloc(int n, int* d_out, const int* d_in, void* stream);

/* number of kernels the most recent call on this ctx launched (bench.py's gpu_launches) */
int b200_sc_last_launches(const b200_sc_ctx* ctx);

const char* b200_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* B200_SC_H */
