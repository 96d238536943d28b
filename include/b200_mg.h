/* b200_mg.h -- C ABI of the multi-GPU layer: ONE process, one host thread per device, NCCL over NVLink.
 *
 * The reference is single-GPU (path_tracer/src/main.cpp:34-47 picks device 0); BASELINE.json's north_star asks for
 * the two hot paths to be partitioned over the GPUs of one box where they shard naturally:
 *
 *   path tracer   iterations are independent -- the seed depends on (iteration, index, depth) only
 *                 (path_tracer/src/pathtrace.cu:57-62) -- so device g of K traces iterations first+g, first+g+K, ...
 *                 of a replicated scene into a private accumulation image, and ONE ncclAllReduce(sum) of the
 *                 3*W*H floats closes the job (or a preview interval).  Same path set as one GPU; the image differs
 *                 only by fp32 re-association.
 *   scan          contiguous 16-byte-aligned shards; every device reduces its shard to one int, ONE ncclAllGather
 *                 of those totals, and each device scans its shard seeded with the exclusive prefix of the totals
 *                 before it (int32 adds are associative mod 2^32: bit-exact).
 *   compact       same shards; ONE ncclAllGather of the kept counts gives every shard's offset in the output.
 *
 * The entry points mirror include/b200_pt.h (b200pt_*) and include/b200_sc.h (b200_scan_i32 / b200_compact_i32):
 * same argument meaning, host pointers in and out, synchronous, errors returned (0 = ok, else cudaError_t,
 * B200_E_* or B200_E_NCCL + ncclResult_t).  With one device no collective is issued.  No CPU fallback.
 * Plain C: pointers and sizes only.
 */
#ifndef B200_MG_H
#define B200_MG_H

#include <stddef.h>
#include <stdint.h>

#include "b200_pt.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { B200_E_NCCL = 20000 }; /* + ncclResult_t */

typedef struct b200mg b200mg;       /* the device group: NCCL communicators, one stream and one scan context per device */
typedef struct b200mg_pt b200mg_pt; /* a renderer replicated over the group */

/* devices == NULL: CUDA devices 0 .. n_devices-1.  n_devices <= 0: every visible device. */
int b200mg_create(int n_devices, const int* devices, b200mg** out);
int b200mg_destroy(b200mg* g);
int b200mg_device_count(const b200mg* g);

/* ---- path tracer (pathtrace.h:6-9 over K devices) ------------------------------------------------------------- */
int b200mg_pt_create(b200mg* g, const b200pt_scene* scene, const b200pt_scene_ext* ext, b200mg_pt** out);
int b200mg_pt_destroy(b200mg_pt* r);
int b200mg_pt_set_options(b200mg_pt* r, const b200pt_options* o);
/* zero every device's accumulation image and statistics */
int b200mg_pt_reset(b200mg_pt* r);
/* restartable rendering: device 0's image becomes host_rgb (3*W*H floats), the others start from zero */
int b200mg_pt_write_sum(b200mg_pt* r, const float* host_rgb);
/* Trace iterations first_iter .. first_iter + n_iters - 1 (1-based): device g takes first_iter + g, + K, ...  One host
 * thread per device enqueues its share; returns when everything is enqueued (asynchronous on the devices). */
int b200mg_pt_render(b200mg_pt* r, int first_iter, int n_iters);
/* ONE ncclAllReduce(sum) of the private images into a second buffer on every device (the private sums stay valid, so
 * rendering can go on afterwards), then the reduced float SUM image (3*W*H floats) to host memory.  Synchronises.
 * allreduce_ms (may be NULL): device time of the collective, max over devices. */
int b200mg_pt_reduce_read_sum(b200mg_pt* r, float* host_rgb, float* allreduce_ms);
/* sendImageToPBO (pathtrace.cu:65-86) of the most recently REDUCED image on device 0, into host memory (W*H*4 bytes) */
int b200mg_pt_preview_u8(b200mg_pt* r, int iter_divisor, void* host_rgba8);
/* statistics summed over the devices (traced_depth: maximum; last_render_ms: maximum) */
int b200mg_pt_get_stats(b200mg_pt* r, b200pt_stats* total);

/* ---- scan / compact over K devices (naive.h / efficient.h / thrust.h signatures, HOST pointers) -------------------- */
/* kernel_ms (may be NULL): device time from the first shard kernel to the last, collectives included, copies excluded
 * (the reference's timing convention, efficient.cu:166-169), max over devices.  exchange_ms (may be NULL): the all-gather. */
int b200mg_scan_i32(b200mg* g, int n, int* odata, const int* idata, float* kernel_ms, float* exchange_ms);
int b200mg_compact_i32(b200mg* g, int n, int* odata, const int* idata, int* count, float* kernel_ms, float* exchange_ms);

const char* b200mg_error_string(int code);

#ifdef __cplusplus
}
#endif
#endif /* B200_MG_H */
