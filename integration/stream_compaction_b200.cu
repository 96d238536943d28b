// integration/stream_compaction_b200.cu -- the reference-side binding for the scan/compact path.
//
// This is the file a maintainer of the reference adds INSTEAD OF naive.cu, efficient.cu and
// thrust.cu (stream_compaction/stream_compaction/): it defines the same symbols those files
// define -- StreamCompaction::{Naive,Efficient,Thrust}::scan / ::timer and Efficient::compact
// (naive.h:7-9, efficient.h:7-11, thrust.h:7-9) -- against the reference's own common.h
// (PerformanceTimer) and forwards to the C ABI of include/b200_sc.h.  cpu.cu and common.cu of
// the reference stay as they are.  Same structure as the reference's host code
// (efficient.cu:152-176): copy in, GPU timer around the kernels only, copy out; host or device
// pointers are both accepted (pathtrace.cu:397 passes device pointers).
//
// oracle/Makefile compiles it against /root/reference together with the reference's UNMODIFIED
// test harness (stream_compaction/src/main.cpp) into oracle/_ref/dropin_sc_harness.
#include <stream_compaction/common.h>
#include <stream_compaction/efficient.h>
#include <stream_compaction/naive.h>
#include <stream_compaction/thrust.h>

#include "b200_sc.h"

namespace {

using StreamCompaction::Common::PerformanceTimer;

b200_sc_ctx* ctx()
{
    static b200_sc_ctx* c = nullptr;
    if (!c && b200_sc_ctx_create(&c) != 0) {
        fprintf(stderr, "b200: no usable CUDA device (there is no CPU fallback)\n");
        exit(EXIT_FAILURE);                      // the reference's own error convention (common.cu:3-15)
    }
    return c;
}

void fail(int code)
{
    fprintf(stderr, "b200: %s\n", b200_error_string(code));
    exit(EXIT_FAILURE);
}

bool on_device(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// n ints in, n ints out (scan) or count ints out (compact); returns count
int run(PerformanceTimer& t, int n, int* odata, const int* idata, int variant, bool compact)
{
    if (n <= 0) return 0;
    const bool din = on_device(idata), dout = on_device(odata);
    int *d_in = const_cast<int*>(idata), *d_out = odata, *d_count = nullptr;
    if (!din) {
        cudaMalloc(&d_in, sizeof(int) * (size_t)n);
        cudaMemcpy(d_in, idata, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice);
    }
    if (!dout) cudaMalloc(&d_out, sizeof(int) * (size_t)n);
    if (compact) cudaMalloc(&d_count, sizeof(int));
    t.startGpuTimer();
    const int r = compact ? b200_compact_i32_dev(ctx(), n, d_out, d_in, d_count, nullptr)
                          : b200_scan_i32_dev(ctx(), n, d_out, d_in, variant, nullptr, nullptr, nullptr);
    t.endGpuTimer();
    if (r) fail(r);
    int count = n;
    if (compact) {
        cudaMemcpy(&count, d_count, sizeof(int), cudaMemcpyDeviceToHost);
        cudaFree(d_count);
    }
    if (!dout) {
        cudaMemcpy(odata, d_out, sizeof(int) * (size_t)count, cudaMemcpyDeviceToHost);
        cudaFree(d_out);
    }
    if (!din) cudaFree(d_in);
    return count;
}

}  // namespace

namespace StreamCompaction {
namespace Naive {
PerformanceTimer& timer() { static PerformanceTimer t; return t; }
void scan(int n, int* odata, const int* idata) { run(timer(), n, odata, idata, B200_SCAN_NAIVE, false); }
}  // namespace Naive
namespace Efficient {
PerformanceTimer& timer() { static PerformanceTimer t; return t; }
void scan(int n, int* odata, const int* idata) { run(timer(), n, odata, idata, B200_SCAN_EFFICIENT, false); }
int compact(int n, int* odata, const int* idata) { return run(timer(), n, odata, idata, 0, true); }
}  // namespace Efficient
namespace Thrust {
PerformanceTimer& timer() { static PerformanceTimer t; return t; }
void scan(int n, int* odata, const int* idata) { run(timer(), n, odata, idata, B200_SCAN_LOOKBACK, false); }
}  // namespace Thrust
}  // namespace StreamCompaction
