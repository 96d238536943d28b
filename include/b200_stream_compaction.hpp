// b200_stream_compaction.hpp -- the reference's C++ surface of the scan / compact library on top
// of the C ABI (include/b200_sc.h).  Header only; link with -lb200sc.
//
// Same namespaces, names, argument meaning and timer contract as the reference
// (stream_compaction/stream_compaction/{common,cpu,naive,efficient,thrust}.h):
//
//   StreamCompaction::Common::PerformanceTimer      common.h:46-130   (B7)
//   ilog2 / ilog2ceil                               common.h:21-31    (B8)
//   StreamCompaction::Naive::scan / timer           naive.h:7-9       (B3)  -> B200_SCAN_NAIVE
//   StreamCompaction::Efficient::scan/compact/timer efficient.h:7-11  (B4, B5) -> B200_SCAN_EFFICIENT, compact
//   StreamCompaction::Thrust::scan / timer          thrust.h:7-9      (B6)  -> B200_SCAN_LOOKBACK
//   StreamCompaction::CPU::scan / compactWithoutScan / compactWithScan / timer   cpu.h:7-13 (B1, B2)
//
// odata / idata may be host or device pointers (the renderer passes device pointers,
// path_tracer/src/pathtrace.cu:397); calls are synchronous; the GPU timer spans the kernels only
// (efficient.cu:166-169).  The CPU namespace is the reference's own serial host API -- it exists
// because callers (the test harness) use it as the ground truth; the GPU entry points never fall
// back to it: without a CUDA device they throw std::runtime_error.
#pragma once
#include <chrono>
#include <stdexcept>
#include <string>

#include "b200_sc.h"

inline int ilog2(int x)
{
    int lg = 0;
    while (x >>= 1) ++lg;
    return lg;
}
inline int ilog2ceil(int x) { return x == 1 ? 0 : ilog2(x - 1) + 1; }

namespace StreamCompaction {
namespace Common {

// Same contract as the reference's class: start/end pairs, std::runtime_error on misuse, elapsed
// milliseconds of the previous operation.  The GPU time is the CUDA-event time the library
// measured around its kernels.
class PerformanceTimer {
public:
    void startCpuTimer()
    {
        if (cpu_started_) throw std::runtime_error("CPU timer already started");
        cpu_started_ = true;
        t0_ = std::chrono::high_resolution_clock::now();
    }
    void endCpuTimer()
    {
        const auto t1 = std::chrono::high_resolution_clock::now();
        if (!cpu_started_) throw std::runtime_error("CPU timer not started");
        cpu_ms_ = (float)std::chrono::duration<double, std::milli>(t1 - t0_).count();
        cpu_started_ = false;
    }
    void startGpuTimer()
    {
        if (gpu_started_) throw std::runtime_error("GPU timer already started");
        gpu_started_ = true;
    }
    void endGpuTimer()
    {
        if (!gpu_started_) throw std::runtime_error("GPU timer not started");
        gpu_started_ = false;
    }
    void setGpuElapsed(float ms) { gpu_ms_ = ms; }
    float getCpuElapsedTimeForPreviousOperation() const { return cpu_ms_; }
    float getGpuElapsedTimeForPreviousOperation() const { return gpu_ms_; }

    PerformanceTimer() = default;
    PerformanceTimer(const PerformanceTimer&) = delete;
    PerformanceTimer& operator=(const PerformanceTimer&) = delete;

private:
    std::chrono::high_resolution_clock::time_point t0_;
    bool cpu_started_ = false, gpu_started_ = false;
    float cpu_ms_ = 0.f, gpu_ms_ = 0.f;
};

inline void check(int code)
{
    if (code != 0) throw std::runtime_error(std::string("StreamCompaction: ") + b200_error_string(code));
}

inline void gpu_scan(PerformanceTimer& t, int n, int* odata, const int* idata, int variant)
{
    float ms = 0.f;
    t.startGpuTimer();
    const int r = b200_scan_i32(n, odata, idata, variant, &ms);
    t.endGpuTimer();
    check(r);
    t.setGpuElapsed(ms);
}

}  // namespace Common

namespace Naive {
inline Common::PerformanceTimer& timer() { static Common::PerformanceTimer t; return t; }
inline void scan(int n, int* odata, const int* idata) { Common::gpu_scan(timer(), n, odata, idata, B200_SCAN_NAIVE); }
}  // namespace Naive

namespace Efficient {
inline Common::PerformanceTimer& timer() { static Common::PerformanceTimer t; return t; }
inline void scan(int n, int* odata, const int* idata) { Common::gpu_scan(timer(), n, odata, idata, B200_SCAN_EFFICIENT); }
inline int compact(int n, int* odata, const int* idata)
{
    float ms = 0.f;
    int count = 0;
    timer().startGpuTimer();
    const int r = b200_compact_i32(n, odata, idata, &count, &ms);
    timer().endGpuTimer();
    Common::check(r);
    timer().setGpuElapsed(ms);
    return count;
}
}  // namespace Efficient

namespace Thrust {
inline Common::PerformanceTimer& timer() { static Common::PerformanceTimer t; return t; }
inline void scan(int n, int* odata, const int* idata) { Common::gpu_scan(timer(), n, odata, idata, B200_SCAN_LOOKBACK); }
}  // namespace Thrust

namespace CPU {
inline Common::PerformanceTimer& timer() { static Common::PerformanceTimer t; return t; }

// exclusive prefix sum, int32 wrap-around (cpu.cu:16-33)
inline void scan(int n, int* odata, const int* idata)
{
    timer().startCpuTimer();
    unsigned run = 0;
    for (int i = 0; i < n; ++i) {
        odata[i] = (int)run;
        run += (unsigned)idata[i];
    }
    timer().endCpuTimer();
}

// order-preserving removal of zeros; returns the number kept (cpu.cu:40-51)
inline int compactWithoutScan(int n, int* odata, const int* idata)
{
    timer().startCpuTimer();
    int kept = 0;
    for (int i = 0; i < n; ++i)
        if (idata[i] != 0) odata[kept++] = idata[i];
    timer().endCpuTimer();
    return kept;
}

// flags -> exclusive scan -> scatter (cpu.cu:58-79)
inline int compactWithScan(int n, int* odata, const int* idata)
{
    if (n <= 0) return 0;
    int* keep = new int[n];
    int* index = new int[n];
    timer().startCpuTimer();
    for (int i = 0; i < n; ++i) keep[i] = idata[i] != 0;
    int run = 0;
    for (int i = 0; i < n; ++i) { index[i] = run; run += keep[i]; }
    for (int i = 0; i < n; ++i)
        if (keep[i]) odata[index[i]] = idata[i];
    timer().endCpuTimer();
    delete[] keep;
    delete[] index;
    return run;
}
}  // namespace CPU
}  // namespace StreamCompaction
