// TEST / BASELINE INFRASTRUCTURE -- not product code.
//
// extern "C" doorway onto the reference's own, unmodified CPU scan/compact
// (stream_compaction/stream_compaction/cpu.cu:16-79), compiled from where it
// lies under /root/reference by oracle/Makefile into oracle/_ref/libref_sc_cpu.so.
// Used (a) to pin oracle/sc_oracle.c against the reference itself and (b) as the
// "reference" CPU baseline timed by bench.py.  OpenMP stays off: the pragmas in
// cpu.cu are inert in the reference build and racy if enabled.
#include <stream_compaction/cpu.h>

extern "C" {

void ref_cpu_scan(int n, int* odata, const int* idata)
{
    StreamCompaction::CPU::scan(n, odata, idata);
}

int ref_cpu_compact_without_scan(int n, int* odata, const int* idata)
{
    return StreamCompaction::CPU::compactWithoutScan(n, odata, idata);
}

int ref_cpu_compact_with_scan(int n, int* odata, const int* idata)
{
    return StreamCompaction::CPU::compactWithScan(n, odata, idata);
}

// the reference's own std::chrono timer reading for the previous call (ms)
float ref_cpu_last_ms(void)
{
    return StreamCompaction::CPU::timer().getCpuElapsedTimeForPreviousOperation();
}

}  // extern "C"
