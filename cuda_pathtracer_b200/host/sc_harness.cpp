// sc_harness.cpp -- self-checking test harness of the scan / compact library, in the shape of the
// reference's stream_compaction/src/main.cpp:14-146 (same sequence of cases, same "passed" /
// "FAIL VALUE" / "FAIL COUNT" verdicts, CPU results as the ground truth, power-of-two and
// non-power-of-two = SIZE - 3), but with the array size and the seed on the command line:
//
//     stream_compaction_test [LG_SIZE = 20] [--seed S = 565] [--cpu-only] [--json]
//
// Data: the reference's generator (testing_helpers.hpp:52-58) with a FIXED seed: srand(S);
// a[i] = rand() % 50 (scan) / % 4 (compact), last element 0.  --cpu-only runs the
// StreamCompaction::CPU cases alone (BASELINE.json configs[0]; needs no GPU).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "b200_mg.h"
#include "b200_stream_compaction.hpp"

namespace {

int failures = 0;

int cmpArrays(int n, const int* a, const int* b)
{
    for (int i = 0; i < n; i++)
        if (a[i] != b[i]) {
            std::printf("    a[%d] = %d, b[%d] = %d\n", i, a[i], i, b[i]);
            return 1;
        }
    return 0;
}
void printDesc(const char* desc) { std::printf("==== %s ====\n", desc); }
void printCmpResult(int n, const int* a, const int* b)
{
    const bool bad = cmpArrays(n, a, b) != 0;
    failures += bad;
    std::printf("    %s \n", bad ? "FAIL VALUE" : "passed");
}
void printCmpLenResult(int n, int expN, const int* a, const int* b)
{
    if (n != expN) std::printf("    expected %d elements, got %d\n", expN, n);
    const char* verdict = (n == -1 || n != expN) ? "FAIL COUNT" : cmpArrays(n, a, b) ? "FAIL VALUE" : "passed";
    failures += std::strcmp(verdict, "passed") != 0;
    std::printf("    %s \n", verdict);
}
void genArray(int n, int* a, int maxval, unsigned seed)
{
    srand(seed);
    for (int i = 0; i < n; i++) a[i] = rand() % maxval;
}
void printArray(int n, const int* a)
{
    std::printf("    [ ");
    for (int i = 0; i < n; i++) {
        if (i + 2 == 15 && n > 16) { i = n - 2; std::printf("... "); }
        std::printf("%3d ", a[i]);
    }
    std::printf("]\n");
}
void printElapsed(float ms, const char* note) { std::printf("   elapsed time: %gms    %s\n", ms, note); }

}  // namespace

int main(int argc, char** argv)
{
    int lg = 20;
    unsigned seed = 565;
    bool cpu_only = false, json = false;
    int gpus = 1;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--seed" && i + 1 < argc) seed = (unsigned)std::atoi(argv[++i]);
        else if (a == "--cpu-only") cpu_only = true;
        else if (a == "--json") json = true;
        else if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
        else if (a[0] != '-') lg = std::atoi(a.c_str());
        else { std::printf("usage: %s [LG_SIZE] [--seed S] [--cpu-only] [--json] [--gpus K]\n", argv[0]); return 1; }
    }
    if (lg < 2 || lg > 30) { std::printf("LG_SIZE must be in [2, 30]\n"); return 1; }
    const int SIZE = 1 << lg, NPOT = SIZE - 3;
    std::vector<int> va(SIZE), vb(SIZE), vc(SIZE);
    int *a = va.data(), *b = vb.data(), *c = vc.data();
    namespace SC = StreamCompaction;
    struct Row { const char* name; float ms; };
    std::vector<Row> rows;
    auto zero = [&](int* p) { std::memset(p, 0, sizeof(int) * (size_t)SIZE); };

    std::printf("\n****************\n** SCAN TESTS **\n****************\n");
    genArray(SIZE - 1, a, 50, seed);
    a[SIZE - 1] = 0;
    printArray(SIZE, a);

    zero(b);
    printDesc("cpu scan, power-of-two");
    SC::CPU::scan(SIZE, b, a);
    rows.push_back({"cpu_scan_pot", SC::CPU::timer().getCpuElapsedTimeForPreviousOperation()});
    printElapsed(rows.back().ms, "(std::chrono Measured)");
    printArray(SIZE, b);

    zero(c);
    printDesc("cpu scan, non-power-of-two");
    SC::CPU::scan(NPOT, c, a);
    rows.push_back({"cpu_scan_npot", SC::CPU::timer().getCpuElapsedTimeForPreviousOperation()});
    printElapsed(rows.back().ms, "(std::chrono Measured)");
    printCmpResult(NPOT, b, c);

    if (!cpu_only) {
        struct Case { const char* desc; const char* key; int n; void (*fn)(int, int*, const int*); SC::Common::PerformanceTimer& (*tm)(); };
        const Case cases[] = {
            {"naive scan, power-of-two", "naive_scan_pot", SIZE, SC::Naive::scan, SC::Naive::timer},
            {"naive scan, non-power-of-two", "naive_scan_npot", NPOT, SC::Naive::scan, SC::Naive::timer},
            {"work-efficient scan, power-of-two", "efficient_scan_pot", SIZE, SC::Efficient::scan, SC::Efficient::timer},
            {"work-efficient scan, non-power-of-two", "efficient_scan_npot", NPOT, SC::Efficient::scan, SC::Efficient::timer},
            {"thrust scan, power-of-two", "thrust_scan_pot", SIZE, SC::Thrust::scan, SC::Thrust::timer},
            {"thrust scan, non-power-of-two", "thrust_scan_npot", NPOT, SC::Thrust::scan, SC::Thrust::timer},
        };
        for (const Case& k : cases) {
            zero(c);
            printDesc(k.desc);
            k.fn(k.n, c, a);
            rows.push_back({k.key, k.tm().getGpuElapsedTimeForPreviousOperation()});
            printElapsed(rows.back().ms, "(CUDA Measured)");
            printCmpResult(k.n, b, c);
        }
    }

    // --gpus K: the same scan over K devices of this box (include/b200_mg.h: contiguous shards, ONE all-gather of one
    // int per device, device-side seeds), host pointers in and out like every other variant of this harness
    b200mg* group = nullptr;
    float exchange_ms = 0.f;
    if (!cpu_only && gpus > 1) {
        const int rc = b200mg_create(gpus, nullptr, &group);
        if (rc) { std::printf("b200mg_create(%d): %s\n", gpus, b200mg_error_string(rc)); return 2; }
        const struct { const char* desc; const char* key; int n; } mcases[] = {
            {"sharded scan, power-of-two", "sharded_scan_pot", SIZE}, {"sharded scan, non-power-of-two", "sharded_scan_npot", NPOT}};
        for (const auto& k : mcases) {
            zero(c);
            printDesc(k.desc);
            float ms = 0.f;
            const int rc2 = b200mg_scan_i32(group, k.n, c, a, &ms, &exchange_ms);
            if (rc2) { std::printf("b200mg_scan_i32: %s\n", b200mg_error_string(rc2)); return 2; }
            rows.push_back({k.key, ms});
            printElapsed(ms, "(CUDA Measured, max over devices, all-gather included)");
            printCmpResult(k.n, b, c);
        }
        rows.push_back({"sharded_scan_all_gather", exchange_ms});
    }

    std::printf("\n*****************************\n** STREAM COMPACTION TESTS **\n*****************************\n");
    genArray(SIZE - 1, a, 4, seed);
    a[SIZE - 1] = 0;
    printArray(SIZE, a);
    int count, expectedCount, expectedNPOT;

    zero(b);
    printDesc("cpu compact without scan, power-of-two");
    count = SC::CPU::compactWithoutScan(SIZE, b, a);
    rows.push_back({"cpu_compact_pot", SC::CPU::timer().getCpuElapsedTimeForPreviousOperation()});
    printElapsed(rows.back().ms, "(std::chrono Measured)");
    expectedCount = count;
    printArray(count, b);
    printCmpLenResult(count, expectedCount, b, b);

    zero(c);
    printDesc("cpu compact without scan, non-power-of-two");
    count = SC::CPU::compactWithoutScan(NPOT, c, a);
    rows.push_back({"cpu_compact_npot", SC::CPU::timer().getCpuElapsedTimeForPreviousOperation()});
    printElapsed(rows.back().ms, "(std::chrono Measured)");
    expectedNPOT = count;
    printArray(count, c);
    printCmpLenResult(count, expectedNPOT, b, c);

    zero(c);
    printDesc("cpu compact with scan");
    count = SC::CPU::compactWithScan(SIZE, c, a);
    rows.push_back({"cpu_compact_with_scan", SC::CPU::timer().getCpuElapsedTimeForPreviousOperation()});
    printElapsed(rows.back().ms, "(std::chrono Measured)");
    printArray(count, c);
    printCmpLenResult(count, expectedCount, b, c);

    if (!cpu_only) {
        zero(c);
        printDesc("work-efficient compact, power-of-two");
        count = SC::Efficient::compact(SIZE, c, a);
        rows.push_back({"efficient_compact_pot", SC::Efficient::timer().getGpuElapsedTimeForPreviousOperation()});
        printElapsed(rows.back().ms, "(CUDA Measured)");
        printCmpLenResult(count, expectedCount, b, c);

        zero(c);
        printDesc("work-efficient compact, non-power-of-two");
        count = SC::Efficient::compact(NPOT, c, a);
        rows.push_back({"efficient_compact_npot", SC::Efficient::timer().getGpuElapsedTimeForPreviousOperation()});
        printElapsed(rows.back().ms, "(CUDA Measured)");
        printCmpLenResult(count, expectedNPOT, b, c);
    }

    if (group) {
        const struct { const char* desc; const char* key; int n; int expected; } mcases[] = {
            {"sharded compact, power-of-two", "sharded_compact_pot", SIZE, expectedCount},
            {"sharded compact, non-power-of-two", "sharded_compact_npot", NPOT, expectedNPOT}};
        for (const auto& k : mcases) {
            zero(c);
            printDesc(k.desc);
            float ms = 0.f;
            const int rc2 = b200mg_compact_i32(group, k.n, c, a, &count, &ms, &exchange_ms);
            if (rc2) { std::printf("b200mg_compact_i32: %s\n", b200mg_error_string(rc2)); return 2; }
            rows.push_back({k.key, ms});
            printElapsed(ms, "(CUDA Measured, max over devices, all-gather included)");
            printCmpLenResult(count, k.expected, b, c);
        }
        b200mg_destroy(group);
    }

    if (json) {
        std::printf("{\"n\": %d, \"seed\": %u, \"gpus\": %d, \"failures\": %d, \"ms\": {", SIZE, seed, gpus, failures);
        for (size_t i = 0; i < rows.size(); ++i) std::printf("%s\"%s\": %.4f", i ? ", " : "", rows[i].name, rows[i].ms);
        std::printf("}}\n");
    }
    return failures ? 2 : 0;
}
