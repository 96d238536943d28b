/* sc_oracle.c -- CPU ORACLE for the scan / stream-compaction path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under cuda_pathtracer_b200/ links, imports or calls this
 * file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may (and only as the checker or the timed CPU baseline).  The product has no CPU fallback.
 *
 * Plain-C restatement of the reference's CPU algorithm, following
 *   stream_compaction/stream_compaction/cpu.cu:16-22   calc_prefix_sum  (serial exclusive scan)
 *   stream_compaction/stream_compaction/cpu.cu:29-33   CPU::scan
 *   stream_compaction/stream_compaction/cpu.cu:40-51   CPU::compactWithoutScan
 *   stream_compaction/stream_compaction/cpu.cu:58-79   CPU::compactWithScan
 * and of the harness's data generator
 *   stream_compaction/src/testing_helpers.hpp:52-58    genArray (glibc rand() % maxval), with the
 *   seed FIXED (the reference seeds with time(nullptr), which is not reproducible).
 *
 * Parity pin: tests/test_oracle_sc.py checks every function here element-for-element against
 * the reference's own cpu.cu compiled from /root/reference (oracle/_ref/libref_sc_cpu.so) and
 * against the committed golden vectors in tests/golden/sc_golden.npz (generated from that
 * reference build by tests/golden/make_sc_golden.py).
 *
 * int32 arithmetic wraps (two's complement): the running sum is kept in unsigned and cast, which
 * is what the reference's `int` additions do on every supported compiler.
 */
#include <stdint.h>
#include <stdlib.h>

/* cpu.cu:16-22 */
static void calc_prefix_sum(int n, int* odata, const int* idata)
{
    uint32_t run = 0u;
    if (n <= 0) return;
    odata[0] = 0;
    for (int i = 1; i < n; ++i) {
        run += (uint32_t)idata[i - 1];
        odata[i] = (int)run;
    }
}

/* cpu.cu:29-33 */
void oracle_scan(int n, int* odata, const int* idata)
{
    calc_prefix_sum(n, odata, idata);
}

/* cpu.cu:40-51 */
int oracle_compact_without_scan(int n, int* odata, const int* idata)
{
    int cnt = 0;
    for (int i = 0; i < n; ++i)
        if (idata[i]) odata[cnt++] = idata[i];
    return cnt;
}

/* cpu.cu:58-79: keep[] -> exclusive scan -> scatter; count = scan[n-1] + keep[n-1] */
int oracle_compact_with_scan(int n, int* odata, const int* idata)
{
    if (n <= 0) return 0;
    int* keep = (int*)malloc(sizeof(int) * (size_t)n);
    int* scan = (int*)malloc(sizeof(int) * (size_t)n);
    if (!keep || !scan) { free(keep); free(scan); return -1; }
    for (int i = 0; i < n; ++i) keep[i] = idata[i] ? 1 : 0;
    calc_prefix_sum(n, scan, keep);
    for (int i = 0; i < n; ++i)
        if (keep[i]) odata[scan[i]] = idata[i];
    int cnt = scan[n - 1] + keep[n - 1];
    free(keep);
    free(scan);
    return cnt;
}

/* testing_helpers.hpp:52-58 with a fixed seed; the harness then forces the last power-of-two
 * element to 0 (stream_compaction/src/main.cpp:22-23,97-98): pass n_pot-1 here and zero it there. */
void oracle_gen_array(int n, int* a, int maxval, unsigned seed)
{
    srand(seed);
    for (int i = 0; i < n; ++i) a[i] = rand() % maxval;
}
