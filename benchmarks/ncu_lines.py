"""Source-line view of an `ncu --page source --csv` dump (SASS rows) of one kernel:

    cuobjdump -xelf all cuda_pathtracer_b200/libb200pt.so && nvdisasm -g -c pt_kernels.sm_100a.cubin > k.sass
    python benchmarks/ncu_lines.py k.sass KERNEL_SUBSTRING gpurun_out/rN_..._src.csv [which = 1] [top = 30]

Joins the per-SASS-instruction counters (instructions executed, thread instructions, stall samples) with the
line table nvdisasm prints for a -lineinfo build and aggregates them per source line and per 20-line bucket:
share of warp instructions, share of stall samples, active threads per instruction."""
import collections
import csv
import re
import sys


def main():
    sass, kern, src = sys.argv[1], sys.argv[2], sys.argv[3]
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    topn = int(sys.argv[5]) if len(sys.argv) > 5 else 30
    amap, cur, on = {}, None, False
    for ln in open(sass):
        if ln.startswith("//---") and ".text." in ln:
            on = kern in ln
            continue
        if not on:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
        if m:
            amap[int(m.group(1), 16)] = cur
    rows, k, hdr, H = [], 0, None, None
    for r in csv.reader(open(src)):
        if r and r[0] == "Kernel Name":
            k += 1
            hdr = None
            continue
        if (k + 1) // 2 != which:          # the dump repeats the "Kernel Name" row: two per kernel
            continue
        if hdr is None:
            hdr = H = r
            continue
        rows.append(r)
    ix = {h: i for i, h in enumerate(H)}

    def num(r, h):
        try:
            return int(float(r[ix[h]]))
        except (ValueError, IndexError):
            return 0

    base = min(int(r[ix["Address"]], 16) for r in rows)
    agg = collections.defaultdict(lambda: [0, 0, 0])
    for r in rows:
        key = amap.get(int(r[ix["Address"]], 16) - base, ("?", 0))
        agg[key][0] += num(r, "Instructions Executed")
        agg[key][1] += num(r, "# Samples")
        agg[key][2] += num(r, "Thread Instructions Executed")
    ti = sum(v[0] for v in agg.values()) or 1
    ts = sum(v[1] for v in agg.values()) or 1
    print("warp instructions", ti, "samples", ts, "threads/inst %.2f" % (sum(v[2] for v in agg.values()) / ti))
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
        print(f"{100 * v[0] / ti:5.1f}%i {100 * v[1] / ts:5.1f}%s {v[2] / max(1, v[0]):5.1f} thr  {key[0]}:{key[1]}")
    print("---- 20-line buckets")
    b = collections.defaultdict(lambda: [0, 0, 0])
    for key, v in agg.items():
        kk = (key[0], key[1] // 20 * 20)
        for j in range(3):
            b[kk][j] += v[j]
    for kk, v in sorted(b.items(), key=lambda kv: -kv[1][0])[:16]:
        print(f"{100 * v[0] / ti:5.1f}%i {100 * v[1] / ts:5.1f}%s {v[2] / max(1, v[0]):5.1f} thr  {kk[0]}:{kk[1]}-{kk[1] + 19}")


if __name__ == "__main__":
    main()
