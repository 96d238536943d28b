"""Scan / compact sweep (BASELINE config 3): device-resident inputs, CUDA-event timing on the
launching stream, algorithmic GB/s (scan 8 B/elem; compact 4n + 4*count) against the measured HBM
peak.  Optionally runs the reference's own GPU scans (oracle/_ref/ref_sc_gpu_bench) beside it.

    python benchmarks/sc_sweep.py --lg 20 22 24 26 28 [--ref] [--json out.json]
"""
import argparse
import json
import os
import subprocess
import sys

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from cuda_pathtracer_b200 import stream_compaction as sc  # noqa: E402


def hbm_peak():
    try:
        return json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def time_op(fn, iters, flush):
    stream = torch.cuda.current_stream()
    for _ in range(3):
        fn()
    times = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1)                      # write > L2 bytes between timed launches
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        fn()
        e1.record(stream)
        e1.synchronize()
        times.append(e0.elapsed_time(e1))
    times.sort()
    return times[len(times) // 2], times[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lg", type=int, nargs="+", default=[20, 22, 24, 26, 28])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    peak, peak_kind = hbm_peak()
    ctx = sc.Context()
    s = torch.cuda.current_stream().cuda_stream
    flush = torch.zeros(64 << 20, dtype=torch.int32, device="cuda")   # 256 MB > 126 MB L2
    rows = []
    for lg in args.lg:
        for delta in (0, 3):
            n = (1 << lg) - delta
            a = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda")
            a[-1] = 0
            b = a % 4
            out = torch.empty_like(a)
            cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
            row = {"n": n, "lg": lg, "npot": bool(delta)}
            for name, variant in (("lookback", sc.LOOKBACK), ("efficient", sc.EFFICIENT), ("naive", sc.NAIVE)):
                if variant == sc.NAIVE and lg > 26:
                    continue
                med, best = time_op(lambda: ctx.scan(n, out, a, variant, stream=s), args.iters, flush)
                row[name + "_ms"] = med
                row[name + "_gbs"] = 8.0 * n / med / 1e6
                row[name + "_frac"] = row[name + "_gbs"] / peak
            med, best = time_op(lambda: ctx.compact(n, out, b, cnt, stream=s), args.iters, flush)
            k = int(cnt.item())
            row["compact_ms"] = med
            row["compact_gbs"] = (4.0 * n + 4.0 * k) / med / 1e6
            row["compact_frac"] = row["compact_gbs"] / peak
            med, best = time_op(lambda: ctx.reduce(n, a, cnt, stream=s), args.iters, flush)
            row["reduce_ms"] = med
            row["reduce_gbs"] = 4.0 * n / med / 1e6
            if args.ref and delta in (0, 3):
                exe = os.path.join(REPO, "oracle", "_ref", "ref_sc_gpu_bench")
                if os.path.exists(exe):
                    r = subprocess.run([exe, str(lg), str(delta), "1" if lg > 26 else "0"],
                                       capture_output=True, text=True)
                    try:
                        row["reference"] = json.loads(r.stdout.strip().splitlines()[-1])
                    except Exception:
                        row["reference"] = {"error": (r.stdout + r.stderr)[-300:]}
            rows.append(row)
            print(json.dumps(row), flush=True)
            del a, b, out
    res = {"peak_gbs": peak, "peak_kind": peak_kind, "rows": rows}
    if args.json:
        json.dump(res, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
