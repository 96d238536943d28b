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


def time_stream(fn_of_k, n_sets, launches, blocker=None):
    """per-launch time of `launches` back-to-back launches cycling through n_sets distinct (input, output) pairs whose
    total size exceeds the L2 (so every launch reads cold data), one CUDA-event pair around the whole train: what a
    launch costs INSIDE a stream of work (the single-launch numbers above it carry ~10 us of launch + event latency)"""
    stream = torch.cuda.current_stream()
    for k in range(n_sets):
        fn_of_k(k)
    torch.cuda.synchronize()
    best = None
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if blocker is not None:                # ~2 ms of GPU work ahead of the train, so that the host (a ctypes call per
            for _ in range(20):                # launch: ~6 us) has queued every launch before the first one starts
                blocker.add_(1)
        e0.record(stream)
        for i in range(launches):
            fn_of_k(i % n_sets)
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1) / launches
        best = ms if best is None else min(best, ms)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lg", type=int, nargs="+", default=[20, 22, 24, 26, 28])
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--stream-max-lg", type=int, default=24, help="in-stream per-launch timing up to this size")
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
            for name, variant in (("lookback", sc.LOOKBACK), ("tile", sc.LOOKBACK_TILE), ("ring", sc.LOOKBACK_RING),
                                  ("wave", sc.LOOKBACK_WAVE), ("efficient", sc.EFFICIENT), ("naive", sc.NAIVE)):
                if (variant == sc.NAIVE and lg > 26) or (variant == sc.LOOKBACK_WAVE and lg > 22):
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
            if lg <= args.stream_max_lg:
                # in-stream: rotate over enough buffer pairs to exceed the 126 MB L2 (8 B per element per pair)
                n_sets = max(2, (300 << 20) // (8 * n) + 1)
                ins = [torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda") for _ in range(n_sets)]
                outs = [torch.empty(n, dtype=torch.int32, device="cuda") for _ in range(n_sets)]
                launches = min(256, max(64, 2 * n_sets))
                for name, variant in (("lookback", sc.LOOKBACK), ("tile", sc.LOOKBACK_TILE), ("ring", sc.LOOKBACK_RING),
                                      ("wave", sc.LOOKBACK_WAVE), ("efficient", sc.EFFICIENT)):
                    if variant == sc.LOOKBACK_WAVE and lg > 22:
                        continue
                    ms = time_stream(lambda k: ctx.scan(n, outs[k], ins[k], variant, stream=s), n_sets, launches, flush)
                    row[name + "_stream_ms"] = ms
                    row[name + "_stream_frac"] = 8.0 * n / ms / 1e6 / peak
                ms = time_stream(lambda k: ctx.compact(n, outs[k], ins[k], cnt, stream=s), n_sets, launches, flush)
                row["compact_stream_ms"] = ms
                row["compact_stream_frac"] = (4.0 * n + 4.0 * int(cnt.item())) / ms / 1e6 / peak
                del ins, outs
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
