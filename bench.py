"""bench.py -- headline benchmark of the B200-native path tracer + scan/compact library.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Metric (BASELINE.json): Mpath-segments/s (and spp/s) on the Cornell box 800x800, depth 8, reference
defaults (Russian roulette, SSAA and DoF on); scan / compact GB/s at 2^28 ints ride along in the
same JSON line.  A STEP is one render call of `--iters-per-step` sample iterations over all
640 000 pixels (default 250: the driver's 20 steps are BASELINE config 2's 5000 iterations) and, at
N > 1, the NCCL all-reduce of the float accumulation image that closes a sharded render.

  value        whole-job path-segments/s with the scene resident in HBM (device-timed, CUDA events,
               max over ranks); a path segment = one live path entering intersection in one bounce
               (SURVEY.md 8d).  N > 1: rank g traces iterations g+1, g+1+N, ... (weak scaling: every
               rank traces --iters-per-step iterations per step); EVERY step ends with the all-reduce
               of the 3*W*H float image (into a second buffer, so the private sums stay valid).
  e2e          the same metric through the C ABI the reference's pathtrace.h maps onto, with HOST
               buffers: every step re-uploads the scene (H2D), renders, (N > 1: all-reduces) and reads
               the -- reduced -- float SUM image back into pinned host memory (D2H), all timed.
  roofline     dominant kernel k_bounce against the measured HBM peak: algorithmic bytes (96 B per
               path segment + 24 B per pixel per iteration) / its summed CUDA-event time.
  cpu_baseline the CPU oracle (oracle/pt_oracle.cpp, a port of the reference's algorithm) on all
               host threads, on a bounded sample of the same workload (rank 0, N = 1 only).
  N = 1 extras reference_cuda_build + dropin (the reference's own CUDA build, and the reference's own
               host loop on this library through integration/pathtrace_b200.cpp, same box, same run),
               scan (2^28 ints: kernels vs HBM peak, e2e with host pointers, the reference's CPU code on one
               core, the reference's GPU Efficient / Thrust scans).
  N > 1 extras config5 (4096 primitives 3840x2160, iterations split, 99.5 MB all-reduce timed by itself),
               scan_sharded / compact_sharded (2^30-3 ints, one all-gather of one int per rank), strong
               (5000 Cornell iterations split over N) and multi_gpu_check (split render == single-rank render).

--impl reference times the reference path on the host cores (the oracle port; the reference has
no CPU renderer of its own) with the same metric / config and prints the same line.
The oracle is only ever the checker / the timed baseline here; the product path is the CUDA library.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

SCENE = "cornell"
ALG_BYTES_PER_SEGMENT = 96.0        # read + write one 48-B path state
ALG_BYTES_PER_PIXEL_ITER = 24.0     # read-modify-write of the 12-B accumulator
REF_ITERS_PER_STEP = 8              # --impl reference: iterations per render call of the CPU arm


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(REPO, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.rows, self.proc, self.idx = [], None, device_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def load_scene(name=None):
    """The scene as this repo's OWN C++ loader reads scenes/<name>.json (host/scene.cpp, `--dump-scene`; byte-identical
    to the reference loader's POD dump, tests/test_host_cli.py); the committed dump when the CLI has not been built."""
    from cuda_pathtracer_b200 import pathtracer as pt
    name = name or SCENE
    exe = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "cuda_path_tracer")
    src = os.path.join(REPO, "scenes", name + ".json")
    if os.path.exists(exe) and os.path.exists(src):
        extra = sum((["--refract", str(i), "1.5"] for i in range(24, 32)), []) if name.startswith("synthetic") else \
            (["--refract", "4", "1.5"] if "glass" in name else [])
        with tempfile.TemporaryDirectory() as tmp:
            out = os.path.join(tmp, "scene.bin")
            r = subprocess.run([exe, src, "--dump-scene", out] + extra, capture_output=True, text=True)
            if r.returncode == 0 and os.path.exists(out):
                return pt.Scene.from_pod_file(out)
    return pt.Scene.from_pod_file(os.path.join(REPO, "tests", "golden", name + ".scene.bin"))


def default_gui(pt):
    return pt.GuiDataContainer()          # reference defaults: RR, SSAA, DoF on; aperture .1; focal 10


def workload_config(sc):
    """the `config` object: identical for both arms (what differs between them is under `run`)"""
    return {"workload": f"{SCENE} {sc.width}x{sc.height}, trace depth {sc.traceDepth}, reference defaults (RR+SSAA+DoF)",
            "scene": SCENE, "resolution": [sc.width, sc.height], "trace_depth": sc.traceDepth,
            "l2": "GPU arm: no flush between steps, the two path pools of a batch (2 x 48 B x 640000 x 32 = 1.97 GB) "
                  "exceed the 126 MB L2; CPU arm: not applicable",
            "parallelism": "replicated scene, iterations split across ranks (stride N), one NCCL all-reduce of the float "
                           "accumulation image per step at N > 1; one GPU / the host threads at N = 1"}


def cpu_baseline(n_iters, threads=None):
    """oracle/pt_oracle.cpp on the host cores: bounded sample of the same workload"""
    from oracle_bindings import PtOracle
    from cuda_pathtracer_b200 import pathtracer as pt
    sc, o = load_scene(), PtOracle()
    threads = threads or (os.cpu_count() or 1)
    t0 = time.perf_counter()
    _, seg = o.render(sc, default_gui(pt), 1, n_iters, threads=threads)
    dt = time.perf_counter() - t0
    return {"value": seg / dt / 1e6, "unit": "Mpath-segments/s", "cores": threads, "kind": "port",
            "spp_per_s": n_iters / dt,
            "sample": f"{n_iters} iterations of {SCENE} 800x800 depth 8 ({seg} segments, {dt:.2f} s) "
                      f"with oracle/pt_oracle.cpp (CPU port of the reference's algorithm) on {threads} threads"}


def run_json_tool(cmd, timeout):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    for line in reversed(r.stdout.strip().splitlines()):
        if line.startswith("{"):
            return json.loads(line)
    raise RuntimeError(f"{os.path.basename(cmd[0])}: rc {r.returncode}, no JSON line ({r.stderr[-200:]})")


def reference_cuda_build(n_iters):
    """The reference's OWN CUDA build (oracle/_ref/ref_pathtracer: unmodified pathtrace.cu etc., built in the build
    container) on this GPU -- the >= 5x comparison north_star names -- and, beside it, `dropin`: the reference's own
    host loop (one pathtrace() call per iteration, main.cpp:142-160) on THIS library through the binding of
    integration/pathtrace_b200.cpp (oracle/_ref/dropin_pathtracer: batch = 1 and a D2H copy of the SUM image per
    call, as the interface demands)."""
    scene = os.path.join(REPO, "scenes", SCENE + ".json")
    out = {}
    for key, name in (("reference_cuda_build", "ref_pathtracer"), ("dropin", "dropin_pathtracer")):
        exe = os.path.join(REPO, "oracle", "_ref", name)
        if not (os.path.exists(exe) and os.path.exists(scene)):
            continue
        try:
            run_json_tool([exe, scene, "--iters", "3"], 120)                                   # warm-up
            d = run_json_tool([exe, scene, "--iters", str(n_iters)], 600)
            out[key] = {"spp_per_s": d["spp_per_s"], "ms_per_iter": d["ms_per_iter"], "iters": n_iters,
                        "how": f"oracle/_ref/{name}, wall clock around its pathtrace() loop ("
                               + ("reference sources unmodified, headless shim)" if key == "reference_cuda_build" else
                                  "the reference's host loop + integration/pathtrace_b200.cpp + libb200pt.so)")}
        except Exception as e:                                     # noqa: BLE001
            out[key] = {"error": str(e)[:200]}
    if "dropin" in out and "reference_cuda_build" in out and "spp_per_s" in out["dropin"] and \
            "spp_per_s" in out["reference_cuda_build"]:
        out["dropin"]["vs_reference_cuda_build"] = out["dropin"]["spp_per_s"] / out["reference_cuda_build"]["spp_per_s"]
    return out


def scan_bench(torch, peak):
    """scan / compact at 2^28 ints (BASELINE metric, second half).
    kernels: device-resident, CUDA events on the launching stream, inputs (1 GiB) far larger than L2;
    e2e: b200_scan_i32 / b200_compact_i32 with HOST pointers (pinned), H2D + kernels + D2H inside the wall-clock timer,
         the call stream_compaction/src/main.cpp:32-138 makes;
    cpu_baseline: the reference's own cpu.cu:16-79 (oracle/_ref/libref_sc_cpu.so) on ONE core (the code is serial);
    reference: the reference's own GPU Efficient / Thrust scans (oracle/_ref/ref_sc_gpu_bench: best of 5 timer()
         readings after a warm-up call)."""
    import numpy as np
    from cuda_pathtracer_b200 import stream_compaction as sc
    n = 1 << 28
    ctx = sc.Context()
    stream = torch.cuda.current_stream()
    g = torch.Generator(device="cuda").manual_seed(565)
    a = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g)
    a[-1] = 0
    b = a % 4
    out = torch.empty_like(a)
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    res = {"n": n}

    def timed(fn, reps=10):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); fn(); e1.record(stream); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    kernels = {"scan_lookback": "k_scan_tma (single pass, persistent TMA ring; what B200_SCAN_LOOKBACK runs from 2^24 aligned ints on)",
               "scan_lookback_tile": "k_scan_lookback (single pass, one register tile per CTA; B200_SCAN_LOOKBACK_TILE)",
               "scan_efficient": "k_reduce_tiles + k_downsweep (work-efficient, 12 B/element of traffic, quoted at 8)"}
    for name, variant in (("scan_lookback", sc.LOOKBACK), ("scan_lookback_tile", sc.LOOKBACK_TILE), ("scan_efficient", sc.EFFICIENT)):
        ms = timed(lambda: ctx.scan(n, out, a, variant, stream=stream.cuda_stream))
        res[name] = {"ms": ms, "gbs": 8.0 * n / ms / 1e6, "frac": 8.0 * n / ms / 1e6 / peak, "kernel": kernels[name]}
    ms = timed(lambda: ctx.compact(n, out, b, cnt, stream=stream.cuda_stream))
    k = int(cnt.item())
    res["compact"] = {"ms": ms, "gbs": (4.0 * n + 4.0 * k) / ms / 1e6, "frac": (4.0 * n + 4.0 * k) / ms / 1e6 / peak,
                      "kept": k, "kernel": "k_compact_tma (single pass on the TMA ring, in-place packing, bulk-stored runs)"}
    ctx.set_ring_min_n(2**31 - 1)                                # the register-tile compaction, for the record
    ms = timed(lambda: ctx.compact(n, out, b, cnt, stream=stream.cuda_stream))
    res["compact_tile"] = {"ms": ms, "gbs": (4.0 * n + 4.0 * k) / ms / 1e6, "frac": (4.0 * n + 4.0 * k) / ms / 1e6 / peak,
                           "kernel": "k_compact_ws (persistent, register tiles, look-back warp; ring switched off)"}
    ctx.set_ring_min_n(0)
    want_scan_tail = int((torch.cumsum(a[-4096:].to(torch.int64), 0)[-1]).item())     # (only used as a smoke value)
    ctx.close()

    # ---- e2e: host pointers through the C ABI ---------------------------------------------------------
    try:
        h_in = torch.empty(n, dtype=torch.int32, pin_memory=True)
        h_out = torch.empty(n, dtype=torch.int32, pin_memory=True)
        h_in.copy_(a)
        torch.cuda.synchronize()
        sc.Thrust.scan(n, h_out, h_in)                            # warm-up (allocates the staging buffers)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            sc.Thrust.scan(n, h_out, h_in)
            ts.append(time.perf_counter() - t0)
        dt = min(ts)
        ok = bool(int(h_out[1]) == int(h_in[0]) and int(h_out[0]) == 0)
        res["e2e_scan"] = {"value": 8.0 * n / dt / 1e9, "unit": "GB/s", "ms": dt * 1e3, "h2d_bytes_per_step": 4 * n,
                           "d2h_bytes_per_step": 4 * n, "kernel_ms": sc.Thrust.timer().getGpuElapsedTimeForPreviousOperation(),
                           "how": "b200_scan_i32(n, host out, host in) with pinned host arrays, wall clock, best of 3", "sane": ok}
        h_in.copy_(b)
        torch.cuda.synchronize()
        sc.Efficient.compact(n, h_out, h_in)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            kk = sc.Efficient.compact(n, h_out, h_in)
            ts.append(time.perf_counter() - t0)
        dt = min(ts)
        res["e2e_compact"] = {"value": (4.0 * n + 4.0 * kk) / dt / 1e9, "unit": "GB/s", "ms": dt * 1e3,
                              "h2d_bytes_per_step": 4 * n, "d2h_bytes_per_step": 4 * kk, "kept": kk,
                              "kernel_ms": sc.Efficient.timer().getGpuElapsedTimeForPreviousOperation(),
                              "how": "b200_compact_i32(n, host out, host in) with pinned host arrays, wall clock, best of 3",
                              "sane": bool(kk == k)}
        del h_out
    except Exception as e:                                         # noqa: BLE001
        res["e2e_scan"] = {"error": str(e)[:200]}
        h_in = None
    del a, b, out
    torch.cuda.empty_cache()

    # ---- the reference's CPU scan / compact on one host core ---------------------------------------------
    try:
        from oracle_bindings import ScOracle, gen_array
        try:
            o, kind = ScOracle(reference=True), "reference"
        except Exception:
            o, kind = ScOracle(), "port"
        arr = gen_array(n, 50)
        t0 = time.perf_counter(); o.scan(arr); t_scan = time.perf_counter() - t0
        own = o.last_ms()
        arr = gen_array(n, 4)
        t0 = time.perf_counter(); kept = o.compact_without_scan(arr); t_cmp = time.perf_counter() - t0
        res["cpu_baseline"] = {"value": 8.0 * n / t_scan / 1e9, "unit": "GB/s", "cores": 1, "host_cores": os.cpu_count(),
                               "kind": kind, "scan_ms": t_scan * 1e3, "scan_ms_own_timer": own,
                               "compact_ms": t_cmp * 1e3, "compact_gbs": (4.0 * n + 4.0 * kept.size) / t_cmp / 1e9,
                               "sample": f"one CPU::scan and one CPU::compactWithoutScan of all 2^28 ints "
                                         f"(stream_compaction/stream_compaction/cpu.cu:16-51, "
                                         f"{'compiled unmodified into oracle/_ref/libref_sc_cpu.so' if kind == 'reference' else 'oracle/sc_oracle.c port'}), "
                                         "serial code: 1 core; wall clock incl. the output array's first touch"}
        del arr, kept
    except Exception as e:                                         # noqa: BLE001
        res["cpu_baseline"] = {"error": str(e)[:200]}

    # ---- the reference's own GPU scans on this box ---------------------------------------------------------
    exe = os.path.join(REPO, "oracle", "_ref", "ref_sc_gpu_bench")
    if os.path.exists(exe):
        try:
            d = run_json_tool([exe, "28", "0", "1", "5"], 900)
            res["reference"] = {"efficient_ms": d["efficient_ms"], "thrust_ms": d["thrust_ms"],
                                "efficient_median_ms": d.get("efficient_median_ms"), "thrust_median_ms": d.get("thrust_median_ms"),
                                "efficient_gbs": 8.0 * n / d["efficient_ms"] / 1e6, "thrust_gbs": 8.0 * n / d["thrust_ms"] / 1e6,
                                "cpu_scan_ms": d["cpu_scan_ms"], "cpu_compact_ms": d["cpu_compact_ms"], "ok": d["ok"],
                                "how": "oracle/_ref/ref_sc_gpu_bench 28 (reference naive/efficient/thrust.cu unmodified): best of "
                                       f"{d.get('timed_calls', 1)} timer() readings after one warm-up call, kernels only as the reference times them"}
            res["reference"]["scan_speedup_vs_thrust"] = d["thrust_ms"] / res["scan_lookback"]["ms"]
            res["reference"]["scan_speedup_vs_efficient"] = d["efficient_ms"] / res["scan_lookback"]["ms"]
        except Exception as e:                                     # noqa: BLE001
            res["reference"] = {"error": str(e)[:200]}
    res["smoke"] = want_scan_tail
    return res


class _DevArray:
    """__cuda_array_interface__ view of a raw device pointer, so torch can wrap the library's
    accumulation image without a copy (for the NCCL all-reduce)."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def run_reference(args, rank):
    """--impl reference: the reference path on the host cores (CPU port; rank 0 only)"""
    if rank != 0:
        return
    from oracle_bindings import PtOracle
    from cuda_pathtracer_b200 import pathtracer as pt
    sc, o = load_scene(), PtOracle()
    gui = default_gui(pt)
    threads = os.cpu_count() or 1
    per = args.ref_iters_per_step
    it = 1
    for _ in range(args.warmup):
        o.render(sc, gui, it, per, threads=threads); it += per
    seg, t0 = 0, time.perf_counter()
    for _ in range(args.steps):
        _, s = o.render(sc, gui, it, per, threads=threads); it += per
        seg += s
    dt = time.perf_counter() - t0
    v = seg / dt / 1e6
    line = {"impl": "reference", "metric": "path_segments_per_s", "value": v, "unit": "Mpath-segments/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "spp_per_s": args.steps * per / dt,
            "config": workload_config(sc),
            "run": {"iters_per_step": per, "host_threads": threads,
                    "note": "each step is a bounded sample of the workload: one render call of "
                            f"{per} full 800x800 iterations on all host threads (thread start-up amortised over the call)"},
            "cpu_baseline": {"value": v, "unit": "Mpath-segments/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} steps x {per} iterations of the full 800x800 workload, "
                                       "oracle/pt_oracle.cpp (the reference has no CPU renderer; this is the CPU "
                                       "port of its algorithm, bit-identical to its CUDA build)"},
            "e2e": {"value": v, "unit": "Mpath-segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---- N > 1 legs ----------------------------------------------------------------------------------------
def _timed_max(torch, dist, world, stream, fn, reps=10, warm=3):
    """median CUDA-event time of fn over `reps` calls, max over ranks"""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def sharded_scan_compact_legs(torch, dist, rank, world, lg=30, npot=3):
    """BASELINE config 3 at N GPUs: contiguous shards, ONE all-gather of one int per rank, device-side seed; bit-exact
    against the whole array scanned / compacted on ONE GPU (this rank's own) by the same library."""
    from cuda_pathtracer_b200 import distributed as D
    from cuda_pathtracer_b200 import stream_compaction as sc
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    n = (1 << lg) - npot
    g = torch.Generator(device="cuda").manual_seed(565)
    full = torch.randint(0, 50, (n,), dtype=torch.int32, device="cuda", generator=g)      # same on every rank
    b, e = D.scan_shard(n, rank, world)
    shard = full[b:e].clone()
    ctx = sc.Context()
    res = torch.empty_like(shard)
    total = torch.zeros(1, dtype=torch.int32, device="cuda")
    totals = torch.zeros(world, dtype=torch.int32, device="cuda")
    seed = torch.zeros(1, dtype=torch.int32, device="cuda")

    def sharded_scan():
        ctx.reduce(e - b, shard, total, stream=sptr)
        dist.all_gather_into_tensor(totals, total)
        # exclusive prefix of the totals before this rank (int32 wrap-around), on the device: no host round trip
        seed.copy_(totals[:rank].sum(dtype=torch.int32).reshape(1) if rank else torch.zeros_like(seed))
        ctx.scan(e - b, res, shard, sc.LOOKBACK, d_seed=seed, stream=sptr)

    ms = _timed_max(torch, dist, world, stream, sharded_scan)
    ms_gather = _timed_max(torch, dist, world, stream, lambda: dist.all_gather_into_tensor(totals, total))
    whole = torch.empty_like(full)
    ctx.scan(n, whole, full, sc.LOOKBACK, stream=sptr)                  # the single-GPU result
    ok_scan = bool(torch.equal(res, whole[b:e]))
    ms_single = _timed_max(torch, dist, 1, stream, lambda: ctx.scan(n, whole, full, sc.LOOKBACK, stream=sptr), reps=5)
    out = {"scan_sharded": {"n": n, "ms": ms, "gbs": 8.0 * n / ms / 1e6, "all_gather_ms": ms_gather, "bit_exact": ok_scan,
                            "single_gpu_ms": ms_single, "speedup_vs_single_gpu": ms_single / ms}}

    flags = full % 4
    fshard = flags[b:e].clone()
    del full
    kept = torch.empty_like(fshard)
    cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
    counts = torch.zeros(world, dtype=torch.int32, device="cuda")

    def sharded_compact():
        ctx.compact(e - b, kept, fshard, cnt, stream=sptr)
        dist.all_gather_into_tensor(counts, cnt)

    ms = _timed_max(torch, dist, world, stream, sharded_compact)
    ctx.compact(n, whole, flags, cnt, stream=sptr)                      # the single-GPU result
    kk = int(cnt[0])
    cl = counts.cpu().tolist()
    off, k = sum(cl[:rank]), cl[rank]
    ok_c = bool(sum(cl) == kk and torch.equal(kept[:k], whole[off:off + k]))
    ms_single = _timed_max(torch, dist, 1, stream, lambda: ctx.compact(n, whole, flags, cnt, stream=sptr), reps=5)
    out["compact_sharded"] = {"n": n, "ms": ms, "gbs": (4.0 * n + 4.0 * kk) / ms / 1e6, "kept": kk, "bit_exact": ok_c,
                              "single_gpu_ms": ms_single, "speedup_vs_single_gpu": ms_single / ms,
                              "note": "survivors stay in their source shard at a known global offset (one all-gather of the "
                                      "counts); moving them into even shards is cuda_pathtracer_b200.distributed.rebalance_compacted"}
    oks = torch.tensor([int(ok_scan and ok_c)], device="cuda")
    dist.all_reduce(oks, op=dist.ReduceOp.MIN)
    out["scan_sharded"]["bit_exact_all_ranks"] = out["compact_sharded"]["bit_exact_all_ranks"] = bool(int(oks[0]))
    ctx.close()
    del whole, flags, fshard, kept, shard, res
    torch.cuda.empty_cache()
    return out


def render_legs(torch, dist, rank, world, local_rank, args):
    """config5, strong and multi_gpu_check (N > 1)"""
    from cuda_pathtracer_b200 import distributed as D
    from cuda_pathtracer_b200 import pathtracer as pt
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    out = {}

    def wrap(r):
        ptr, nfl = r.sum_device_ptr()
        return torch.as_tensor(_DevArray(ptr, nfl), device=f"cuda:{local_rank}")

    def sync_all():
        torch.cuda.synchronize()
        dist.barrier()

    def max_ms(ms):
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- multi_gpu_check: 32 iterations of cornell_small split over N == the same 32 on this rank alone ----
    sc = load_scene("cornell_small")
    r = pt.Renderer(sc)
    r.set_options(default_gui(pt))
    img = wrap(r)
    first, count, stride = D.iteration_shard(1, 32, rank, world)
    r.render(first, count, stride, stream=sptr, sync=False)
    got = img.clone()
    dist.all_reduce(got)
    torch.cuda.synchronize()
    r.reset()
    r.render(1, 32, 1)
    solo = torch.from_numpy(r.read_sum().reshape(-1)).cuda()
    ok = torch.tensor([int(torch.allclose(got, solo, rtol=2e-6, atol=1e-6))], device="cuda")
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    out["multi_gpu_check"] = "ok" if int(ok[0]) else "MISMATCH"
    out["multi_gpu_check_detail"] = {"scene": "cornell_small", "iters": 32, "split": f"stride {world}",
                                     "max_abs_diff_vs_single_rank": float((got - solo).abs().max()),
                                     "tolerance": "allclose rtol 2e-6 atol 1e-6 on the float SUM image, all ranks"}
    r.close()
    del got, solo

    # ---- strong scaling: BASELINE config 2's 5000 iterations split over N, one all-reduce ----
    sc = load_scene("cornell")
    r = pt.Renderer(sc)
    r.set_options(default_gui(pt))
    img = wrap(r)
    red = torch.empty_like(img)
    first, count, stride = D.iteration_shard(1, args.strong_iters, rank, world)
    r.render(first, min(count, 64), stride, stream=sptr, sync=False)     # warm-up
    red.copy_(img); dist.all_reduce(red)
    sync_all()
    r.reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    t0 = time.perf_counter()
    e0.record(stream)
    r.render(first, count, stride, stream=sptr, sync=False)
    red.copy_(img)
    dist.all_reduce(red)
    e1.record(stream)
    sync_all()
    wall = time.perf_counter() - t0
    ms = max_ms(e0.elapsed_time(e1))
    seg = torch.tensor([float(r.stats()["path_segments"])], dtype=torch.float64, device="cuda")
    dist.all_reduce(seg)
    out["strong"] = {"iters_total": args.strong_iters, "iters_this_rank": count, "ms": ms, "wall_ms": wall * 1e3,
                     "spp_per_s": args.strong_iters / (ms * 1e-3), "mpath_segments_per_s": float(seg[0]) / (ms * 1e-3) / 1e6,
                     "workload": "cornell 800x800 depth 8, BASELINE config 2's iterations split over the ranks (stride N), "
                                 "one all-reduce of the 7.68 MB SUM image; device time, max over ranks"}
    r.close()
    del red

    # ---- config 5: 4096 primitives 3840x2160, iterations split over N, one 99.5 MB all-reduce per step ----
    sc = load_scene("synthetic_4096")
    r = pt.Renderer(sc)
    r.set_options(default_gui(pt))
    img = wrap(r)
    red = torch.empty_like(img)
    per, steps = args.config5_iters, args.config5_steps
    it = [rank + 1]

    def step5():
        r.render(it[0], per, world, stream=sptr, sync=False)
        it[0] += per * world
        red.copy_(img)
        dist.all_reduce(red)

    step5()
    sync_all()
    r.reset()
    sync_all()
    e0.record(stream)
    for _ in range(steps):
        step5()
    e1.record(stream)
    sync_all()
    ms = max_ms(e0.elapsed_time(e1))
    seg = torch.tensor([float(r.stats()["path_segments"])], dtype=torch.float64, device="cuda")
    dist.all_reduce(seg)
    ar = _timed_max(torch, dist, world, stream, lambda: dist.all_reduce(red), reps=5, warm=2)
    out["config5"] = {"workload": f"synthetic_4096 {sc.width}x{sc.height} depth {sc.traceDepth} (BASELINE configs[4])",
                      "iters_per_step_per_rank": per, "steps": steps, "ms_per_step": ms / steps,
                      "spp_per_s": steps * per * world / (ms * 1e-3),
                      "mpath_segments_per_s": float(seg[0]) / (ms * 1e-3) / 1e6,
                      "allreduce_ms": ar, "allreduce_bytes": int(red.numel()) * 4,
                      "allreduce_busbw_gbs": 2.0 * (world - 1) / world * red.numel() * 4 / (ar * 1e-3) / 1e9,
                      "allreduce_share_of_step": ar / (ms / steps)}
    r.close()
    del red
    torch.cuda.empty_cache()
    return out


def cxx_host_leg(torch, dist, rank, world, args):
    """The same two sharded paths through the C++ host (north_star: "host code is C++"): ONE process, one host thread per
    device, NCCL (include/b200_mg.h) -- `cuda_path_tracer --gpus N` and `stream_compaction_test --gpus N`, run by rank 0
    while the other ranks wait on the rendezvous store (a CPU-side wait: their GPUs stay idle)."""
    import datetime
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    store = dist.distributed_c10d._get_default_store()
    out = {}
    if rank == 0:
        exe = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "cuda_path_tracer")
        try:
            d = run_json_tool([exe, os.path.join(REPO, "scenes", "cornell.json"), "--gpus", str(world), "--iters",
                               str(args.strong_iters), "--no-png", "--json"], 600)
            out["render"] = {"iters": d["iters"], "seconds": d["seconds"], "spp_per_s": d["spp_per_s"],
                             "mpath_segments_per_s": d["mpath_segments_per_s"], "last_allreduce_ms": d["last_allreduce_ms"],
                             "how": f"cuda_path_tracer scenes/cornell.json --gpus {world} --iters {args.strong_iters}: wall clock around "
                                    "render + all-reduce + D2H of the SUM image (first-call allocation included)"}
        except Exception as e:                                     # noqa: BLE001
            out["render"] = {"error": str(e)[:200]}
        exe = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "stream_compaction_test")
        try:
            d = run_json_tool([exe, "28", "--gpus", str(world), "--json"], 900)
            n, ms = d["n"], d["ms"]
            out["scan_compact"] = {"n": n, "failures": d["failures"], "sharded_scan_pot_ms": ms.get("sharded_scan_pot"),
                                   "sharded_scan_gbs": 8.0 * n / ms["sharded_scan_pot"] / 1e6 if ms.get("sharded_scan_pot") else None,
                                   "sharded_compact_pot_ms": ms.get("sharded_compact_pot"), "all_gather_ms": ms.get("sharded_scan_all_gather"),
                                   "single_gpu_scan_pot_ms": ms.get("thrust_scan_pot"),
                                   "how": f"stream_compaction_test 28 --gpus {world}: host arrays in and out, device time from the first shard "
                                          "kernel to the last (all-gather included, copies excluded), max over devices; every result compared "
                                          "with CPU::scan / CPU::compactWithoutScan by the harness"}
        except Exception as e:                                     # noqa: BLE001
            out["scan_compact"] = {"error": str(e)[:200]}
        store.set("cxx_host_done", "1")
    else:
        store.wait(["cxx_host_done"], datetime.timedelta(minutes=25))
    return {"cxx_host": out} if rank == 0 else {}


def run_own(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from cuda_pathtracer_b200 import pathtracer as pt

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    peak, peak_kind = hbm_peak()
    sc = load_scene()
    P = sc.width * sc.height
    I = args.iters_per_step
    r = pt.Renderer(sc)
    gui = default_gui(pt)
    gui.batch = args.batch
    r.set_options(gui)
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream
    img_ptr, img_n = r.sum_device_ptr()
    img = torch.as_tensor(_DevArray(img_ptr, img_n), device=f"cuda:{local_rank}")
    red = torch.empty_like(img) if world > 1 else None      # the all-reduced image (the private sums stay valid)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    it = [rank + 1]

    def step():
        r.render(it[0], I, world, stream=sptr, sync=False)
        it[0] += I * world
        if world > 1:                                          # the one collective of the path, every step
            red.copy_(img)
            dist.all_reduce(red)

    for _ in range(args.warmup):
        step()
    barrier()
    r.reset()
    r.set_profiling(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = e0.elapsed_time(e1)
    st = r.stats()
    r.set_profiling(False)
    tot = torch.tensor([float(st["path_segments"]), float(st["kernel_launches"])], dtype=torch.float64, device="cuda")
    tmax = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tot)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    segments, launches, ms = float(tot[0]), int(tot[1]), float(tmax[0])

    # ---- e2e: host buffers, H2D scene upload + (N > 1: all-reduce) + D2H read of the (reduced) image, all timed ----
    host_img_t = torch.empty((sc.height, sc.width, 3), dtype=torch.float32, pin_memory=True)
    host_img = host_img_t.numpy()
    h2d = 272 * len(sc.geoms) + 48 * len(sc.materials) + 84
    d2h = host_img.nbytes

    def e2e_step():
        r.update_scene(sc, stream=sptr)
        r.reset()
        r.render(it[0], I, world, stream=sptr, sync=False)
        it[0] += I * world
        if world > 1:
            red.copy_(img)
            dist.all_reduce(red)
            host_img_t.view(-1).copy_(red)                     # D2H of the REDUCED image (synchronous: pinned destination)
            torch.cuda.synchronize()
        else:
            r.read_sum(host_img)
        return r.stats()["path_segments"]

    for _ in range(2):
        e2e_step()
    barrier()
    t0, seg_e2e = time.perf_counter(), 0
    for _ in range(args.steps):
        seg_e2e += e2e_step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    e2e_t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    e2e_s = torch.tensor([float(seg_e2e)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(e2e_s)
    e2e_value = float(e2e_s[0]) / float(e2e_t[0]) / 1e6

    bounce_ms, bounce_n = st["bounce_kernel_ms"], st["bounce_kernel_launches"]
    r.close()
    del img
    extras = {}
    if world > 1 and not args.skip_extras:
        extras.update(render_legs(torch, dist, rank, world, local_rank, args))
        extras.update(sharded_scan_compact_legs(torch, dist, rank, world, lg=args.sharded_lg))
        extras.update(cxx_host_leg(torch, dist, rank, world, args))
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    traffic, ncu = None, {}  # DRAM bytes / warp instructions per launch of the dominant kernel from the committed ncu capture of this workload
    prof_json = "k_bounce_traffic.json" if SCENE == "cornell" else f"{SCENE}_traffic.json"
    try:
        ncu = json.load(open(os.path.join(REPO, "profiles", prof_json)))
        traffic = ncu["bytes_per_launch"]
    except Exception:
        pass
    alg_bytes = st["path_segments"] * ALG_BYTES_PER_SEGMENT + st["iterations"] * P * ALG_BYTES_PER_PIXEL_ITER
    achieved = alg_bytes / (bounce_ms * 1e-3) / 1e9 if bounce_ms > 0 else 0.0
    cfg = workload_config(sc)
    line = {
        "metric": "path_segments_per_s", "value": segments / (ms * 1e-3) / 1e6, "unit": "Mpath-segments/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "spp_per_s": args.steps * I * world / (ms * 1e-3),
        "segments_per_sample_path": segments / (args.steps * I * world * P),
        "config": cfg,
        "run": {"iters_per_step": I, "iterations_timed": args.steps * I * world, "timed_region_s": ms * 1e-3,
                "batch": args.batch or "library default (32M paths in flight, at most 32 iterations)"},
        "e2e": {"value": e2e_value, "unit": "Mpath-segments/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "includes_allreduce": world > 1},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": {"kernel": "k_bounce (closest hit + shade + look-back compaction, one launch per bounce)" if SCENE.startswith("cornell")
                               else "the bounce kernels of this scene (closest hit with dynamic ray fetch + k_shade)",
                     "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "peak_kind": peak_kind, "traffic": traffic,
                     "alg_bytes_per_launch": alg_bytes / bounce_n if bounce_n else None,
                     "avg_launch_ms": bounce_ms / bounce_n if bounce_n else None, "launches": bounce_n,
                     "share_of_step": bounce_ms / ms if ms > 0 else None,
                     "note": "algorithmic bytes = 96 B per path segment + 24 B per pixel per iteration; the kernel "
                             "is FP32-issue bound at this scene size (IEEE div/sqrt sequences kept for bit parity), "
                             "see DESIGN.md"},
    }
    if ncu.get("warp_inst_per_launch") and bounce_n and clocks and clocks.get("sm_mhz"):
        # the ceiling that actually binds the bounce kernels: warp-instruction issue slots (4 schedulers per SM, one
        # instruction per clock each).  Instructions per launch come from the committed ncu capture (same workload, same
        # batch structure), the launch time and the SM clock are measured live.
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        peak_issue = sms * 4 * float(clocks["sm_mhz"]) * 1e6 / 1e9
        ach_issue = ncu["warp_inst_per_launch"] / (bounce_ms * 1e-3 / bounce_n) / 1e9
        line["roofline_issue"] = {"kernel": ncu.get("kernel", "k_bounce"), "bound": "issue", "achieved": ach_issue, "peak": peak_issue,
                                  "unit": "Gwarp-inst/s", "frac": ach_issue / peak_issue,
                                  "warp_inst_per_launch": ncu["warp_inst_per_launch"],
                                  "ncu_issue_active_pct": ncu.get("issue_active_pct_time_weighted"),
                                  "ncu_threads_per_warp_inst": ncu.get("threads_per_inst_inst_weighted"),
                                  "source": f"profiles/{prof_json} (ncu --set full) x live CUDA-event launch time"}
    line.update(extras)
    if world == 1 and not args.skip_extras:
        line["cpu_baseline"] = cpu_baseline(args.cpu_iters)
        line.update(reference_cuda_build(args.ref_cuda_iters))
        line["scan"] = scan_bench(torch, peak)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--iters-per-step", type=int, default=250,
                    help="iterations per render call; 20 steps x 250 = BASELINE config 2's 5000 iterations")
    ap.add_argument("--ref-iters-per-step", type=int, default=REF_ITERS_PER_STEP)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--scene", default="cornell",
                    help="scenes/<scene>.json; the contract benchmark is the default (BASELINE configs[1]); "
                         "e.g. --scene synthetic_4096 --iters-per-step 8 --skip-extras is configs[4] (per-GPU share)")
    ap.add_argument("--cpu-iters", type=int, default=8)
    ap.add_argument("--ref-cuda-iters", type=int, default=200)
    ap.add_argument("--strong-iters", type=int, default=5000)
    ap.add_argument("--config5-iters", type=int, default=6, help="iterations per rank per step of the config5 leg (N > 1)")
    ap.add_argument("--config5-steps", type=int, default=3)
    ap.add_argument("--sharded-lg", type=int, default=30)
    ap.add_argument("--skip-extras", action="store_true",
                    help="no cpu_baseline / reference CUDA build / drop-in / scan legs (N = 1), no config5 / sharded / strong legs (N > 1)")
    args = ap.parse_args()
    global SCENE
    SCENE = args.scene
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_own(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
