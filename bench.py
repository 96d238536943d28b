"""bench.py -- headline benchmark of the B200-native path tracer + scan/compact library.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Metric (BASELINE.json): Mpath-segments/s (and spp/s) on the Cornell box 800x800, depth 8, reference
defaults (Russian roulette, SSAA and DoF on); scan / compact GB/s at 2^28 ints ride along in the
same JSON line.  A STEP is one render call of `--iters-per-step` sample iterations over all
640 000 pixels (one wavefront pass of the hot path over one batch of camera paths).

  value        whole-job path-segments/s with the scene resident in HBM (device-timed, CUDA events,
               max over ranks); a path segment = one live path entering intersection in one bounce
               (SURVEY.md 8d).  N > 1: rank g traces iterations g+1, g+1+N, ... (weak scaling: every
               rank traces --iters-per-step iterations per step) and one NCCL all-reduce of the
               float accumulation image closes the timed region.
  e2e          the same metric through the C ABI the reference's pathtrace.h maps onto, with HOST
               buffers: every step re-uploads the scene (H2D, pinned), renders, and reads the float
               SUM image back into pinned host memory (D2H), all inside the timed region.
  roofline     dominant kernel k_bounce against the measured HBM peak: algorithmic bytes (96 B per
               path segment + 24 B per pixel per iteration) / its summed CUDA-event time.
  cpu_baseline the CPU oracle (oracle/pt_oracle.cpp, a port of the reference's algorithm) on all
               host threads, on a bounded sample of the same workload (rank 0, N = 1 only).

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
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

SCENE = "cornell"
ALG_BYTES_PER_SEGMENT = 96.0        # read + write one 48-B path state
ALG_BYTES_PER_PIXEL_ITER = 24.0     # read-modify-write of the 12-B accumulator


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
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def load_scene():
    from cuda_pathtracer_b200 import pathtracer as pt
    return pt.Scene.from_pod_file(os.path.join(REPO, "tests", "golden", SCENE + ".scene.bin"))


def default_gui(pt):
    return pt.GuiDataContainer()          # reference defaults: RR, SSAA, DoF on; aperture .1; focal 10


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


def reference_cuda_build(n_iters):
    """the reference's OWN CUDA build (oracle/_ref/ref_pathtracer: unmodified pathtrace.cu etc.,
    built in the build container) on this GPU: the >= 5x comparison north_star names"""
    exe = os.path.join(REPO, "oracle", "_ref", "ref_pathtracer")
    scene = os.path.join(REPO, "scenes", SCENE + ".json")
    if not (os.path.exists(exe) and os.path.exists(scene)):
        return None
    try:
        subprocess.run([exe, scene, "--iters", "3"], capture_output=True, text=True, timeout=120)   # warm-up
        r = subprocess.run([exe, scene, "--iters", str(n_iters)], capture_output=True, text=True, timeout=600)
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return {"spp_per_s": d["spp_per_s"], "ms_per_iter": d["ms_per_iter"], "iters": n_iters,
                "how": "oracle/_ref/ref_pathtracer (reference sources unmodified, headless shim), wall clock "
                       "around its pathtrace() loop"}
    except Exception as e:                                     # noqa: BLE001
        return {"error": str(e)[:200]}


def scan_bench(torch, peak):
    """scan / compact at 2^28 ints (BASELINE metric, second half): device-resident, CUDA events on
    the launching stream, inputs (1 GiB) far larger than L2."""
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

    for name, variant in (("scan_lookback", sc.LOOKBACK), ("scan_efficient", sc.EFFICIENT)):
        ms = timed(lambda: ctx.scan(n, out, a, variant, stream=stream.cuda_stream))
        res[name] = {"ms": ms, "gbs": 8.0 * n / ms / 1e6, "frac": 8.0 * n / ms / 1e6 / peak}
    ms = timed(lambda: ctx.compact(n, out, b, cnt, stream=stream.cuda_stream))
    k = int(cnt.item())
    res["compact"] = {"ms": ms, "gbs": (4.0 * n + 4.0 * k) / ms / 1e6, "frac": (4.0 * n + 4.0 * k) / ms / 1e6 / peak,
                      "kept": k}
    ctx.close()
    del a, b, out
    torch.cuda.empty_cache()
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
    sc, o, gui = load_scene(), PtOracle(), None
    gui = default_gui(pt)
    threads = os.cpu_count() or 1
    it = 1
    for _ in range(args.warmup):
        o.render(sc, gui, it, 1, threads=threads); it += 1
    seg, t0 = 0, time.perf_counter()
    for _ in range(args.steps):
        _, s = o.render(sc, gui, it, 1, threads=threads); it += 1
        seg += s
    dt = time.perf_counter() - t0
    v = seg / dt / 1e6
    line = {"impl": "reference", "metric": "path_segments_per_s", "value": v, "unit": "Mpath-segments/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "spp_per_s": args.steps / dt,
            "config": {"workload": f"{SCENE} 800x800, trace depth 8, reference defaults (RR+SSAA+DoF)",
                       "iters_per_step": 1, "host_threads": threads},
            "cpu_baseline": {"value": v, "unit": "Mpath-segments/s", "cores": threads, "kind": "port",
                             "sample": f"{args.steps} steps x 1 iteration of the full 800x800 workload, "
                                       "oracle/pt_oracle.cpp (the reference has no CPU renderer; this is the CPU "
                                       "port of its algorithm, bit-identical to its CUDA build)"},
            "e2e": {"value": v, "unit": "Mpath-segments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    it = [rank + 1]

    def step():
        r.render(it[0], I, world, stream=sptr, sync=False)
        it[0] += I * world

    for _ in range(args.warmup):
        step()
    if world > 1:
        dist.all_reduce(img)                                   # warm NCCL up
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
    if world > 1:
        dist.all_reduce(img)                                   # the one collective of the path
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

    # ---- e2e: host buffers, H2D scene upload + D2H image read inside the timed region -------------
    host_img = torch.empty((sc.height, sc.width, 3), dtype=torch.float32, pin_memory=True).numpy()
    h2d = 272 * len(sc.geoms) + 48 * len(sc.materials) + 84
    d2h = host_img.nbytes

    def e2e_step():
        r.update_scene(sc, stream=sptr)
        r.reset()
        r.render(it[0], I, world, stream=sptr, sync=False)
        it[0] += I * world
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
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    traffic, ncu = None, {}  # DRAM bytes / warp instructions per k_bounce launch from the committed ncu capture of this workload
    try:
        ncu = json.load(open(os.path.join(REPO, "profiles", "k_bounce_traffic.json")))
        traffic = ncu["bytes_per_launch"]
    except Exception:
        pass
    alg_bytes = st["path_segments"] * ALG_BYTES_PER_SEGMENT + st["iterations"] * P * ALG_BYTES_PER_PIXEL_ITER
    achieved = alg_bytes / (bounce_ms * 1e-3) / 1e9 if bounce_ms > 0 else 0.0
    line = {
        "metric": "path_segments_per_s", "value": segments / (ms * 1e-3) / 1e6, "unit": "Mpath-segments/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "spp_per_s": args.steps * I * world / (ms * 1e-3),
        "segments_per_sample_path": segments / (args.steps * I * world * P),
        "config": {"workload": f"{SCENE} {sc.width}x{sc.height}, trace depth {sc.traceDepth}, reference defaults (RR+SSAA+DoF)",
                   "iters_per_step": I, "batch": args.batch or "library default (32M paths in flight, at most 32 iterations)",
                   "l2": "no flush: the two path pools of a batch (2 x 48 B x 640000 x batch) exceed the 126 MB L2",
                   "parallelism": "replicated scene, iterations split across ranks (stride N), one NCCL "
                                  "all-reduce of the float accumulation image" if world > 1 else "single GPU"},
        "e2e": {"value": e2e_value, "unit": "Mpath-segments/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
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
    if SCENE != "cornell":
        line["roofline"]["traffic"] = None               # the committed ncu capture is of the Cornell workload
    if SCENE == "cornell" and ncu.get("warp_inst_per_launch") and bounce_n and clocks.get("sm_mhz"):
        # the ceiling that actually binds k_bounce: warp-instruction issue slots (4 schedulers per SM, one instruction
        # per clock each).  Instructions per launch come from the committed ncu capture (same workload, same batch
        # structure), the launch time and the SM clock are measured live.
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        peak_issue = sms * 4 * float(clocks["sm_mhz"]) * 1e6 / 1e9
        ach_issue = ncu["warp_inst_per_launch"] / (bounce_ms * 1e-3 / bounce_n) / 1e9
        line["roofline_issue"] = {"kernel": "k_bounce", "bound": "issue", "achieved": ach_issue, "peak": peak_issue,
                                  "unit": "Gwarp-inst/s", "frac": ach_issue / peak_issue,
                                  "warp_inst_per_launch": ncu["warp_inst_per_launch"],
                                  "ncu_issue_active_pct": ncu.get("issue_active_pct_time_weighted"),
                                  "ncu_threads_per_warp_inst": ncu.get("threads_per_inst_inst_weighted"),
                                  "source": "profiles/k_bounce_traffic.json (ncu --set full) x live CUDA-event launch time"}
    if world == 1 and not args.skip_extras:
        line["cpu_baseline"] = cpu_baseline(args.cpu_iters)
        ref = reference_cuda_build(args.ref_cuda_iters)
        if ref is not None:
            line["reference_cuda_build"] = ref
        line["scan"] = scan_bench(torch, peak)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--iters-per-step", type=int, default=48)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--scene", default="cornell",
                    help="tests/golden/<scene>.scene.bin; the contract benchmark is the default (BASELINE configs[1]); "
                         "e.g. --scene synthetic_4096 --iters-per-step 8 --skip-extras is configs[4] (per-GPU share)")
    ap.add_argument("--cpu-iters", type=int, default=8)
    ap.add_argument("--ref-cuda-iters", type=int, default=100)
    ap.add_argument("--skip-extras", action="store_true", help="no cpu_baseline / reference CUDA build / scan legs")
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
