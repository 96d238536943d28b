"""Throughput of the path tracer on every path-tracer configuration BASELINE.json names (and the
mesh scene of SURVEY 8f N3), with the reference's own CUDA build timed beside it on the same GPU.

    python benchmarks/pt_configs.py [--only NAME] [--no-ref] > gpurun_out/pt_configs.jsonl

Per configuration: W warm-up renders, then `iters` iterations timed with CUDA events on the
launching stream (the library's own last_render_ms is the same measurement), the float SUM image
stays on the device (reading it back is part of bench.py's e2e leg, not of this table).
Reference: oracle/_ref/ref_pathtracer (unmodified reference sources + headless shim) on the same
scene file and flags, wall clock around its pathtrace() loop, fewer iterations where one
iteration takes seconds.  Path segments are counted by this library (same definition as SURVEY
8d); the reference traces the same paths (bit-identical images, tests/test_pt_gpu.py), so its
segments/s = our segments per iteration x its iterations/s.
One JSON line per configuration.
"""
import argparse
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

from cuda_pathtracer_b200 import pathtracer as pt  # noqa: E402
from pt_cases import gui_from_flags  # noqa: E402

GLASS = ["--refract", "4", "1.5"]
SYN = sum((["--refract", str(i), "1.5"] for i in range(24, 32)), [])

# name, scene, flags (reference shim spelling), first-bounce cache, iterations (ours), iterations (reference), reference extra
CONFIGS = [
    ("config2 cornell 800x800 d8 defaults", "cornell", [], False, 192, 60, []),
    ("config2b cornell_mirror 800x800 d8 defaults", "cornell_mirror", [], False, 192, 60, []),
    ("config4 glass 1920x1080 d12 sort+cache (SSAA/DoF off)", "cornell_glass_1080p", ["--sort", "--no-ssaa", "--no-dof"], True, 96, 20, GLASS),
    ("config4b glass 1920x1080 d12 sort only (SSAA/DoF off)", "cornell_glass_1080p", ["--sort", "--no-ssaa", "--no-dof"], False, 96, 0, GLASS),
    ("config4c glass 1920x1080 d12 no sort, no cache, defaults", "cornell_glass_1080p", [], False, 96, 20, GLASS),
    ("config5 synthetic 4096 prims 3840x2160 d8 (one GPU's share)", "synthetic_4096", [], False, 8, 2, SYN),
    ("N3 mesh_room 1200x1200 d8 (4 OBJ meshes, BVH, 2 textures)", "mesh_room", [], False, 96, 10, []),
    ("N3b mesh_room sort", "mesh_room", ["--sort"], False, 48, 10, []),
]


def ours(scene_name, flags, cache, iters, geom_bvh=0):
    sc = pt.Scene.from_pod_file(os.path.join(REPO, "tests", "golden", scene_name + ".scene.bin"))
    r = pt.Renderer(sc)
    g = gui_from_flags(flags, batch=0)
    g.firstBounceCache = cache
    g.geomBVH = geom_bvh
    r.set_options(g)
    r.render(1, min(iters, 8))                                   # warm-up (allocations, caches)
    r.reset()
    r.render(1, iters)
    st = r.stats()
    ms = st["last_render_ms"]
    out = {"iters": iters, "ms": ms, "spp_per_s": iters / ms * 1e3, "segments": st["path_segments"],
           "mseg_per_s": st["path_segments"] / ms / 1e3, "launches": st["kernel_launches"],
           "resolution": [sc.width, sc.height], "depth": sc.traceDepth}
    r.close()
    return out


def reference(scene_name, flags, extra, iters):
    exe = os.path.join(REPO, "oracle", "_ref", "ref_pathtracer")
    scene = os.path.join(REPO, "scenes", scene_name + ".json")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_pathtracer not built"}
    try:
        subprocess.run([exe, scene, "--iters", "2"] + flags + extra, capture_output=True, text=True, timeout=300)
        r = subprocess.run([exe, scene, "--iters", str(iters)] + flags + extra, capture_output=True, text=True, timeout=900)
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return {"iters": iters, "spp_per_s": d["spp_per_s"], "ms_per_iter": d["ms_per_iter"]}
    except Exception as e:                                       # noqa: BLE001
        return {"error": str(e)[:200]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--geom-bvh", type=int, default=0, help="0 auto, 1 always, -1 never (linear scan)")
    args = ap.parse_args()
    for name, scene, flags, cache, iters, ref_iters, extra in CONFIGS:
        if args.only and args.only not in name:
            continue
        row = {"config": name, "scene": scene, "flags": flags, "first_bounce_cache": cache, "geom_bvh": args.geom_bvh,
               "b200": ours(scene, flags, cache, iters, args.geom_bvh)}
        if ref_iters and not args.no_ref:
            ref = reference(scene, flags, extra, ref_iters)
            if "spp_per_s" in ref:
                seg_per_iter = row["b200"]["segments"] / row["b200"]["iters"]
                ref["mseg_per_s"] = seg_per_iter * ref["spp_per_s"] / 1e6
                row["speedup_spp"] = row["b200"]["spp_per_s"] / ref["spp_per_s"]
            row["reference_cuda_build"] = ref
        print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
