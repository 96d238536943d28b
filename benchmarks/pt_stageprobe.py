"""Diagnostic (tier T1): run one iteration stage by stage, the reference's own kernels
(oracle/_ref/ref_stage_probe) beside this library's on IDENTICAL inputs, and report the first
bitwise difference per stage.   python benchmarks/pt_stageprobe.py SCENE ITER [flags]"""
import os, subprocess, sys, tempfile
import numpy as np
import torch
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from cuda_pathtracer_b200 import pathtracer as pt

PROBE = os.path.join(REPO, "oracle/_ref/ref_stage_probe")


def run_probe(*args):
    subprocess.run([PROBE] + [str(a) for a in args], check=True)


def diff_report(name, a, b):
    a32, b32 = a.view(np.uint32), b.view(np.uint32)
    ne = np.argwhere((a32 != b32).any(axis=1)).ravel()
    print(f"  {name}: {len(ne)} / {len(a)} records differ", end="")
    if len(ne):
        i = ne[0]
        print(f"; first at {i}: mine {a[i].tolist()} ref {b[i].tolist()}", end="")
    print(flush=True)
    return len(ne)


def main():
    scene_name, it = sys.argv[1], int(sys.argv[2])
    flags = sys.argv[3:]
    scene_bin = scene_name if scene_name.endswith(".bin") else os.path.join(REPO, "tests/golden", scene_name + ".scene.bin")
    sc = pt.Scene.from_pod_file(scene_bin)
    r = pt.Renderer(sc)
    g = pt.GuiDataContainer()
    g.SSAA = "--no-ssaa" not in flags; g.DoF = "--no-dof" not in flags; g.russianRoulette = "--no-rr" not in flags
    r.set_options(g)
    tmp = tempfile.mkdtemp()
    f = lambda n: os.path.join(tmp, n)
    P = sc.width * sc.height
    # raygen
    run_probe("raygen", scene_bin, it, f("seg_ref.bin"), *flags)
    seg_ref = np.fromfile(f("seg_ref.bin"), dtype=np.float32).reshape(P, 12)
    d_seg = torch.zeros(P * 12, dtype=torch.float32, device="cuda")
    r.debug_raygen(it, d_seg.data_ptr()); torch.cuda.synchronize()
    diff_report("raygen", d_seg.cpu().numpy().reshape(P, 12), seg_ref)
    segs = seg_ref
    for bounce in range(sc.traceDepth):
        n = len(segs)
        if n == 0:
            break
        print(f"bounce {bounce}: {n} live")
        segs.tofile(f("seg_in.bin"))
        run_probe("intersect", scene_bin, f("seg_in.bin"), f("is_ref.bin"), *flags)
        is_ref = np.fromfile(f("is_ref.bin"), dtype=np.float32).reshape(n, 7)
        d_seg = torch.from_numpy(segs.copy()).cuda().reshape(-1)
        d_is = torch.zeros(n * 7, dtype=torch.float32, device="cuda")
        r.debug_intersect(n, d_seg.data_ptr(), d_is.data_ptr()); torch.cuda.synchronize()
        diff_report("intersect", d_is.cpu().numpy().reshape(n, 7), is_ref)
        run_probe("shade", scene_bin, it, f("seg_in.bin"), f("is_ref.bin"), f("seg_out.bin"), *flags)
        out_ref = np.fromfile(f("seg_out.bin"), dtype=np.float32).reshape(n, 12)
        d_is = torch.from_numpy(is_ref.copy()).cuda().reshape(-1)
        r.debug_shade(it, n, d_is.data_ptr(), d_seg.data_ptr()); torch.cuda.synchronize()
        diff_report("shade", d_seg.cpu().numpy().reshape(n, 12), out_ref)
        live = out_ref[:, 10].view(np.int32) != 0
        segs = np.ascontiguousarray(out_ref[live])


if __name__ == "__main__":
    main()
