"""Generates per-stage golden vectors from the reference's OWN kernels (run on a B200, under
gpurun): oracle/_ref/ref_stage_probe launches generateRayFromCamera / computeIntersections /
shadeMaterials (unmodified pathtrace.cu) stage by stage; the arrays of the first bounces are
saved to gpurun_out/<scene>.<tag>.stage.npz and then copied into tests/golden/.
    python tests/golden/make_pt_stage_golden.py            # on the GPU box
"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
PROBE = os.path.join(REPO, "oracle", "_ref", "ref_stage_probe")
CASES = [("cornell_tiny", 3, [], "default", 4), ("cornell_tiny", 5, ["--no-ssaa", "--no-dof", "--no-rr"], "plain", 3),
         ("cornell_glass_small", 2, [], "default", 3), ("mesh_room_small", 2, [], "default", 4),
         ("mesh_room_small", 2, ["--no-bvh"], "nobvh", 2)]


def probe(*args):
    subprocess.run([PROBE] + [str(a) for a in args], check=True)


def main():
    out_dir = os.path.join(REPO, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    for scene, it, flags, tag, n_bounces in CASES:
        scene_bin = os.path.join(HERE, scene + ".scene.bin")
        tmp = tempfile.mkdtemp()
        f = lambda n: os.path.join(tmp, n)
        probe("raygen", scene_bin, it, f("seg.bin"), *flags)
        segs = np.fromfile(f("seg.bin"), dtype=np.float32).reshape(-1, 12)
        arrays = {"iteration": np.int32(it), "raygen": segs}
        for b in range(n_bounces):
            segs.tofile(f("in.bin"))
            probe("intersect", scene_bin, f("in.bin"), f("is.bin"), *flags)
            probe("shade", scene_bin, it, f("in.bin"), f("is.bin"), f("out.bin"), *flags)
            isects = np.fromfile(f("is.bin"), dtype=np.float32).reshape(-1, 7)
            out = np.fromfile(f("out.bin"), dtype=np.float32).reshape(-1, 12)
            arrays[f"b{b}_in"], arrays[f"b{b}_isect"], arrays[f"b{b}_out"] = segs, isects, out
            segs = np.ascontiguousarray(out[out[:, 10].view(np.int32) != 0])   # stable partition
        np.savez_compressed(os.path.join(out_dir, f"{scene}.{tag}.stage.npz"), **arrays)
        print("wrote", scene, tag, {k: v.shape for k, v in arrays.items() if k != "iteration"})


if __name__ == "__main__":
    main()
