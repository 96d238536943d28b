"""Diagnostic: render the small scenes with this library and compare the float SUM image with
the reference CUDA build's (gpurun_out/*.refsum.bin or tests/golden/*.refsum.bin).
    python benchmarks/pt_compare.py [--batch B]
"""
import argparse
import os
import sys

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests", "golden"))
from cuda_pathtracer_b200 import pathtracer as pt  # noqa: E402
from make_pt_golden import RENDERS  # noqa: E402


def gui_from_flags(flags, batch):
    g = pt.GuiDataContainer()
    g.SSAA = "--no-ssaa" not in flags
    g.DoF = "--no-dof" not in flags
    g.russianRoulette = "--no-rr" not in flags
    g.sortbyMaterial = "--sort" in flags
    g.batch = batch
    return g


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1)
    args = ap.parse_args()
    for scene, iters, flags, tag in RENDERS:
        name = f"{scene}.{tag}.{iters}it.refsum.bin"
        path = next((p for p in (os.path.join(REPO, "gpurun_out", name), os.path.join(REPO, "tests", "golden", name))
                     if os.path.exists(p)), None)
        sc = pt.Scene.from_pod_file(os.path.join(REPO, "tests", "golden", scene + ".scene.bin"))
        r = pt.Renderer(sc)
        r.set_options(gui_from_flags(flags, args.batch))
        r.render(1, iters)
        mine = r.read_sum()
        st = r.stats()
        line = f"{name}: segments {st['path_segments']} ms {st['last_render_ms']:.3f} mean {mine.mean()/iters:.5f}"
        if path:
            ref = np.fromfile(path, dtype=np.float32).reshape(mine.shape)
            d = np.abs(mine - ref) / iters
            neq = int((mine != ref).sum())
            mse = float((d ** 2).mean())
            psnr = 99.0 if mse == 0 else 10 * np.log10(1.0 / mse)
            line += f" | ref mean {ref.mean()/iters:.5f} unequal {neq}/{mine.size} MAE {d.mean():.3e} PSNR {psnr:.1f}"
            if neq:
                idx = np.argwhere(mine != ref)[:5]
                line += " first " + str([(tuple(i), float(mine[tuple(i)]), float(ref[tuple(i)])) for i in idx])
        print(line, flush=True)
        r.close()


if __name__ == "__main__":
    main()
