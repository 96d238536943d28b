"""Diagnostic: compare against the reference CUDA build for individual option flags."""
import os, subprocess, sys
import numpy as np
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from cuda_pathtracer_b200 import pathtracer as pt
scene = sys.argv[1] if len(sys.argv) > 1 else "cornell_small"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 1
for flags in (["--no-ssaa"], ["--no-dof"], ["--no-rr"], ["--no-ssaa", "--no-dof"]):
    out = os.path.join(REPO, "gpurun_out", "probe.bin")
    subprocess.run([os.path.join(REPO, "oracle/_ref/ref_pathtracer"), os.path.join(REPO, "scenes", scene + ".json"),
                    "--iters", str(iters), "--out", out] + flags, check=True, stdout=subprocess.DEVNULL)
    sc = pt.Scene.from_pod_file(os.path.join(REPO, "tests/golden", scene + ".scene.bin"))
    r = pt.Renderer(sc)
    g = pt.GuiDataContainer(); g.batch = 1
    g.SSAA = "--no-ssaa" not in flags; g.DoF = "--no-dof" not in flags; g.russianRoulette = "--no-rr" not in flags
    r.set_options(g); r.render(1, iters)
    mine = r.read_sum(); ref = np.fromfile(out, dtype=np.float32).reshape(mine.shape)
    ne = np.argwhere((mine != ref).any(axis=2))
    print(flags, "unequal pixels", len(ne), "first", ne[:3].tolist(), "nonzero mine/ref", int((mine != 0).any(axis=2).sum()), int((ref != 0).any(axis=2).sum()),
          "first nonzero ref", np.argwhere((ref != 0).any(axis=2))[:2].tolist(), "mine", np.argwhere((mine != 0).any(axis=2))[:2].tolist(), flush=True)
    r.close()
