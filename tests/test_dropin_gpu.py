"""Drop-in boundary tests: the reference's OWN host code (unmodified scene loader + headless main
loop; unmodified scan test harness stream_compaction/src/main.cpp) linked against THIS repo's
libraries through the bindings in integration/ -- built in the build container by oracle/Makefile
into oracle/_ref/dropin_* (the reference sources do not travel; the binaries do)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(REPO, "oracle", "_ref")


def need(name):
    p = os.path.join(REF, name)
    if not os.path.exists(p):
        pytest.skip(f"{name} was not built (needs /root/reference at build time)")
    return p


@pytest.mark.parametrize("scene,flags", [("cornell_tiny", []), ("cornell_small", ["--no-ssaa", "--no-dof"]),
                                         ("cornell_glass_small", ["--refract", "4", "1.5"]),
                                         # OBJ meshes + BVH + textures built by the reference's own loader
                                         ("mesh_room_small", []), ("mesh_room_small", ["--sort"]),
                                         # guiData->useBVHtree = false (and useBBox = false) through the binding
                                         ("mesh_room_small_notex", ["--no-bvh"]),
                                         ("mesh_room_small_notex", ["--no-bvh", "--no-bbox", "--sort"]),
                                         # guiData->useThrustPartition = true (pathtrace.cu:499-503) on both sides
                                         ("cornell_small", ["--thrust-partition"]),
                                         ("cornell_glass_small", ["--refract", "4", "1.5", "--thrust-partition", "--sort"])])
def test_pathtrace_h_dropin_is_bit_identical(tmp_path, scene, flags):
    ref, mine = need("ref_pathtracer"), need("dropin_pathtracer")
    scene_json = os.path.join(REPO, "scenes", scene + ".json")
    outs = []
    for exe, tag in ((ref, "ref"), (mine, "b200")):
        o, p = str(tmp_path / f"{tag}.sum"), str(tmp_path / f"{tag}.pbo")
        subprocess.run([exe, scene_json, "--iters", "8", "--out", o, "--pbo", p] + flags, check=True,
                       stdout=subprocess.DEVNULL)
        outs.append((np.fromfile(o, dtype=np.uint32), np.fromfile(p, dtype=np.uint8)))
    assert outs[0][0].size > 0 and np.array_equal(outs[0][0], outs[1][0])      # float SUM image, bitwise
    assert np.array_equal(outs[0][1], outs[1][1])                              # uchar4 preview


def test_reference_scan_harness_passes_on_this_library():
    """stream_compaction/src/main.cpp, unmodified (SIZE = 2^28, POT and NPOT, time-seeded data):
    every GPU variant is compared with the reference's own CPU scan / compact by the harness."""
    exe = need("dropin_sc_harness")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600, input="\n")
    out = r.stdout
    assert "FAIL" not in out, out[-2000:]
    assert out.count("passed") >= 12, out[-2000:]
