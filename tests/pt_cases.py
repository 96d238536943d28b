"""Shared case tables of the path-tracer tests."""
import os

import numpy as np

from cuda_pathtracer_b200 import pathtracer as pt

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# (scene, iterations, reference-renderer flags, tag) -> tests/golden/<scene>.<tag>.<it>it.refsum.bin,
# float SUM images written by the reference's own CUDA build (tests/golden/make_pt_golden.py)
RENDERS = [
    ("cornell_small", 16, [], "default"),
    ("cornell_small", 16, ["--no-ssaa", "--no-dof", "--no-rr"], "plain"),
    ("cornell_small_mirror", 16, [], "default"),
    ("cornell_glass_small", 16, [], "default"),
    ("synthetic_600_small", 4, [], "default"),
    ("mesh_room_small", 8, [], "default"),                # triangle meshes, BVH, textures
    ("mesh_room_small_notex", 8, ["--no-bvh"], "nobvh"),  # useBVHtree = false: meshIntersectionTest
    ("synthetic_4096_small", 2, [], "default"),           # BASELINE config 5's scene at 480x270 (k_isect_dyn, 1383-node tree)
    ("cornell_small", 16, ["--thrust-partition"], "thrustpart"),   # useThrustPartition
]
# material sort on (thrust::sort_by_key in the reference, pathtrace.cu:479-491): the paths differ
# from the unsorted ones because the shading seed is the slot index AFTER sorting
SORT_RENDERS = [
    ("cornell_glass_small", 16, ["--sort"], "sort"),
    ("synthetic_600_small", 4, ["--sort"], "sort"),
    ("cornell_small", 16, ["--sort", "--no-ssaa", "--no-dof"], "sortplain"),
    ("mesh_room_small", 8, ["--sort", "--no-ssaa", "--no-dof"], "sortplain"),
    ("mesh_room_small_notex", 8, ["--no-bvh", "--no-bbox", "--sort", "--no-ssaa", "--no-dof"], "nobvhsort"),
    ("synthetic_4096_small", 2, ["--sort"], "sort"),      # config 5's 64 materials through the material bucketing
    ("cornell_glass_small", 16, ["--thrust-partition", "--sort"], "thrustpartsort"),
]
# per-stage vectors from the reference's own kernels (tests/golden/make_pt_stage_golden.py)
STAGES = [("cornell_tiny", "default", []), ("cornell_tiny", "plain", ["--no-ssaa", "--no-dof", "--no-rr"]),
          ("cornell_glass_small", "default", []), ("mesh_room_small", "default", []),
          ("mesh_room_small", "nobvh", ["--no-bvh"])]


def load_scene(name):
    return pt.Scene.from_pod_file(os.path.join(GOLD, name + ".scene.bin"))


def gui_from_flags(flags, batch=0):
    g = pt.GuiDataContainer()
    g.SSAA = "--no-ssaa" not in flags
    g.DoF = "--no-dof" not in flags
    g.russianRoulette = "--no-rr" not in flags
    g.sortbyMaterial = "--sort" in flags
    g.useBVHtree = "--no-bvh" not in flags
    g.useBBox = "--no-bbox" not in flags
    g.useThrustPartition = "--thrust-partition" in flags
    g.batch = batch
    return g


def texture_reads_in_bounds(scene, isects):
    """Rows of ShadeableIntersection records whose shading does NOT make the reference's Texture::get_color
    read outside the image (sceneStructs.h:175-188 has no bounds check; this library clamps the index).  With
    useBVHtree = false the texture coordinate is az*uv0 + uv1 + uv2 (intersections.cu:164), which leaves
    [0, 1] -- what the reference shades there is whatever lies next to the texture in device memory."""
    isects = np.ascontiguousarray(isects, dtype=np.float32)
    ok = np.ones(len(isects), dtype=bool)
    mat = isects[:, 4].view(np.int32)
    for m, material in enumerate(scene.materials):
        tid = material.texture_id
        rows = (mat == m) & (isects[:, 0] > 0)
        if tid < 0 or not rows.any():
            continue
        t = scene.textures[tid]
        w, h = np.float32(t.width), np.float32(t.height)
        u, v = isects[rows, 5], isects[rows, 6]
        x = np.minimum(w * u, w - np.float32(1)).astype(np.int64)
        y = np.minimum(h * (np.float32(1) - v), h - np.float32(1)).astype(np.int64)
        idx = y * t.width + x
        ok[rows] = (idx >= 0) & (idx < t.width * t.height)
    return ok


def ref_sum(scene, iters, tag, shape):
    return np.fromfile(os.path.join(GOLD, f"{scene}.{tag}.{iters}it.refsum.bin"), dtype=np.float32).reshape(shape)


def bits_equal(a, b):
    """Bitwise equality of float32 arrays; a NaN matches a NaN whatever its payload (the GPU's
    canonical NaN is 0x7fffffff, x86's 0xffc00000: total internal reflection in glm::refract
    yields a NaN direction in the reference, interactions.cu:68)."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    if a.shape != b.shape:
        return False
    same = a.view(np.uint32) == b.view(np.uint32)
    return bool(np.all(same | (np.isnan(a) & np.isnan(b))))


def image_metrics(mine_sum, ref_sum_img, iters):
    """north_star's tolerance is stated on the per-pixel MEAN image (sum / iterations), peak 1.0."""
    d = (mine_sum.astype(np.float64) - ref_sum_img.astype(np.float64)) / iters
    mae = float(np.abs(d).mean())
    mse = float((d ** 2).mean())
    psnr = float("inf") if mse == 0 else 10.0 * np.log10(1.0 / mse)
    return mae, psnr
