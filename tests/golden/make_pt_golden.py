"""Generates the path-tracer golden fixtures.

Step 1 (here, CPU only): scene POD dumps produced by the reference's OWN loader
(oracle/_ref/ref_scene_dump = /root/reference/path_tracer/src/scene.cpp + BVH_tree.cpp +
utilities.cpp compiled unmodified) for the scenes under scenes/:
    python tests/golden/make_pt_golden.py scenes
Step 2 (on a B200, under gpurun): float SUM images rendered by the reference's own CUDA build
(oracle/_ref/ref_pathtracer, unmodified pathtrace.cu / intersections.cu / interactions.cu) on
the small scenes, written to gpurun_out/ and then copied into tests/golden/ by
    python tests/golden/make_pt_golden.py renders [--only-missing] [--collect]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(REPO, "oracle", "_ref")

# scene -> extra args (--refract MATERIAL_ID IOR: material ids follow the alphabetical key order)
SCENES = {
    "cornell": [],
    "cornell_mirror": [],
    "cornell_small": [],
    "cornell_small_mirror": [],
    "cornell_tiny": [],
    "cornell_glass_small": ["--refract", "4", "1.5"],
    "cornell_glass_1080p": ["--refract", "4", "1.5"],
    "synthetic_600_small": sum((["--refract", str(i), "1.5"] for i in range(24, 32)), []),
    "mesh_room_small": [],
    "mesh_room_small_notex": [],
    "mesh_room": [],
    "mesh_stress": [],      # loader / BVH-builder stress (scenes/make_scenes.py stress_assets)
    "jpeg_textures": [],
    "png_textures": [],     # PNG decoding parity with stb_image (scenes/make_png_fixtures.py)    # JPEG decoding parity with stb_image (scenes/make_jpeg_fixtures.py)
    "synthetic_4096": sum((["--refract", str(i), "1.5"] for i in range(24, 32)), []),
    "synthetic_4096_small": sum((["--refract", str(i), "1.5"] for i in range(24, 32)), []),
}
# (scene, iterations, extra renderer flags, tag)
RENDERS = [
    ("cornell_small", 16, [], "default"),
    ("cornell_small", 16, ["--no-ssaa", "--no-dof", "--no-rr"], "plain"),
    ("cornell_small_mirror", 16, [], "default"),
    ("cornell_glass_small", 16, [], "default"),
    ("synthetic_600_small", 4, [], "default"),
    # material sort on (tests/pt_cases.py SORT_RENDERS)
    ("cornell_glass_small", 16, ["--sort"], "sort"),
    ("synthetic_600_small", 4, ["--sort"], "sort"),
    ("cornell_small", 16, ["--sort", "--no-ssaa", "--no-dof"], "sortplain"),
    # triangle meshes + BVH + textures (the room.json path)
    ("mesh_room_small", 8, [], "default"),
    ("mesh_room_small", 8, ["--sort", "--no-ssaa", "--no-dof"], "sortplain"),
    # useBVHtree = false (and useBBox = false): meshIntersectionTest instead of the BVH walk
    # (untextured: that mode's texture coordinates leave [0, 1] and the reference then reads outside the image)
    ("mesh_room_small_notex", 8, ["--no-bvh"], "nobvh"),
    ("mesh_room_small_notex", 8, ["--no-bvh", "--no-bbox", "--sort", "--no-ssaa", "--no-dof"], "nobvhsort"),
    # BASELINE config 5's scene (4096 primitives, 64 materials) at 480x270, plain and sorted
    ("synthetic_4096_small", 2, [], "default"),
    ("synthetic_4096_small", 2, ["--sort"], "sort"),
    # useThrustPartition = true (pathtrace.cu:416-420,499-503): thrust::stable_partition instead of mark / scan / keep
    ("cornell_small", 16, ["--thrust-partition"], "thrustpart"),
    ("cornell_glass_small", 16, ["--thrust-partition", "--sort"], "thrustpartsort"),
]


def scenes():
    for name, extra in SCENES.items():
        subprocess.run([os.path.join(REF, "ref_scene_dump"), os.path.join(REPO, "scenes", name + ".json"),
                        os.path.join(HERE, name + ".scene.bin")] + extra, check=True, stdout=subprocess.DEVNULL)
        print("wrote", name + ".scene.bin")


def renders(collect, only_missing=False):
    out = os.path.join(REPO, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    for scene, iters, flags, tag in RENDERS:
        name = f"{scene}.{tag}.{iters}it.refsum.bin"
        if only_missing and os.path.exists(os.path.join(HERE, name)):
            continue
        if collect:
            if os.path.exists(os.path.join(out, name)):
                os.replace(os.path.join(out, name), os.path.join(HERE, name))
            continue
        subprocess.run([os.path.join(REF, "ref_pathtracer"), os.path.join(REPO, "scenes", scene + ".json"),
                        "--iters", str(iters), "--out", os.path.join(out, name)] + flags + SCENES[scene], check=True)


if __name__ == "__main__":
    if sys.argv[1] == "scenes":
        scenes()
    else:
        renders("--collect" in sys.argv, "--only-missing" in sys.argv)
