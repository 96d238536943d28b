"""Small baseline-JPEG textures + the scene that names them (loader parity fixtures).

    python scenes/make_jpeg_fixtures.py          # writes scenes/Textures/j_*.jpg, scenes/jpeg_textures.json

The files are committed; this script only documents how they were made (Pillow's / OpenCV's libjpeg encoder).
They cover what the host decoder (cuda_pathtracer_b200/host/jpeg.cpp) must reproduce byte for byte
against stb_image, the reference's texture reader (sceneStructs.h:170-175): every chroma layout with
its own upsampling filter (4:4:4, 4:2:2, 4:2:0, 4:4:0, 4:1:1 -> the generic path), grayscale, widths
and heights that are not multiples of the MCU, one-pixel-wide images, restart intervals, optimised
(non-default) Huffman tables, low quality (large quantisers -> clamping in the inverse DCT), and progressive files.
tests/golden/jpeg_textures.scene.bin is the dump of the reference's own loader for the scene
(tests/golden/make_pt_golden.py); tests/test_host_cli.py compares this repo's loader with it."""
import json
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
TEX = os.path.join(HERE, "Textures")


def picture(w, h, seed):
    """smooth gradients + a few hard edges + noise: exercises DC-only and dense AC blocks alike"""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.stack([127 + 120 * np.sin(x / 5.0 + seed), 127 + 120 * np.cos(y / 3.0), (x * 7 + y * 13) % 256], -1)
    img[h // 3: h // 2, w // 4: w // 2] = (255, 0, 32)
    img += rng.normal(0, 12, img.shape)
    img[: max(1, h // 8)] = rng.integers(0, 256, (max(1, h // 8), w, 3))
    return np.clip(img, 0, 255).astype(np.uint8)


CASES = [
    # name, w, h, mode, save options
    ("j_444.jpg", 37, 21, "RGB", dict(quality=90, subsampling="4:4:4")),
    ("j_422.jpg", 50, 19, "RGB", dict(quality=85, subsampling="4:2:2")),
    ("j_420.jpg", 33, 35, "RGB", dict(quality=75, subsampling="4:2:0")),
    ("j_420_opt.jpg", 61, 47, "RGB", dict(quality=60, subsampling="4:2:0", optimize=True)),
    ("j_420_rst.jpg", 64, 48, "RGB", dict(quality=80, subsampling="4:2:0", restart_marker_blocks=3)),
    ("j_420_lowq.jpg", 40, 40, "RGB", dict(quality=5, subsampling="4:2:0")),
    ("j_440.jpg", 29, 31, "RGB", dict(cv2_440=True)),            # Pillow cannot write 4:4:0, OpenCV can
    ("j_411.jpg", 45, 18, "RGB", dict(quality=88, subsampling="4:1:1")),
    ("j_gray.jpg", 23, 17, "L", dict(quality=90)),
    ("j_1wide.jpg", 1, 9, "RGB", dict(quality=90, subsampling="4:2:0")),
    ("j_1high.jpg", 19, 1, "RGB", dict(quality=90, subsampling="4:2:0")),
    ("j_2wide_422.jpg", 2, 5, "RGB", dict(quality=90, subsampling="4:2:2")),
    # progressive (SOF2): spectral selection + successive approximation, end-of-band runs, restart intervals
    ("j_prog_420.jpg", 53, 41, "RGB", dict(quality=80, subsampling="4:2:0", progressive=True)),
    ("j_prog_444.jpg", 37, 21, "RGB", dict(quality=92, subsampling="4:4:4", progressive=True)),
    ("j_prog_422_opt.jpg", 50, 19, "RGB", dict(quality=60, subsampling="4:2:2", progressive=True, optimize=True)),
    ("j_prog_gray.jpg", 23, 17, "L", dict(quality=85, progressive=True)),
    ("j_prog_rst.jpg", 64, 48, "RGB", dict(quality=75, subsampling="4:2:0", progressive=True, restart_marker_blocks=2)),
    ("j_prog_lowq.jpg", 40, 40, "RGB", dict(quality=8, subsampling="4:2:0", progressive=True)),
]


def main():
    os.makedirs(TEX, exist_ok=True)
    mats = {"light": {"TYPE": "Emitting", "RGB": [1.0, 1.0, 1.0], "EMITTANCE": 5.0}}
    for k, (name, w, h, mode, opts) in enumerate(CASES):
        img = Image.fromarray(picture(w, h, k + 1))
        if mode == "L":
            img = img.convert("L")
        if opts.get("cv2_440"):
            import cv2
            cv2.imwrite(os.path.join(TEX, name), np.asarray(img)[:, :, ::-1],
                        [cv2.IMWRITE_JPEG_QUALITY, 88, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_440])
        else:
            img.save(os.path.join(TEX, name), "JPEG", **opts)
        mats["tex_%02d" % k] = {"TYPE": "Diffuse", "RGB": [0.9, 0.9, 0.9], "TEXTURE_FILE": name}
    objs = [{"TYPE": "cube", "MATERIAL": "light", "TRANS": [0.0, 10.0, 0.0], "ROTAT": [0.0, 0.0, 0.0], "SCALE": [3.0, 0.3, 3.0]}]
    for k in range(len(CASES)):
        objs.append({"TYPE": "cube", "MATERIAL": "tex_%02d" % k, "TRANS": [k - 5.0, 1.0, 0.0],
                     "ROTAT": [0.0, 10.0 * k, 0.0], "SCALE": [0.8, 0.8, 0.8]})
    scene = {"Materials": mats,
             "Camera": {"RES": [64, 48], "FOVY": 45.0, "ITERATIONS": 4, "DEPTH": 4, "FILE": "jpeg_textures",
                        "EYE": [0.0, 5.0, 10.5], "LOOKAT": [0.0, 5.0, 0.0], "UP": [0.0, 1.0, 0.0]},
             "Objects": objs}
    with open(os.path.join(HERE, "jpeg_textures.json"), "w") as f:
        json.dump(scene, f, indent=1)


if __name__ == "__main__":
    main()
