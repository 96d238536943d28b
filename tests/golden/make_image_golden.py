"""Golden for the image writer: the reference's OWN saveImage path -- main.cpp:88-112's loop on
its Image class, image.cpp:22-49 compiled unmodified into oracle/_ref/ref_pathtracer -- run on a seeded float SUM
image.  Runs on the CPU (no CUDA call is made in --convert mode):
    python tests/golden/make_image_golden.py
writes tests/golden/ref_saveimage.png / .hdr; tests/test_host_cli.py feeds the same SUM image to this repo's
writer (host/image_io.cpp) and compares."""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
W, H, SAMPLES = 53, 37, 7


def sum_image():
    rng = np.random.default_rng(565)
    s = (rng.random((H, W, 3), dtype=np.float32) * np.float32(10) - np.float32(1.5)).astype(np.float32)
    s[1, 2] = [SAMPLES * 1.0, SAMPLES * 0.999999, -0.0]
    s[2, 3] = [SAMPLES * 255.5 / 255.0, SAMPLES * 0.5 / 255.0, SAMPLES * 254.99999 / 255.0]
    return s


if __name__ == "__main__":
    src = "/tmp/ref_saveimage_sum.bin"
    sum_image().tofile(src)
    base = os.path.join(HERE, "ref_saveimage")
    subprocess.run([os.path.join(REPO, "oracle", "_ref", "ref_pathtracer"), "--convert", src, str(W), str(H),
                    str(SAMPLES), base, "--hdr"], check=True)
    print("wrote", base + ".png", base + ".hdr")
