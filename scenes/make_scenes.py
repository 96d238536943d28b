"""Writes the scene files of BASELINE.json's configs in the reference's JSON scene format
(path_tracer/INSTRUCTION.md:225-300; keys actually parsed: path_tracer/src/scene.cpp:40-199).

    python scenes/make_scenes.py            # (re)writes scenes/*.json

cornell.json        config 2: the Cornell box of the reference's scenes/cornell.json (same values:
                    7 primitives, 5 materials, 800x800, depth 8, 5000 iterations)
cornell_mirror.json same box with REFLECTIVE 1.0 on the sphere (the shipped scene's "specular"
                    material is diffuse to the reference loader)
cornell_small.json  160x120 variant for fast parity tests (also *_mirror)
cornell_glass_1080p.json   config 4: 1920x1080, depth 12, mirror + glass + diffuse spheres.
                    The glass material is marked by the extension keys REFRACTIVE / IOR which the
                    reference loader ignores: the reference baseline patches material id
                    `glass` after load (oracle/ref_headless_main.cpp --refract).
synthetic_4096.json config 5: 4096 random spheres/cubes, 64 materials, 3840x2160, depth 8.
"""
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))


def cube(mat, t, r, s):
    return {"TYPE": "cube", "MATERIAL": mat, "TRANS": t, "ROTAT": r, "SCALE": s}


def sphere(mat, t, r, s):
    return {"TYPE": "sphere", "MATERIAL": mat, "TRANS": t, "ROTAT": r, "SCALE": s}


def cornell_materials(sphere_mat):
    m = {
        "light": {"TYPE": "Emitting", "RGB": [1.0, 1.0, 1.0], "EMITTANCE": 5.0},
        "diffuse_white": {"TYPE": "Diffuse", "RGB": [0.98, 0.98, 0.98]},
        "diffuse_red": {"TYPE": "Diffuse", "RGB": [0.85, 0.35, 0.35]},
        "diffuse_green": {"TYPE": "Diffuse", "RGB": [0.35, 0.85, 0.35]},
    }
    m["specular_white"] = sphere_mat
    return m


def cornell_box():
    return [
        cube("light", [0.0, 10.0, 0.0], [0.0, 0.0, 0.0], [3.0, 0.3, 3.0]),
        cube("diffuse_white", [0.0, 0.0, 0.0], [0.0, 0.0, 0.0], [10.0, 0.01, 10.0]),
        cube("diffuse_white", [0.0, 10.0, 0.0], [0.0, 0.0, 90.0], [0.01, 10.0, 10.0]),
        cube("diffuse_white", [0.0, 5.0, -5.0], [0.0, 90.0, 0.0], [0.01, 10.0, 10.0]),
        cube("diffuse_red", [-5.0, 5.0, 0.0], [0.0, 0.0, 0.0], [0.01, 10.0, 10.0]),
        cube("diffuse_green", [5.0, 5.0, 0.0], [0.0, 0.0, 0.0], [0.01, 10.0, 10.0]),
    ]


def camera(res, iters, depth, name, eye=(0.0, 5.0, 10.5), lookat=(0.0, 5.0, 0.0)):
    return {"RES": list(res), "FOVY": 45.0, "ITERATIONS": iters, "DEPTH": depth, "FILE": name,
            "EYE": list(eye), "LOOKAT": list(lookat), "UP": [0.0, 1.0, 0.0]}


def cornell(res, iters, depth, name, mirror):
    mat = {"TYPE": "Specular", "RGB": [0.98, 0.98, 0.98], "ROUGHNESS": 0.0}
    if mirror:
        mat["REFLECTIVE"] = 1.0
    return {"Materials": cornell_materials(mat), "Camera": camera(res, iters, depth, name),
            "Objects": cornell_box() + [sphere("specular_white", [-1.0, 4.0, -1.0], [0.0, 0.0, 0.0], [3.0, 3.0, 3.0])]}


def cornell_glass(res, iters, depth, name):
    mats = cornell_materials({"TYPE": "Specular", "RGB": [0.98, 0.98, 0.98], "REFLECTIVE": 1.0})
    mats["glass"] = {"TYPE": "Refractive", "RGB": [0.95, 0.97, 0.99], "SPECRGB": [0.98, 0.98, 0.98],
                     "REFRACTIVE": 1.0, "IOR": 1.5}
    mats["diffuse_blue"] = {"TYPE": "Diffuse", "RGB": [0.35, 0.45, 0.85]}
    objs = cornell_box() + [
        sphere("specular_white", [-2.6, 2.0, -1.5], [0.0, 0.0, 0.0], [3.2, 3.2, 3.2]),
        sphere("glass", [0.4, 3.2, 1.2], [0.0, 0.0, 0.0], [2.6, 2.6, 2.6]),
        sphere("diffuse_blue", [2.8, 1.5, -1.0], [0.0, 0.0, 0.0], [3.0, 3.0, 3.0]),
    ]
    return {"Materials": mats, "Camera": camera(res, iters, depth, name), "Objects": objs}


class SplitMix64:
    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def uniform(self, lo, hi):
        return round(lo + (hi - lo) * (self.next() >> 11) / float(1 << 53), 4)


def synthetic(n_prims, res, iters, depth, name, seed=565):
    """random spheres/cubes, 64 materials (8 emissive, 16 mirror,
    8 glass, 32 diffuse), floor + back wall."""
    rng = SplitMix64(seed)
    mats = {}
    for i in range(8):
        mats["m%02d_light" % i] = {"RGB": [1.0, 1.0, 1.0], "EMITTANCE": rng.uniform(4.0, 8.0)}
    for i in range(8, 24):
        c = [rng.uniform(0.2, 0.95) for _ in range(3)]
        mats["m%02d_mirror" % i] = {"RGB": c, "SPECRGB": c, "REFLECTIVE": 1.0}
    for i in range(24, 32):
        c = [rng.uniform(0.8, 0.99) for _ in range(3)]
        mats["m%02d_glass" % i] = {"RGB": c, "SPECRGB": c, "REFRACTIVE": 1.0, "IOR": 1.5}
    for i in range(32, 64):
        mats["m%02d_diffuse" % i] = {"RGB": [rng.uniform(0.2, 0.95) for _ in range(3)]}
    names = sorted(mats)
    objs = [cube("m40_diffuse", [0.0, -0.5, -20.0], [0.0, 0.0, 0.0], [80.0, 1.0, 80.0]),
            cube("m41_diffuse", [0.0, 20.0, -42.0], [0.0, 0.0, 0.0], [80.0, 50.0, 1.0])]
    while len(objs) < n_prims:
        is_sphere = rng.next() & 1
        t = [rng.uniform(-24.0, 24.0), rng.uniform(0.5, 20.0), rng.uniform(-40.0, 0.0)]
        m = names[rng.next() % 64]
        if is_sphere:
            s = rng.uniform(0.3, 1.2)
            objs.append(sphere(m, t, [0.0, 0.0, 0.0], [s, s, s]))
        else:
            objs.append(cube(m, t, [rng.uniform(0.0, 360.0) for _ in range(3)],
                             [rng.uniform(0.3, 1.2) for _ in range(3)]))
    cam = camera(res, iters, depth, name, eye=(0.0, 8.0, 14.0), lookat=(0.0, 8.0, 0.0))
    return {"Materials": mats, "Camera": cam, "Objects": objs}


def write(name, scene):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(scene, f, indent=1 if len(scene["Objects"]) < 100 else None)
        f.write("\n")


def main():
    write("cornell.json", cornell((800, 800), 5000, 8, "cornell", False))
    write("cornell_mirror.json", cornell((800, 800), 5000, 8, "cornell_mirror", True))
    write("cornell_small.json", cornell((160, 120), 64, 8, "cornell_small", False))
    write("cornell_small_mirror.json", cornell((160, 120), 64, 8, "cornell_small_mirror", True))
    write("cornell_tiny.json", cornell((64, 48), 16, 8, "cornell_tiny", True))
    write("cornell_glass_1080p.json", cornell_glass((1920, 1080), 2000, 12, "cornell_glass_1080p"))
    write("cornell_glass_small.json", cornell_glass((192, 108), 64, 12, "cornell_glass_small"))
    write("synthetic_4096.json", synthetic(4096, (3840, 2160), 256, 8, "synthetic_4096"))
    write("synthetic_600_small.json", synthetic(600, (128, 72), 16, 8, "synthetic_600_small"))


if __name__ == "__main__":
    main()
