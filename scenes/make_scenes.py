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
synthetic_4096_small.json the same scene at 480x270 (parity fixture of config 5).
synthetic_4096.json config 5: 4096 random spheres/cubes, 64 materials, 3840x2160, depth 8.
mesh_room*.json     triangle meshes + textures in the layout of the reference's scenes/room.json
                    (cube walls, "mesh" objects with OBJ_FILE under Models/, materials with
                    TEXTURE_FILE under Textures/).  The OBJ models (grid quad, icosphere, torus) and the
                    binary PPM textures are generated here as well.
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


# ---- triangle meshes and textures (the room.json analogue) -----------------------------------------
import math


def fmt(x):
    return "%.6f" % x


def write_obj(name, verts, uvs, normals, faces):
    """faces: tuples of (v, vt, vn) 0-based index triples, counter-clockwise seen from outside"""
    os.makedirs(os.path.join(HERE, "Models"), exist_ok=True)
    with open(os.path.join(HERE, "Models", name), "w") as f:
        f.write("# generated by scenes/make_scenes.py\n")
        for v in verts:
            f.write("v " + " ".join(fmt(c) for c in v) + "\n")
        for t in uvs:
            f.write("vt " + " ".join(fmt(c) for c in t) + "\n")
        for n in normals:
            f.write("vn " + " ".join(fmt(c) for c in n) + "\n")
        for face in faces:
            f.write("f " + " ".join("%d/%d/%d" % (a + 1, b + 1, c + 1) for a, b, c in face) + "\n")


def grid_quad(n):
    """unit square in the xy plane facing +z, n x n cells, QUAD faces (the loader fans them)"""
    verts, uvs = [], []
    for j in range(n + 1):
        for i in range(n + 1):
            verts.append((i / n - 0.5, j / n - 0.5, 0.0))
            uvs.append((i / n, j / n))
    faces = []
    for j in range(n):
        for i in range(n):
            a, b, c, d = j * (n + 1) + i, j * (n + 1) + i + 1, (j + 1) * (n + 1) + i + 1, (j + 1) * (n + 1) + i
            faces.append([(a, a, 0), (b, b, 0), (c, c, 0), (d, d, 0)])
    return verts, uvs, [(0.0, 0.0, 1.0)], faces


def icosphere(levels):
    t = (1.0 + math.sqrt(5.0)) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
         (8, 6, 7), (9, 8, 1)]

    def norm(p):
        l = math.sqrt(sum(c * c for c in p))
        return tuple(c / l for c in p)

    v = [norm(p) for p in v]
    for _ in range(levels):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                cache[key] = len(v)
                v.append(norm(tuple((v[a][k] + v[b][k]) / 2 for k in range(3))))
            return cache[key]

        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    verts = [tuple(0.5 * c for c in p) for p in v]
    uvs = [(0.5 + math.atan2(p[2], p[0]) / (2 * math.pi), 0.5 - math.asin(max(-1.0, min(1.0, p[1]))) / math.pi) for p in v]
    faces = [[(a, a, a), (b, b, b), (c, c, c)] for a, b, c in f]
    return verts, uvs, v, faces


def torus(nu, nv, R=0.35, r=0.15):
    verts, uvs, normals = [], [], []
    for i in range(nu):
        for j in range(nv):
            u, w = 2 * math.pi * i / nu, 2 * math.pi * j / nv
            verts.append(((R + r * math.cos(w)) * math.cos(u), r * math.sin(w), (R + r * math.cos(w)) * math.sin(u)))
            normals.append((math.cos(w) * math.cos(u), math.sin(w), math.cos(w) * math.sin(u)))
            uvs.append((i / nu, j / nv))
    faces = []
    for i in range(nu):
        for j in range(nv):
            a, b = i * nv + j, ((i + 1) % nu) * nv + j
            c, d = ((i + 1) % nu) * nv + (j + 1) % nv, i * nv + (j + 1) % nv
            faces.append([(a, a, a), (d, d, d), (c, c, c)])
            faces.append([(a, a, a), (c, c, c), (b, b, b)])
    return verts, uvs, normals, faces


def write_ppm(name, w, h, pixel):
    os.makedirs(os.path.join(HERE, "Textures"), exist_ok=True)
    with open(os.path.join(HERE, "Textures", name), "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(bytes(c for y in range(h) for x in range(w) for c in pixel(x, y)))


def mesh_assets():
    write_obj("grid_quad.obj", *grid_quad(4))
    write_obj("icosphere.obj", *icosphere(2))
    write_obj("torus.obj", *torus(16, 12))
    write_ppm("checker.ppm", 64, 64,
              lambda x, y: (230, 230, 60) if ((x // 8) + (y // 8)) % 2 else (40, 60, 200))
    write_ppm("gradient.ppm", 48, 32, lambda x, y: (5 * x + 10, 7 * y + 20, (x * y) % 256))


def mesh_room(res, iters, depth, name, textured=True):
    mats = {
        "light": {"TYPE": "Emitting", "RGB": [1.0, 1.0, 1.0], "EMITTANCE": 12.0},
        "diffuse_white": {"TYPE": "Diffuse", "RGB": [0.98, 0.98, 0.98]},
        "diffuse_red": {"TYPE": "Diffuse", "RGB": [0.85, 0.65, 0.65], "SPECRGB": [0.98, 0.98, 0.98], "REFL": 1.0},
        "diffuse_blue": {"TYPE": "Diffuse", "RGB": [0.65, 0.65, 0.85]},
        "checker_wall": {"TYPE": "Diffuse", "RGB": [0.88, 0.88, 0.88], "TEXTURE_FILE": "checker.ppm"},
        "gradient_skin": {"TYPE": "Diffuse", "RGB": [0.98, 0.98, 0.98], "SPECRGB": [0.5, 0.58, 0.9],
                          "TEXTURE_FILE": "gradient.ppm"},
        "mirror": {"TYPE": "Specular", "RGB": [0.98, 0.98, 0.98], "REFLECTIVE": 1.0},
        "plain_green": {"TYPE": "Diffuse", "RGB": [0.35, 0.85, 0.35]},
    }

    if not textured:
        # useBVHtree = false blends texture coordinates as az*uv0 + uv1 + uv2 (intersections.cu:164): they leave
        # [0, 1] and the reference's Texture::get_color then reads outside the image (sceneStructs.h:175-188) --
        # whole renders of that mode are only comparable without textures
        for m in mats.values():
            m.pop("TEXTURE_FILE", None)

    def mesh(mat, obj, t, r, s):
        return {"TYPE": "mesh", "MATERIAL": mat, "TRANS": t, "ROTAT": r, "SCALE": s, "OBJ_FILE": obj}

    objs = [
        cube("light", [0.0, 10.0, 0.0], [0.0, 0.0, 0.0], [3.0, 0.1, 3.0]),
        cube("diffuse_white", [0.0, 0.0, 0.0], [0.0, 0.0, 0.0], [15.0, 0.01, 15.0]),
        cube("diffuse_white", [0.0, 10.0, 0.0], [0.0, 0.0, 90.0], [0.01, 15.0, 15.0]),
        cube("diffuse_white", [0.0, 5.0, -5.0], [0.0, 90.0, 0.0], [0.01, 10.0, 15.0]),
        cube("diffuse_red", [-7.0, 5.0, 0.0], [0.0, 0.0, 0.0], [0.01, 10.0, 15.0]),
        cube("diffuse_blue", [7.0, 5.0, 0.0], [0.0, 0.0, 0.0], [0.01, 10.0, 15.0]),
        mesh("checker_wall", "grid_quad.obj", [0.0, 5.0, -4.8], [0.0, 0.0, 0.0], [9.0, 7.0, 1.0]),
        mesh("gradient_skin", "icosphere.obj", [-2.5, 2.0, -1.0], [0.0, 30.0, 0.0], [3.6, 3.6, 3.6]),
        mesh("plain_green", "torus.obj", [2.6, 2.2, 0.5], [35.0, 20.0, 10.0], [5.0, 5.0, 5.0]),
        sphere("mirror", [0.3, 1.2, 2.0], [0.0, 0.0, 0.0], [2.4, 2.4, 2.4]),
        mesh("gradient_skin", "torus.obj", [-3.5, 6.5, -2.0], [80.0, 0.0, 25.0], [4.0, 4.0, 4.0]),
    ]
    return {"Materials": mats, "Camera": camera(res, iters, depth, name), "Objects": objs}


def write(name, scene):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(scene, f, indent=1 if len(scene["Objects"]) < 100 else None)
        f.write("\n")


def stress_assets():
    """OBJ files that exercise the loader rather than the renderer: polygons of 3-6 corners (fan
    triangulation), negative (relative) indices, floats with exponents / explicit signs / 17
    significant digits (tinyobjloader's own decimal parser differs from strtod in the last place),
    doubled blanks, g/s/usemtl/comment lines, clustered and grid-snapped vertices (many equal
    centroids and bucket ties for the BVH build)."""
    import random
    rnd = random.Random(7)

    def fmt(x):
        r = rnd.random()
        if r < 0.3:
            return "%.17g" % x
        if r < 0.5:
            return "%.9e" % x
        if r < 0.6:
            return "%+.4f" % x
        if r < 0.7:
            return "%.12f" % x
        return repr(float(x))

    for name, n in (("stress_cloud", 260), ("stress_grid", 140), ("stress_few", 9)):
        lines, nv = [], 0
        for i in range(n):
            cx, cy, cz = [rnd.choice([0, 0, 1, -1, .5]) * rnd.random() * 2 for _ in range(3)]
            k = rnd.choice([3, 3, 3, 4, 5, 6])
            for j in range(k):
                if name == "stress_grid":
                    lines.append("v %s %s %s" % (fmt(round(cx * 4) / 4 + (j % 2) * 0.25), fmt(round(cy * 4) / 4 + (j // 2) * 0.25),
                                                 fmt(round(cz * 4) / 4)))
                else:
                    lines.append("v %s %s %s" % (fmt(cx + rnd.uniform(-.1, .1)), fmt(cy + rnd.uniform(-.1, .1)),
                                                 fmt(cz + rnd.uniform(-.1, .1))))
                lines.append("vt %s %s" % (fmt(rnd.random()), fmt(rnd.random())))
                lines.append("vn %s %s %s" % (fmt(rnd.uniform(-1, 1)), fmt(rnd.uniform(-1, 1)), fmt(rnd.uniform(-1, 1))))
            if rnd.random() < 0.5:
                lines.append("f " + " ".join("%d/%d/%d" % ((nv + j + 1,) * 3) for j in range(k)))
            else:
                lines.append("f " + "  ".join("%d/%d/%d" % ((j - k,) * 3) for j in range(k)))
            nv += k
            if i % 50 == 0:
                lines += ["g grp%d" % i, "# comment", "s off", "usemtl foo"]
        with open(os.path.join(HERE, "Models", name + ".obj"), "w") as f:
            f.write("\n".join(lines) + "\n")


def mesh_stress():
    sc = mesh_room((96, 96), 4, 8, "mesh_stress")
    objs = [o for o in sc["Objects"] if o["TYPE"] != "mesh"]
    proto = [o for o in sc["Objects"] if o["TYPE"] == "mesh"][0]
    for nm, t, r, s_ in (("stress_cloud", [0, 5, 0], [10, 20, 30], [1.5, 1, 2]), ("stress_few", [1, 2, 3], [0, 0, 0], [1, 1, 1]),
                         ("stress_grid", [-2, 4, 1], [45, 0, 90], [2, 2, 2]), ("stress_few", [0, 0, 0], [0, 90, 0], [3, .1, 3])):
        o = dict(proto)
        o.update(OBJ_FILE=nm + ".obj", TRANS=t, ROTAT=r, SCALE=s_)
        objs.append(o)
    sc["Objects"] = objs
    return sc


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
    # config 5's 4096 primitives (same seed, same scene) at a resolution the reference build and the CPU oracle finish
    # in seconds: the parity case of that configuration (tests/pt_cases.py)
    write("synthetic_4096_small.json", synthetic(4096, (480, 270), 16, 8, "synthetic_4096_small"))
    mesh_assets()
    write("mesh_room.json", mesh_room((1200, 1200), 2000, 8, "mesh_room"))
    write("mesh_room_small.json", mesh_room((160, 160), 32, 8, "mesh_room_small"))
    write("mesh_room_small_notex.json", mesh_room((160, 160), 32, 8, "mesh_room_small_notex", textured=False))
    stress_assets()
    write("mesh_stress.json", mesh_stress())


if __name__ == "__main__":
    main()
