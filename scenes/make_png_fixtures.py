"""Small PNG textures + the scene that names them (loader parity fixtures for cuda_pathtracer_b200/host/png.cpp).

    python scenes/make_png_fixtures.py          # writes scenes/Textures/p_*.png, scenes/png_textures.json

The files are committed; this script documents how they were made (Pillow).  They cover every output convention of
stb_image 2.06, the reference's texture reader: RGB / RGBA / grey / grey+alpha at 8 bits, palette with and without
tRNS, 1/2/4-bit grey (scaled) and palette (not scaled), colour-key transparency on grey and RGB, Adam7 interlacing
(also on widths below 8 and with sub-byte depths), stored / fixed / dynamic deflate blocks (compress_level 0, 1, 9).
tests/golden/png_textures.scene.bin is the dump of the reference's own loader (tests/golden/make_pt_golden.py)."""
import json
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
TEX = os.path.join(HERE, "Textures")


def picture(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.stack([127 + 120 * np.sin(x / 4.0 + seed), 127 + 120 * np.cos(y / 3.0), (x * 9 + y * 5) % 256], -1)
    img[h // 3: h // 2, w // 4: w // 2] = (255, 0, 32)
    img += rng.normal(0, 20, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def write_png(name, samples, depth, color, palette, trns, grey_key, interlace, rng):
    """samples: H x W x C array of sample values (< 2^depth); rows get filter None or Sub at random; three IDAT chunks"""
    import struct
    import zlib

    def chunk(t, body):
        return struct.pack(">I", len(body)) + t + body + struct.pack(">I", zlib.crc32(t + body) & 0xFFFFFFFF)

    def pack_rows(sub):
        hh, ww, cc = sub.shape
        out = b""
        for y in range(hh):
            vals = sub[y].reshape(-1).astype(np.uint32)
            if depth == 8:
                row = bytes(vals.astype(np.uint8))
                bpp = cc
            else:
                per = 8 // depth
                pad = (-len(vals)) % per
                v = np.concatenate([vals, np.zeros(pad, np.uint32)]).reshape(-1, per)
                row = bytes(sum(v[:, k] << (8 - depth * (k + 1)) for k in range(per)).astype(np.uint8))
                bpp = 1
            if rng.integers(0, 2):
                out += b"\x01" + bytes((row[i] - (row[i - bpp] if i >= bpp else 0)) & 255 for i in range(len(row)))
            else:
                out += b"\x00" + row
        return out

    h, w, c = samples.shape
    if interlace:
        raw = b""
        for xo, yo, xs, ys in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = samples[yo::ys, xo::xs]
            if sub.shape[0] and sub.shape[1]:
                raw += pack_rows(sub)
    else:
        raw = pack_rows(samples)
    z = zlib.compress(raw, 6)
    cut = [len(z) // 3, 2 * len(z) // 3]
    body = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 1 if interlace else 0))
    body += chunk(b"gAMA", struct.pack(">I", 45455))                      # an ancillary chunk to skip
    if palette is not None:
        body += chunk(b"PLTE", bytes(palette.reshape(-1)))
    if trns is not None:
        body += chunk(b"tRNS", trns)
    if grey_key is not None:
        body += chunk(b"tRNS", struct.pack(">H", grey_key))
    for part in (z[:cut[0]], z[cut[0]:cut[1]], z[cut[1]:]):
        body += chunk(b"IDAT", part)
    body += chunk(b"IEND", b"")
    with open(os.path.join(TEX, name), "wb") as f:
        f.write(body)
    return name


def main():
    os.makedirs(TEX, exist_ok=True)
    cases = []

    def save(name, img, **opts):
        img.save(os.path.join(TEX, name), "PNG", **opts)
        cases.append(name)

    rgb = Image.fromarray(picture(37, 21, 1))
    save("p_rgb8.png", rgb, compress_level=9)
    save("p_rgb8_stored.png", rgb, compress_level=0)
    save("p_rgb8_fast.png", Image.fromarray(picture(64, 5, 2)), compress_level=1)
    rgba = np.dstack([picture(19, 23, 4), (np.arange(19 * 23).reshape(23, 19) * 3 % 256).astype(np.uint8)])
    save("p_rgba8.png", Image.fromarray(rgba, "RGBA"))
    save("p_l8.png", Image.fromarray(picture(30, 14, 5)).convert("L"))
    save("p_la8.png", Image.fromarray(rgba, "RGBA").convert("LA"))
    pal = Image.fromarray(picture(41, 27, 6)).quantize(colors=200)
    save("p_pal8.png", pal)
    palt = Image.fromarray(picture(17, 17, 10)).quantize(colors=32)
    save("p_pal8_trns.png", palt, transparency=bytes([(i * 37) % 256 for i in range(20)]))
    grey = Image.fromarray(picture(29, 10, 11)).convert("L")
    save("p_l8_key.png", grey, transparency=int(np.asarray(grey)[3, 3]))
    rgbk = Image.fromarray(picture(16, 16, 12))
    save("p_rgb8_key.png", rgbk, transparency=tuple(int(v) for v in np.asarray(rgbk)[5, 5]))
    save("p_1wide.png", Image.fromarray(picture(1, 9, 13)))
    # hand-written files: sub-byte depths and Adam7, rows filtered with None / Sub only.  (stb_image 2.06 takes the
    # PRIOR row of a sub-byte image from the wrong place -- `prior` is computed before `cur` is moved to the packed
    # bytes, stb_image.h:3982-3991 -- so Up / Average / Paeth rows of such files decode to whatever the heap held:
    # the reference is undefined there, this repo follows the PNG specification.)
    rng = np.random.default_rng(99)
    for depth in (1, 2, 4):
        for color, tag in ((0, "l"), (3, "pal")):
            w, h = 29 + depth, 7 + depth
            idx = rng.integers(0, 1 << depth, (h, w), dtype=np.uint8)
            pal = rng.integers(0, 256, (1 << depth, 3), dtype=np.uint8) if color == 3 else None
            trns = bytes(rng.integers(0, 256, 1 << depth, dtype=np.uint8)) if (color == 3 and depth == 2) else None
            key = 1 if (color == 0 and depth == 4) else None
            cases.append(write_png("p_%s%d.png" % (tag, depth), idx[:, :, None], depth, color, pal, trns, key, False, rng))
            cases.append(write_png("p_%s%d_adam7.png" % (tag, depth), idx[:, :, None], depth, color, pal, trns, key, True, rng))
    cases.append(write_png("p_rgb8_adam7.png", picture(45, 33, 3), 8, 2, None, None, None, True, rng))
    cases.append(write_png("p_rgba8_adam7_small.png", np.dstack([picture(5, 3, 4), np.full((3, 5), 77, np.uint8)]), 8, 6, None, None, None, True, rng))
    mats = {"light": {"TYPE": "Emitting", "RGB": [1.0, 1.0, 1.0], "EMITTANCE": 5.0}}
    objs = [{"TYPE": "cube", "MATERIAL": "light", "TRANS": [0.0, 10.0, 0.0], "ROTAT": [0.0, 0.0, 0.0], "SCALE": [3.0, 0.3, 3.0]}]
    for k, name in enumerate(cases):
        mats["tex_%02d" % k] = {"TYPE": "Diffuse", "RGB": [0.9, 0.9, 0.9], "TEXTURE_FILE": name}
        objs.append({"TYPE": "cube", "MATERIAL": "tex_%02d" % k, "TRANS": [k - 8.0, 1.0, 0.0], "ROTAT": [0.0, 5.0 * k, 0.0],
                     "SCALE": [0.8, 0.8, 0.8]})
    scene = {"Materials": mats,
             "Camera": {"RES": [64, 48], "FOVY": 45.0, "ITERATIONS": 4, "DEPTH": 4, "FILE": "png_textures",
                        "EYE": [0.0, 5.0, 10.5], "LOOKAT": [0.0, 5.0, 0.0], "UP": [0.0, 1.0, 0.0]},
             "Objects": objs}
    with open(os.path.join(HERE, "png_textures.json"), "w") as f:
        json.dump(scene, f, indent=1)
    print(len(cases), "PNG fixtures")


if __name__ == "__main__":
    main()
