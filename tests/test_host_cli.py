"""Host side above the C ABI: the C++ scene loader, the image writer and the headless CLI
(cuda_pathtracer_b200/host/ -> cuda_pathtracer_b200/bin/cuda_path_tracer), the drop-in for the
reference executable (`cuda_path_tracer SCENE.json`, path_tracer/src/main.cpp:34-112).

CPU tests: the loader's PODs are BYTE-IDENTICAL to dumps written by the reference's own, unmodified
scene.cpp (tests/golden/*.scene.bin, made by oracle/ref_scene_dump.cpp); the PNG writer decodes to
the pixels saveImage()/Image::savePNG define (x mirrored, clamp, *255.f, truncation).
GPU tests: the CLI's float SUM image equals the reference CUDA build's bit for bit, output naming,
restartable rendering."""
import glob
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

from pt_cases import GOLD, bits_equal, ref_sum

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "cuda_path_tracer")
SCENES = ["cornell", "cornell_mirror", "cornell_small", "cornell_small_mirror", "cornell_tiny",
          "cornell_glass_small", "cornell_glass_1080p", "synthetic_600_small", "synthetic_4096", "synthetic_4096_small",
          "mesh_room", "mesh_room_small", "mesh_room_small_notex", "mesh_stress",   # OBJ meshes + BVH + textures ("B2PTSCN2")
          "jpeg_textures",   # + six progressive ones
          "png_textures"]    # 25 PNGs (scenes/make_png_fixtures.py) decoded by host/png.cpp vs stb_image   # twelve baseline JPEGs (scenes/make_jpeg_fixtures.py) decoded by host/jpeg.cpp vs stb_image


def scene_path(name):
    return os.path.join(REPO, "scenes", name + ".json")


def decode_png(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    at, idat, w, h = 8, b"", None, None
    while at < len(raw):
        n, kind = struct.unpack(">I4s", raw[at:at + 8])
        body = raw[at + 8:at + 8 + n]
        crc, = struct.unpack(">I", raw[at + 8 + n:at + 12 + n])
        assert crc == zlib.crc32(kind + body) & 0xFFFFFFFF
        if kind == b"IHDR":
            w, h, depth, ctype, comp, filt, inter = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, filt, inter) == (8, 2, 0, 0, 0)
        elif kind == b"IDAT":
            idat += body
        at += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert not rows[:, 0].any()                              # filter type 0 on every row
    return rows[:, 1:].reshape(h, w, 3)


def expected_png_pixels(sum_img, samples):
    """main.cpp:88-101 + image.cpp:22-36: mean, x mirror, clamp to [0,1], * 255.f, truncate"""
    mean = sum_img.astype(np.float32) / np.float32(samples)
    v = np.clip(mean, np.float32(0), np.float32(1)) * np.float32(255)
    v = np.where(np.isnan(v), np.float32(0), v)
    return v.astype(np.uint8)[:, ::-1, :]


@pytest.mark.parametrize("name", SCENES)
def test_loader_is_byte_identical_to_the_reference_loader(tmp_path, name):
    out = str(tmp_path / "scene.bin")
    subprocess.run([EXE, scene_path(name), "--dump-scene", out], check=True, stdout=subprocess.DEVNULL)
    assert open(out, "rb").read() == open(os.path.join(GOLD, name + ".scene.bin"), "rb").read()


def test_loader_against_the_reference_tree_when_present(tmp_path):
    """the reference's own shipped cornell.json through both loaders (build container only)"""
    ref_scene = "/root/reference/path_tracer/scenes/cornell.json"
    dump = os.path.join(REPO, "oracle", "_ref", "ref_scene_dump")
    if not (os.path.exists(ref_scene) and os.path.exists(dump)):
        pytest.skip("needs /root/reference and oracle/_ref/ref_scene_dump")
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    subprocess.run([dump, ref_scene, a], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([EXE, ref_scene, "--dump-scene", b], check=True, stdout=subprocess.DEVNULL)
    assert open(a, "rb").read() == open(b, "rb").read()


REF_ROOM = os.path.join(REPO, "oracle", "_ref", "ref_scenes", "room.json")     # staged by oracle/Makefile


def test_reference_room_scene_loads_unmodified(tmp_path):
    """the reference's own scenes/room.json (4 OBJ meshes, chair.jpg 1024x1024 + wallpaper.jpg 500x250,
    both baseline 4:2:0 JPEG) through both loaders: every byte equal, texels included"""
    dump = os.path.join(REPO, "oracle", "_ref", "ref_scene_dump")
    if not (os.path.exists(REF_ROOM) and os.path.exists(dump)):
        pytest.skip("needs oracle/_ref (reference scene files + ref_scene_dump, build container)")
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    subprocess.run([dump, REF_ROOM, a], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([EXE, REF_ROOM, "--dump-scene", b], check=True, stdout=subprocess.DEVNULL)
    raw = open(a, "rb").read()
    assert len(raw) > 3 * 1024 * 1024 and raw == open(b, "rb").read()


def test_texture_decoders_reject_what_they_do_not_decode(tmp_path):
    """arithmetic-coded JPEG, 16-bit PNG (stb_image 2.06 refuses both too), truncated data: a load error naming the
    file (exit 255 as for any load error), never garbage texels; a progressive JPEG loads"""
    from PIL import Image
    tex = tmp_path / "Textures"
    tex.mkdir()
    img = Image.fromarray((np.arange(48 * 40 * 3) % 251).astype(np.uint8).reshape(40, 48, 3))
    img.save(str(tex / "prog.jpg"), "JPEG", progressive=True)
    img.save(str(tex / "ok.jpg"), "JPEG")
    whole = (tex / "ok.jpg").read_bytes()
    (tex / "cut.jpg").write_bytes(whole[:len(whole) // 3])
    sof = whole.index(b"\xff\xc0")
    (tex / "arith.jpg").write_bytes(whole[:sof] + b"\xff\xc9" + whole[sof + 2:])      # SOF9: arithmetic coding
    Image.fromarray((np.arange(12 * 10 * 3) * 97 % 65536).astype(np.uint16).reshape(10, 12 * 3)[:, :12]).save(str(tex / "deep.png"))   # 16-bit grey
    img.save(str(tex / "ok.png"), "PNG")
    whole_png = (tex / "ok.png").read_bytes()
    (tex / "cut.png").write_bytes(whole_png[:len(whole_png) // 2])
    for name, code in (("ok.jpg", 0), ("prog.jpg", 0), ("cut.jpg", 255), ("arith.jpg", 255), ("ok.png", 0), ("deep.png", 255),
                       ("cut.png", 255)):
        scene = {"Materials": {"m": {"TYPE": "Diffuse", "RGB": [1, 1, 1], "TEXTURE_FILE": name}},
                 "Camera": {"RES": [8, 8], "FOVY": 45.0, "ITERATIONS": 1, "DEPTH": 1, "FILE": "t", "EYE": [0, 0, 5],
                            "LOOKAT": [0, 0, 0], "UP": [0, 1, 0]},
                 "Objects": [{"TYPE": "cube", "MATERIAL": "m", "TRANS": [0, 0, 0], "ROTAT": [0, 0, 0], "SCALE": [1, 1, 1]}]}
        import json
        (tmp_path / "s.json").write_text(json.dumps(scene))
        r = subprocess.run([EXE, str(tmp_path / "s.json"), "--dump-scene", str(tmp_path / "o.bin")], capture_output=True, text=True)
        assert r.returncode == code, (name, r.stdout, r.stderr)
        if code:
            assert name in (r.stdout + r.stderr)


def test_texture_decoders_survive_corrupt_files(tmp_path):
    """120 mutated copies of the JPEG / PNG fixtures (flipped bytes, truncation, insertions): the loader either decodes
    or reports a load error (exit 255) -- no crash, no hang.  (The same decoders ran 1500 mutations clean under
    AddressSanitizer + UBSan when they were written.)"""
    import json
    import random
    rnd = random.Random(7)
    files = sorted(glob.glob(os.path.join(REPO, "scenes", "Textures", "p_*.png")) +
                   glob.glob(os.path.join(REPO, "scenes", "Textures", "j_*.jpg")))
    assert len(files) > 30
    (tmp_path / "Textures").mkdir()
    for k in range(120):
        src = rnd.choice(files)
        raw = bytearray(open(src, "rb").read())
        mode = rnd.randrange(3)
        if mode == 0:
            for _ in range(rnd.randrange(1, 6)):
                raw[rnd.randrange(len(raw))] = rnd.randrange(256)
        elif mode == 1:
            raw = raw[:rnd.randrange(1, len(raw))]
        else:
            at = rnd.randrange(len(raw))
            raw[at:at] = bytes(rnd.randrange(256) for _ in range(rnd.randrange(1, 9)))
        name = "t" + os.path.splitext(src)[1]
        (tmp_path / "Textures" / name).write_bytes(bytes(raw))
        scene = {"Materials": {"m": {"TYPE": "Diffuse", "RGB": [1, 1, 1], "TEXTURE_FILE": name}},
                 "Camera": {"RES": [8, 8], "FOVY": 45.0, "ITERATIONS": 1, "DEPTH": 1, "FILE": "t", "EYE": [0, 0, 5],
                            "LOOKAT": [0, 0, 0], "UP": [0, 1, 0]},
                 "Objects": [{"TYPE": "cube", "MATERIAL": "m", "TRANS": [0, 0, 0], "ROTAT": [0, 0, 0], "SCALE": [1, 1, 1]}]}
        (tmp_path / "s.json").write_text(json.dumps(scene))
        r = subprocess.run([EXE, str(tmp_path / "s.json"), "--dump-scene", str(tmp_path / "o.bin")], capture_output=True,
                           timeout=30)
        assert r.returncode in (0, 255), (os.path.basename(src), mode, k, r.returncode)


def test_material_ids_follow_alphabetical_key_order(tmp_path):
    """nlohmann's object is a std::map, so ids are NOT in file order"""
    out = str(tmp_path / "s.bin")
    subprocess.run([EXE, scene_path("cornell_tiny"), "--dump-scene", out], check=True, stdout=subprocess.DEVNULL)
    from cuda_pathtracer_b200 import pathtracer as pt
    sc = pt.Scene.from_pod_file(out)
    # cornell: diffuse_green 0, diffuse_red 1, diffuse_white 2, light 3, specular_white 4
    assert [round(m.emittance) for m in sc.materials] == [0, 0, 0, 5, 0]
    assert tuple(round(c, 2) for c in sc.materials[0].color) == (0.35, 0.85, 0.35)


def test_cli_exit_codes(tmp_path):
    assert subprocess.run([EXE], capture_output=True).returncode == 1                 # usage (main.cpp:38-42)
    assert subprocess.run([EXE, str(tmp_path / "missing.json")], capture_output=True).returncode == 255
    assert subprocess.run([EXE, str(tmp_path / "scene.txt")], capture_output=True).returncode == 255
    bad = tmp_path / "bad.json"
    bad.write_text('{"Materials": {}, "Objects": [{"TYPE": "cube"}]}')
    assert subprocess.run([EXE, str(bad)], capture_output=True).returncode == 255
    assert subprocess.run([EXE, scene_path("cornell_tiny"), "--bogus"], capture_output=True).returncode == 1


def test_png_writer_pixels(tmp_path):
    rng = np.random.default_rng(565)
    h, w, samples = 37, 53, 7
    s = (rng.random((h, w, 3), dtype=np.float32) * np.float32(10) - np.float32(1.5)).astype(np.float32)
    s[0, 0] = [np.nan, 7.0, 0.0]
    s[1, 2] = [samples * 1.0, samples * 0.999999, -0.0]
    src = str(tmp_path / "sum.bin")
    s.tofile(src)
    base = str(tmp_path / "img")
    subprocess.run([EXE, "--convert", src, str(w), str(h), str(samples), base, "--hdr"], check=True,
                   stdout=subprocess.DEVNULL)
    assert np.array_equal(decode_png(base + ".png"), expected_png_pixels(s, samples))
    hdr = open(base + ".hdr", "rb").read()
    assert hdr.startswith(b"#?RADIANCE\n") and hdr.endswith(hdr[-4 * w * h:]) and len(hdr) > 4 * w * h


def decode_hdr(path):
    """Radiance RGBE, flat or new-style run-length scanlines -> (h, w, 4) bytes (r, g, b, e)"""
    raw = open(path, "rb").read()
    assert raw.startswith(b"#?RADIANCE\n")
    head_end = raw.index(b"\n\n") + 2
    assert b"FORMAT=32-bit_rle_rgbe" in raw[:head_end]
    line_end = raw.index(b"\n", head_end)
    dims = raw[head_end:line_end].split()
    assert dims[0] == b"-Y" and dims[2] == b"+X"
    h, w = int(dims[1]), int(dims[3])
    at, out = line_end + 1, np.zeros((h, w, 4), dtype=np.uint8)
    for y in range(h):
        if 8 <= w < 32768 and raw[at] == 2 and raw[at + 1] == 2 and (raw[at + 2] << 8 | raw[at + 3]) == w:
            at += 4
            for c in range(4):
                x = 0
                while x < w:
                    n = raw[at]; at += 1
                    if n > 128:
                        out[y, x:x + n - 128, c] = raw[at]; at += 1; x += n - 128
                    else:
                        out[y, x:x + n, c] = np.frombuffer(raw, dtype=np.uint8, count=n, offset=at); at += n; x += n
        else:
            out[y] = np.frombuffer(raw, dtype=np.uint8, count=4 * w, offset=at).reshape(w, 4); at += 4 * w
    assert at == len(raw)
    return out


def test_image_writer_against_the_reference_saveimage(tmp_path):
    """N2 pinned to the reference itself: tests/golden/ref_saveimage.{png,hdr} were written by the reference's OWN
    saveImage loop + Image::savePNG / saveHDR (main.cpp:88-112, image.cpp:22-49, stb_image_write) from the seeded SUM
    image of tests/golden/make_image_golden.py; this repo's writer must decode to the same pixels (the files' own
    bytes differ: other deflate / filter / run-length choices, same content).  When the reference build is present
    (oracle/_ref) it is also run live."""
    import sys
    from PIL import Image
    sys.path.insert(0, GOLD)
    import make_image_golden as g
    src, base = str(tmp_path / "sum.bin"), str(tmp_path / "mine")
    g.sum_image().tofile(src)
    subprocess.run([EXE, "--convert", src, str(g.W), str(g.H), str(g.SAMPLES), base, "--hdr"], check=True,
                   stdout=subprocess.DEVNULL)
    mine_png = decode_png(base + ".png")
    assert np.array_equal(mine_png, np.array(Image.open(os.path.join(GOLD, "ref_saveimage.png"))))
    assert np.array_equal(decode_hdr(base + ".hdr"), decode_hdr(os.path.join(GOLD, "ref_saveimage.hdr")))
    ref_exe = os.path.join(REPO, "oracle", "_ref", "ref_pathtracer")
    if os.path.exists(ref_exe):
        live = str(tmp_path / "ref")
        subprocess.run([ref_exe, "--convert", src, str(g.W), str(g.H), str(g.SAMPLES), live, "--hdr"], check=True,
                       stdout=subprocess.DEVNULL)
        assert np.array_equal(mine_png, np.array(Image.open(live + ".png")))
        assert np.array_equal(decode_hdr(base + ".hdr"), decode_hdr(live + ".hdr"))


# ---- on a B200 ---------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("scene,iters,flags,tag", [("cornell_small", 16, [], "default"),
                                                    ("cornell_glass_small", 16, ["--sort"], "sort"),
                                                    # useThrustPartition: golden from the reference build run with it
                                                    ("cornell_small", 16, ["--thrust-partition"], "thrustpart"),
                                                    # OBJ + BVH + textures through the C++ loader
                                                    ("mesh_room_small", 8, [], "default"),
                                                    ("mesh_room_small", 8, ["--sort", "--no-ssaa", "--no-dof"], "sortplain")])
def test_cli_render_matches_reference_cuda_build(tmp_path, scene, iters, flags, tag):
    sum_out, base = str(tmp_path / "sum.bin"), str(tmp_path / "out")
    subprocess.run([EXE, scene_path(scene), "--iters", str(iters), "--batch", "1", "--sum-out", sum_out,
                    "--out", base] + flags, check=True, stdout=subprocess.DEVNULL)
    got = np.fromfile(sum_out, dtype=np.float32)
    png = decode_png(base + ".png")
    h, w = png.shape[:2]
    got = got.reshape(h, w, 3)
    assert bits_equal(got, ref_sum(scene, iters, tag, got.shape))
    assert np.array_equal(png, expected_png_pixels(got, iters))


@pytest.mark.gpu
def test_reference_room_scene_renders_bit_identical_to_the_reference_cuda_build(tmp_path):
    """The reference's scenes/room.json, unmodified (1200x1200, OBJ meshes + BVH + JPEG textures), through
    this repo's CLI and through the reference's own CUDA build on this box: the float SUM images are equal
    bit for bit.  Needs oracle/_ref (travels with the snapshot from the build container)."""
    ref_exe = os.path.join(REPO, "oracle", "_ref", "ref_pathtracer")
    if not (os.path.exists(REF_ROOM) and os.path.exists(ref_exe)):
        pytest.skip("needs oracle/_ref (reference scene files + the reference CUDA build)")
    want, got = str(tmp_path / "ref.bin"), str(tmp_path / "got.bin")
    subprocess.run([ref_exe, REF_ROOM, "--iters", "3", "--out", want], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([EXE, REF_ROOM, "--iters", "3", "--batch", "1", "--no-png", "--sum-out", got], check=True,
                   stdout=subprocess.DEVNULL)
    a, b = np.fromfile(want, dtype=np.float32), np.fromfile(got, dtype=np.float32)
    assert a.size == 1200 * 1200 * 3 and bits_equal(a, b)
    assert np.count_nonzero(a) > 100000                 # a real (if dark: few paths reach the lamp in 3 samples) picture


@pytest.mark.gpu
def test_cli_default_output_name_and_iterations(tmp_path):
    """no flags: renders Camera.ITERATIONS samples and writes <FILE>.<time>.<N>samp.png into cwd"""
    r = subprocess.run([EXE, scene_path("cornell_tiny")], cwd=str(tmp_path), capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    files = glob.glob(str(tmp_path / "*.png"))
    assert len(files) == 1
    name = os.path.basename(files[0])
    import json
    cam = json.load(open(scene_path("cornell_tiny")))["Camera"]
    parts = name.split(".")
    assert parts[0] == cam["FILE"] and parts[-2] == f"{cam['ITERATIONS']}samp" and parts[1].endswith("z")
    assert "Reading scene from" in r.stdout and "Saved" in r.stdout


@pytest.mark.gpu
def test_cli_restartable_rendering(tmp_path):
    """N4: 6 iterations == 4 iterations, stop, resume with 2 more (batch 1: bit-identical)"""
    a, b, c = (str(tmp_path / n) for n in ("a.bin", "b.bin", "c.bin"))
    common = [EXE, scene_path("cornell_tiny"), "--batch", "1", "--no-png"]
    subprocess.run(common + ["--iters", "6", "--sum-out", a], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(common + ["--iters", "4", "--sum-out", b], check=True, stdout=subprocess.DEVNULL)
    subprocess.run(common + ["--iters", "2", "--resume-sum", b, "--samples-done", "4", "--sum-out", c], check=True,
                   stdout=subprocess.DEVNULL)
    assert bits_equal(np.fromfile(a, dtype=np.float32), np.fromfile(c, dtype=np.float32))


# ---- scan / compact harness (include/b200_stream_compaction.hpp + host/sc_harness.cpp) ------------------
SC_EXE = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "stream_compaction_test")


def test_sc_harness_cpu_namespace_config0():
    """BASELINE configs[0]: StreamCompaction::CPU scan + compact on n = 2^20 (POT and NPOT) on the host"""
    import json
    r = subprocess.run([SC_EXE, "20", "--cpu-only", "--json"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "FAIL" not in r.stdout and r.stdout.count("passed") == 4
    row = json.loads(r.stdout.strip().splitlines()[-1])
    assert row["n"] == 1 << 20 and row["failures"] == 0 and row["ms"]["cpu_scan_pot"] > 0


def test_sc_header_cpu_namespace_matches_golden_vectors(tmp_path):
    """the header's CPU::scan / compact* against the reference's own cpu.cu outputs (sc_golden.npz)"""
    gold = np.load(os.path.join(GOLD, "sc_golden.npz"))
    src = tmp_path / "t.cpp"
    src.write_text(r'''
#include <cstdio>
#include <vector>
#include "b200_stream_compaction.hpp"
int main(int argc, char** argv) {
    FILE* f = fopen(argv[1], "rb"); int n; if (fread(&n, 4, 1, f) != 1) return 1;
    std::vector<int> in(n), out(n), out2(n);
    if (fread(in.data(), 4, n, f) != (size_t)n) return 1; fclose(f);
    int mode = argv[3][0] - '0', count = n;
    if (mode == 0) StreamCompaction::CPU::scan(n, out.data(), in.data());
    else if (mode == 1) count = StreamCompaction::CPU::compactWithoutScan(n, out.data(), in.data());
    else count = StreamCompaction::CPU::compactWithScan(n, out.data(), in.data());
    f = fopen(argv[2], "wb"); fwrite(out.data(), 4, count, f); fclose(f);
    if (ilog2ceil(1) != 0 || ilog2ceil(5) != 3 || ilog2(8) != 3 || ilog2ceil(8) != 3) return 3;
    return 0;
}''')
    exe = str(tmp_path / "t")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(REPO, "include"), str(src), "-o", exe,
                    "-L", os.path.join(REPO, "cuda_pathtracer_b200"), "-lb200sc",
                    "-Wl,-rpath," + os.path.join(REPO, "cuda_pathtracer_b200")], check=True)

    def run(a, mode):
        fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
        with open(fin, "wb") as f:
            f.write(struct.pack("<i", a.size))
            f.write(np.ascontiguousarray(a, dtype=np.int32).tobytes())
        subprocess.run([exe, fin, fout, str(mode)], check=True)
        return np.fromfile(fout, dtype=np.int32)

    assert np.array_equal(run(gold["scan_in"], 0), gold["scan_pot"])
    assert np.array_equal(run(gold["scan_in"][:-3], 0), gold["scan_npot"])
    assert np.array_equal(run(gold["wrap_in"], 0), gold["wrap_scan"])
    for mode in (1, 2):
        assert np.array_equal(run(gold["compact_in"], mode), gold["compact_pot"])
        assert np.array_equal(run(gold["compact_in"][:-3], mode), gold["compact_npot"])


@pytest.mark.gpu
@pytest.mark.parametrize("lg", [8, 20, 26])
def test_sc_harness_all_variants_pass(lg):
    r = subprocess.run([SC_EXE, str(lg), "--json"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert "FAIL" not in r.stdout and r.stdout.count("passed") == 12
