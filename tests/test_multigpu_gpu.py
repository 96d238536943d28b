"""The multi-GPU layer of include/b200_mg.h (one process, one host thread per device, NCCL) and the `--gpus K` flags of
the two command-line programs.  With ONE device the layer issues no collective and must reproduce the single-GPU
library exactly (runs on any GPU box); the K >= 2 cases need that many devices and are skipped otherwise
(`gpurun --gpus 2`)."""
import glob
import json
import os
import subprocess

import numpy as np
import pytest

from oracle_bindings import PtOracle, ScOracle, gen_array
from pt_cases import bits_equal, gui_from_flags, load_scene

pytestmark = pytest.mark.gpu

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "cuda_path_tracer")
SC_EXE = os.path.join(REPO, "cuda_pathtracer_b200", "bin", "stream_compaction_test")


def n_devices():
    import torch
    return torch.cuda.device_count()


def group_sizes():
    return [k for k in (1, 2, 3, 4, 8) if k <= n_devices()]


@pytest.mark.parametrize("k", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("n", [0, 1, 5, 4099, (1 << 20) - 3, 1 << 22])
def test_sharded_scan_and_compact_bit_exact(k, n):
    if k > n_devices():
        pytest.skip(f"needs {k} GPUs")
    from cuda_pathtracer_b200.multigpu import DeviceGroup
    g = DeviceGroup(k)
    assert g.device_count == k
    oracle = ScOracle()
    a = gen_array(max(1, 1 << 22), 50)[:n].copy()
    a[::7] = 2 ** 29                                      # the running sum wraps (int32)
    o = np.full(n + 8, -7, dtype=np.int32)
    g.scan(n, o, a)
    assert np.array_equal(o[:n], oracle.scan(a)) and (o[n:] == -7).all()
    b = (gen_array(max(1, 1 << 22), 4)[:n]).copy()
    want = oracle.compact_without_scan(b)
    o[:] = -7
    cnt, ms, ex = g.compact(n, o, b)
    assert cnt == want.size and np.array_equal(o[:cnt], want) and (o[n:] == -7).all()
    g.close()


@pytest.mark.parametrize("k", [1, 2, 4, 8])
def test_group_render_matches_single_gpu(k):
    """iterations split over K devices + ONE all-reduce == the same iterations on one device, up to the fp32
    re-association of the per-pixel sums; same path set (segment count); K = 1: bit for bit with batch = 1"""
    if k > n_devices():
        pytest.skip(f"needs {k} GPUs")
    from cuda_pathtracer_b200 import pathtracer as pt
    from cuda_pathtracer_b200.multigpu import DeviceGroup, GroupRenderer
    sc, gui = load_scene("cornell_small"), gui_from_flags([], batch=1)
    solo = pt.Renderer(sc)
    solo.set_options(gui)
    solo.render(1, 32)
    want, seg = solo.read_sum(), solo.stats()["path_segments"]
    solo.close()
    g = DeviceGroup(k)
    r = GroupRenderer(g, sc)
    r.set_options(gui)
    r.render(1, 20)
    part, _ = r.reduce_read_sum()                         # a preview read-back mid-run: rendering goes on afterwards
    r.render(21, 12)
    got, ar_ms = r.reduce_read_sum()
    assert r.stats()["path_segments"] == seg and r.stats()["iterations"] == 32
    assert np.allclose(got, want, rtol=2e-6, atol=1e-6)
    if k == 1:
        assert bits_equal(got, want)
    assert not np.allclose(part, want, rtol=2e-6, atol=1e-6)
    # sendImageToPBO of the reduced image
    pbo = r.preview_u8(32)
    mean = got.astype(np.float32) / np.float32(32)
    ref = np.clip((mean.astype(np.float64) * 255.0).astype(np.int64), 0, 255).astype(np.uint8)
    assert np.array_equal(pbo[..., :3], ref) and not pbo[..., 3].any()
    r.reset()
    r.render(1, 1)
    one, _ = r.reduce_read_sum()
    assert one.sum() < got.sum() / 8
    r.close()
    g.close()


@pytest.mark.parametrize("k", [2, 4, 8])
def test_cli_gpus_flag(tmp_path, k):
    """`cuda_path_tracer SCENE.json --gpus K` == `--gpus 1` within the re-association tolerance; `--preview-every`
    writes the mid-run images (N4: the reference's 'S'-key save, main.cpp:176-182)"""
    if k > n_devices():
        pytest.skip(f"needs {k} GPUs")
    scene = os.path.join(REPO, "scenes", "cornell_small.json")
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    subprocess.run([EXE, scene, "--iters", "24", "--no-png", "--sum-out", a], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([EXE, scene, "--iters", "24", "--gpus", str(k), "--sum-out", b, "--out", str(tmp_path / "img"),
                        "--preview-every", "10", "--json"], check=True, capture_output=True, text=True)
    x, y = np.fromfile(a, dtype=np.float32), np.fromfile(b, dtype=np.float32)
    assert np.allclose(x, y, rtol=2e-6, atol=1e-6)
    line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert line["gpus"] == k and line["iters"] == 24
    names = sorted(os.path.basename(p) for p in glob.glob(str(tmp_path / "img*.png")))
    assert names == ["img.10samp.png", "img.20samp.png", "img.png"]


def test_cli_preview_every_single_gpu(tmp_path):
    scene = os.path.join(REPO, "scenes", "cornell_tiny.json")
    a, b = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    subprocess.run([EXE, scene, "--iters", "9", "--batch", "1", "--no-png", "--sum-out", a], check=True, stdout=subprocess.DEVNULL)
    subprocess.run([EXE, scene, "--iters", "9", "--batch", "1", "--sum-out", b, "--out", str(tmp_path / "p"),
                    "--preview-every", "4"], check=True, stdout=subprocess.DEVNULL)
    assert bits_equal(np.fromfile(a, dtype=np.float32), np.fromfile(b, dtype=np.float32))     # chunking changes nothing
    names = sorted(os.path.basename(p) for p in glob.glob(str(tmp_path / "p*.png")))
    assert names == ["p.4samp.png", "p.8samp.png", "p.png"]


@pytest.mark.parametrize("k", [2, 4, 8])
def test_scan_harness_gpus_flag(k):
    if k > n_devices():
        pytest.skip(f"needs {k} GPUs")
    r = subprocess.run([SC_EXE, "22", "--gpus", str(k), "--json"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["failures"] == 0 and line["gpus"] == k
    assert {"sharded_scan_pot", "sharded_scan_npot", "sharded_compact_pot", "sharded_compact_npot"} <= set(line["ms"])
