"""GPU parity tests of the wavefront path tracer, through the C ABI (include/b200_pt.h).

Tiers:
  T1  every stage (camera rays, closest hit, shading) on identical inputs, BIT FOR BIT against
      vectors produced by the reference's own kernels (tests/golden/*.stage.npz) and against the
      CPU oracle;
  T3  whole renders against float SUM images written by the reference's own CUDA build
      (tests/golden/*.refsum.bin): bit-identical with batch = 1 (the reference's accumulation
      order), and within north_star's tolerance -- mean absolute error <= 1e-3 and PSNR >= 45 dB
      on the mean image -- with the default batched wavefront;
  full-size properties at BASELINE's Cornell 800x800: iteration-split additivity, batch
      invariance, segment accounting, against the oracle on a sample of pixels' worth of work.
"""
import os
import subprocess

import numpy as np
import pytest

from oracle_bindings import PtOracle
from pt_cases import (GOLD, RENDERS, SORT_RENDERS, STAGES, bits_equal, gui_from_flags, image_metrics, load_scene, ref_sum,
                      texture_reads_in_bounds)

pytestmark = pytest.mark.gpu

MAE_TOL = 1e-3      # north_star: mean absolute error of the mean image
PSNR_TOL = 45.0     # north_star: PSNR in dB, peak 1.0


@pytest.fixture(scope="module")
def pt():
    from cuda_pathtracer_b200 import pathtracer as m
    return m


@pytest.fixture(scope="module")
def torch():
    import torch as t
    return t


@pytest.fixture(scope="module")
def oracle():
    return PtOracle()


def dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda().reshape(-1)


# ---- T1: per-stage ----------------------------------------------------------------------------------
@pytest.mark.parametrize("scene,tag,flags", STAGES)
def test_stages_vs_reference_kernels(pt, torch, scene, tag, flags):
    d = np.load(os.path.join(GOLD, f"{scene}.{tag}.stage.npz"))
    sc, it = load_scene(scene), int(d["iteration"])
    r = pt.Renderer(sc)
    r.set_options(gui_from_flags(flags))
    P = sc.width * sc.height
    seg = torch.zeros(P * 12, dtype=torch.float32, device="cuda")
    r.debug_raygen(it, seg.data_ptr())
    torch.cuda.synchronize()
    n = len(d["raygen"])
    assert bits_equal(seg.cpu().numpy().reshape(P, 12)[:n], d["raygen"])
    b = 0
    while f"b{b}_in" in d.files:
        seg_in, isect, seg_out = d[f"b{b}_in"], d[f"b{b}_isect"], d[f"b{b}_out"]
        m = len(seg_in)
        d_seg, d_is = dev(torch, seg_in), torch.zeros(m * 7, dtype=torch.float32, device="cuda")
        r.debug_intersect(m, d_seg.data_ptr(), d_is.data_ptr())
        torch.cuda.synchronize()
        assert bits_equal(d_is.cpu().numpy().reshape(m, 7), isect), f"intersect, bounce {b}"
        d_is = dev(torch, isect)
        r.debug_shade(it, m, d_is.data_ptr(), d_seg.data_ptr())
        torch.cuda.synchronize()
        ok = texture_reads_in_bounds(sc, isect)              # all rows, except with --no-bvh (see pt_cases.py)
        assert ok.all() or "--no-bvh" in flags
        assert bits_equal(d_seg.cpu().numpy().reshape(m, 12)[ok], seg_out[ok]), f"shade, bounce {b}"
        b += 1
    r.close()


@pytest.mark.parametrize("scene,flags", [("cornell_small", []), ("synthetic_600_small", []),
                                          ("cornell_glass_small", ["--no-ssaa", "--no-dof"])])
def test_stages_vs_oracle_all_bounces(pt, torch, oracle, scene, flags):
    """lock-step wavefront: each stage of every bounce against the CPU oracle on the same inputs"""
    sc, gui, it = load_scene(scene), gui_from_flags(flags), 7
    r = pt.Renderer(sc)
    r.set_options(gui)
    P = sc.width * sc.height
    seg = torch.zeros(P * 12, dtype=torch.float32, device="cuda")
    r.debug_raygen(it, seg.data_ptr())
    torch.cuda.synchronize()
    segs = oracle.raygen(sc, gui, it)
    assert bits_equal(seg.cpu().numpy().reshape(P, 12), segs)
    for bounce in range(sc.traceDepth):
        n = len(segs)
        if n == 0:
            break
        want_is = oracle.intersect(sc, segs)
        d_seg, d_is = dev(torch, segs), torch.zeros(n * 7, dtype=torch.float32, device="cuda")
        r.debug_intersect(n, d_seg.data_ptr(), d_is.data_ptr())
        torch.cuda.synchronize()
        assert bits_equal(d_is.cpu().numpy().reshape(n, 7), want_is), f"intersect, bounce {bounce}"
        want_out = oracle.shade(sc, gui, it, want_is, segs)
        r.debug_shade(it, n, d_is.data_ptr(), d_seg.data_ptr())
        torch.cuda.synchronize()
        assert bits_equal(d_seg.cpu().numpy().reshape(n, 12), want_out), f"shade, bounce {bounce}"
        segs = np.ascontiguousarray(want_out[want_out[:, 10].view(np.int32) != 0])
    r.close()


# ---- T3: whole renders ------------------------------------------------------------------------------
@pytest.mark.parametrize("scene,iters,flags,tag", RENDERS)
def test_render_bit_exact_vs_reference_cuda(pt, scene, iters, flags, tag):
    sc = load_scene(scene)
    r = pt.Renderer(sc)
    r.set_options(gui_from_flags(flags, batch=1))       # the reference's accumulation order
    r.render(1, iters)
    img = r.read_sum()
    assert bits_equal(img, ref_sum(scene, iters, tag, img.shape))
    st = r.stats()
    # ray generation + one fused launch per bounce; tree-walking scenes (meshes, > 24 spheres / cubes) run closest
    # hit and shading as two launches per bounce
    walks = len(sc.triangles) > 0 or sum(1 for g in sc.geoms if g.type != 2) > 24
    assert st["iterations"] == iters and st["kernel_launches"] == iters * (1 + (2 if walks else 1) * sc.traceDepth)
    r.close()


@pytest.mark.parametrize("scene,iters,flags,tag", RENDERS)
@pytest.mark.parametrize("batch", [0, 3])
def test_render_batched_within_north_star_tolerance(pt, scene, iters, flags, tag, batch):
    sc = load_scene(scene)
    r = pt.Renderer(sc)
    r.set_options(gui_from_flags(flags, batch=batch))
    r.render(1, iters)
    img = r.read_sum()
    ref = ref_sum(scene, iters, tag, img.shape)
    mae, psnr = image_metrics(img, ref, iters)
    assert mae <= MAE_TOL and psnr >= PSNR_TOL
    # in fact only the order of the per-pixel float additions differs
    assert np.allclose(img, ref, rtol=2e-6, atol=1e-6)
    r.close()


# ---- geometry BVH: prunes, never changes the hit ----------------------------------------------------------
@pytest.mark.parametrize("scene,iters,flags,tag", RENDERS + SORT_RENDERS)
@pytest.mark.parametrize("geom_bvh", [1, -1])
def test_geometry_bvh_on_and_off_are_bit_identical_to_the_reference(pt, scene, iters, flags, tag, geom_bvh):
    """forced on (even for the 7-primitive Cornell box and the mesh scene's six cubes) and forced off
    (linear scan, also for the 600-primitive scene): same float SUM image as the reference's CUDA build"""
    sc = load_scene(scene)
    r = pt.Renderer(sc)
    g = gui_from_flags(flags, batch=1)
    g.geomBVH = geom_bvh
    r.set_options(g)
    r.render(1, iters)
    img = r.read_sum()
    assert bits_equal(img, ref_sum(scene, iters, tag, img.shape))
    r.close()


def test_geometry_bvh_full_size_many_primitives(pt):
    """BASELINE config 5's scene (4096 primitives) at a reduced resolution: BVH == linear scan, bit for bit,
    incl. after an update_scene that moves primitives (the tree is rebuilt)"""
    sc = load_scene("synthetic_4096")
    sc.camera.resolution[0], sc.camera.resolution[1] = 256, 144
    imgs = []
    for mode in (1, -1):
        r = pt.Renderer(sc)
        g = gui_from_flags([], batch=1)
        g.geomBVH = mode
        r.set_options(g)
        r.render(1, 2)
        a = r.read_sum()
        for k in range(0, len(sc.geoms), 7):                 # shift a subset of the primitives
            sc.geoms[k].transform[12] += 0.25
            sc.geoms[k].inverseTransform[12] -= 0.25 * sc.geoms[k].inverseTransform[0]
            sc.geoms[k].inverseTransform[13] -= 0.25 * sc.geoms[k].inverseTransform[1]
            sc.geoms[k].inverseTransform[14] -= 0.25 * sc.geoms[k].inverseTransform[2]
        r.update_scene(sc)
        r.reset()
        r.render(3, 2)
        imgs.append((a, r.read_sum(), r.stats()["path_segments"]))
        r.close()
        sc = load_scene("synthetic_4096")
        sc.camera.resolution[0], sc.camera.resolution[1] = 256, 144
    assert bits_equal(imgs[0][0], imgs[1][0]) and bits_equal(imgs[0][1], imgs[1][1]) and imgs[0][2] == imgs[1][2]
    assert not bits_equal(imgs[0][0], imgs[0][1])


# ---- material sort (A9) and first-bounce cache ----------------------------------------------------------
def have_golden(scene, iters, tag):
    return os.path.exists(os.path.join(GOLD, f"{scene}.{tag}.{iters}it.refsum.bin"))


@pytest.mark.parametrize("scene,iters,flags,tag", SORT_RENDERS)
def test_sorted_render_bit_exact_vs_reference_cuda(pt, oracle, scene, iters, flags, tag):
    """sort-by-material on: bit-identical to the reference's CUDA build (thrust::sort_by_key) and to
    the oracle's stable sort, with batch = 1; same paths (segment count) with the default batch"""
    sc = load_scene(scene)
    gui = gui_from_flags(flags, batch=1)
    r = pt.Renderer(sc)
    r.set_options(gui)
    r.render(1, iters)
    img = r.read_sum()
    want, segments = oracle.render(sc, gui, 1, iters)
    assert bits_equal(img, want)
    assert r.stats()["path_segments"] == segments
    if have_golden(scene, iters, tag):
        assert bits_equal(img, ref_sum(scene, iters, tag, img.shape))
    r.reset()
    r.set_options(gui_from_flags(flags, batch=0))
    r.render(1, iters)
    assert r.stats()["path_segments"] == segments
    assert np.allclose(r.read_sum(), want, rtol=2e-6, atol=1e-6)
    r.close()


def test_sort_with_more_than_256_materials(pt, oracle):
    """material ids beyond one radix digit: the LSD passes of the round-1 kernels (k_sort_hist / k_sort_scatter) instead of
    the one-pass bucketing.  The glass scene's materials replicated to 700 entries, every primitive moved to a copy of its
    material with a different (large) id -- same shading, another sort order -- bit for bit against the oracle's stable sort"""
    sc = load_scene("cornell_glass_small")
    n, M = len(sc.materials), 700
    assert M % n == 0
    mats = (pt.Material * M)()
    for i in range(M):
        mats[i] = sc.materials[i % n]
    sc.materials = mats
    for gi, g in enumerate(sc.geoms):
        g.materialid = g.materialid + n * ((gi * 37 + 5) % (M // n))
    assert max(g.materialid for g in sc.geoms) > 255
    gui = gui_from_flags(["--sort"], batch=1)
    r = pt.Renderer(sc)
    r.set_options(gui)
    r.render(1, 6)
    img, seg = r.read_sum(), r.stats()["path_segments"]
    want, segments = oracle.render(sc, gui, 1, 6)
    assert seg == segments and bits_equal(img, want)
    # ... and with the first-bounce cache (the cached order by material comes from the last radix pass)
    gui2 = gui_from_flags(["--sort", "--no-ssaa", "--no-dof"], batch=1)
    gui2.firstBounceCache = True
    r.reset()
    r.set_options(gui2)
    r.render(1, 5)
    want2, seg2 = oracle.render(sc, gui2, 1, 5)
    assert r.stats()["path_segments"] == seg2 and bits_equal(r.read_sum(), want2)
    r.close()


def test_sorted_golden_fixtures_present():
    missing = [c for c in SORT_RENDERS if not have_golden(c[0], c[1], c[3])]
    assert not missing, f"reference-CUDA goldens missing for {missing}: run tests/golden/make_pt_golden.py renders"


@pytest.mark.parametrize("scene,flags", [("cornell_small", ["--no-ssaa", "--no-dof"]),
                                          ("cornell_glass_small", ["--no-ssaa", "--no-dof", "--sort"]),
                                          ("synthetic_600_small", ["--no-ssaa", "--no-dof"])])
@pytest.mark.parametrize("batch", [1, 3])
def test_first_bounce_cache_changes_nothing(pt, scene, flags, batch):
    """the cache replaces bounce 0's closest-hit search by a read of the identical record: the image
    must be bit-identical with batch = 1 (and the same paths otherwise); it is ignored (not an
    error) while SSAA or DoF is on, because the camera rays then differ per iteration"""
    sc = load_scene(scene)
    r = pt.Renderer(sc)
    gui = gui_from_flags(flags, batch=batch)
    r.set_options(gui)
    r.render(1, 9)
    plain, seg = r.read_sum(), r.stats()["path_segments"]
    r.reset()
    gui.firstBounceCache = True
    r.set_options(gui)
    r.render(1, 4)                       # fills the cache in the first batch, uses it afterwards
    r.render(5, 5)
    cached = r.read_sum()
    assert r.stats()["path_segments"] == seg
    if batch == 1:
        assert bits_equal(cached, plain)
    else:
        assert np.allclose(cached, plain, rtol=2e-6, atol=1e-6)
    # a scene update drops the cache (the reference re-inits on every camera change)
    r.update_scene(sc)
    r.reset()
    r.render(1, 9)
    assert np.allclose(r.read_sum(), plain, rtol=2e-6, atol=1e-6)
    # with SSAA on the option is ignored
    r.reset()
    gui.SSAA = True
    r.set_options(gui)
    r.render(1, 3)
    want = pt.Renderer(sc)
    gui.firstBounceCache = False
    want.set_options(gui)
    want.render(1, 3)
    assert np.allclose(r.read_sum(), want.read_sum(), rtol=2e-6, atol=1e-6)
    want.close()
    r.close()


def test_segments_and_traced_depth_vs_oracle(pt, oracle):
    sc, gui = load_scene("cornell_small"), gui_from_flags([], batch=4)
    r = pt.Renderer(sc)
    r.set_options(gui)
    r.render(3, 6, 2)                                    # iterations 3, 5, ..., 13
    img = r.read_sum()
    want, segments = oracle.render(sc, gui, 3, 6, stride=2)
    st = r.stats()
    assert st["path_segments"] == segments
    assert 1 <= st["traced_depth"] <= sc.traceDepth
    assert np.allclose(img, want, rtol=2e-6, atol=1e-6)
    r.reset()
    assert r.stats()["path_segments"] == 0 and not r.read_sum().any()
    r.close()


def test_reference_shaped_api(pt, torch):
    """InitDataContainer / pathtraceInit / pathtrace / pathtraceFree (pathtrace.h:6-9): one call per
    iteration, the SUM image left in scene.state_image, the 8-bit preview in the caller's buffer."""
    scene, iters, flags, tag = RENDERS[0]
    sc = load_scene(scene)
    gui = pt.GuiDataContainer()
    pt.InitDataContainer(gui)
    pt.pathtraceFree()                                   # main.cpp:144 calls Free before the first Init
    pt.pathtraceInit(sc)
    pbo = torch.zeros(sc.width * sc.height * 4, dtype=torch.uint8, device="cuda")
    for it in range(1, iters + 1):
        pt.pathtrace(pbo.data_ptr(), 0, it)
    assert bits_equal(sc.state_image, ref_sum(scene, iters, tag, sc.state_image.shape))
    assert 1 <= gui.TracedDepth <= sc.traceDepth
    # sendImageToPBO (pathtrace.cu:65-86): clamp(int(pix / iter * 255.0), 0, 255), w = 0
    mean = sc.state_image.astype(np.float32) / np.float32(iters)
    want = np.clip((mean.astype(np.float64) * 255.0).astype(np.int64), 0, 255).astype(np.uint8)
    got = pbo.cpu().numpy().reshape(sc.height, sc.width, 4)
    assert np.array_equal(got[..., :3], want) and not got[..., 3].any()
    pt.pathtraceFree()
    with pytest.raises(RuntimeError):
        pt.pathtrace(None, 0, 1)


def test_errors_are_returned_not_fatal(pt):
    from cuda_pathtracer_b200 import B200Error
    sc = load_scene("cornell_tiny")
    sc.geoms[0].type = 2                                 # MESH whose triangle range is not in the scene
    sc.geoms[0].tri_end_idx = 5
    with pytest.raises(B200Error):
        pt.Renderer(sc)
    sc = load_scene("cornell_tiny")
    sc.geoms[0].type = 7                                 # unknown GeomType
    with pytest.raises(B200Error):
        pt.Renderer(sc)
    sc = load_scene("cornell_tiny")
    sc.materials[0].texture_id = 0
    with pytest.raises(B200Error):
        pt.Renderer(sc)
    sc = load_scene("cornell_tiny")
    r = pt.Renderer(sc)
    gui = pt.GuiDataContainer()
    gui.batch = -1
    with pytest.raises(B200Error):
        r.set_options(gui)
    r.close()


# ---- full size: BASELINE config 2 (Cornell 800x800, depth 8) -------------------------------------------
def test_full_size_cornell_properties(pt, oracle):
    sc = load_scene("cornell")
    assert (sc.width, sc.height, sc.traceDepth) == (800, 800, 8)
    r = pt.Renderer(sc)
    g1 = gui_from_flags([], batch=1)
    r.set_options(g1)
    r.render(1, 8)
    a = r.read_sum()
    seg_a = r.stats()["path_segments"]
    # (1) batch invariance: same paths, only the float addition order may differ
    r.reset()
    r.set_options(gui_from_flags([], batch=0))
    r.render(1, 8)
    b = r.read_sum()
    assert r.stats()["path_segments"] == seg_a
    assert np.allclose(a, b, rtol=2e-6, atol=1e-6)
    # (2) iteration-split additivity (what the multi-GPU split relies on): odd + even == all
    r.reset()
    r.render(1, 4, 2)
    r.render(2, 4, 2)
    c = r.read_sum()
    assert r.stats()["path_segments"] == seg_a
    assert np.allclose(a, c, rtol=2e-6, atol=1e-6)
    # (3) about three segments per camera path on this scene, image mean ~0.14
    assert 2.5 < seg_a / (8 * 640000) < 3.6
    assert 0.10 < a.mean() / 8 < 0.18
    # (4) one full-size iteration against the CPU oracle, bit for bit
    r.reset()
    r.set_options(g1)
    r.render(5, 1)
    want, seg = oracle.render(sc, g1, 5, 1)
    assert r.stats()["path_segments"] == seg
    assert bits_equal(r.read_sum(), want)
    r.close()


# ---- full size: BASELINE config 4 (glass / mirror / diffuse 1920x1080, depth 12, material sort + first-bounce cache) ----
def test_full_size_config4_sort_and_cache_vs_oracle(pt, oracle):
    """cornell_glass_1080p as BASELINE.json states it: sort by material ON, first-bounce cache ON (hence SSAA and DoF
    off), full resolution.  Two iterations with batch = 1: the first fills the cache, the second uses it.  Bit for bit
    against the CPU oracle (which is pinned to the reference build's sorted renders on the small glass scene)."""
    sc = load_scene("cornell_glass_1080p")
    assert (sc.width, sc.height, sc.traceDepth) == (1920, 1080, 12)
    gui = gui_from_flags(["--sort", "--no-ssaa", "--no-dof"], batch=1)
    gui.firstBounceCache = True
    r = pt.Renderer(sc)
    r.set_options(gui)
    r.render(1, 2)
    img, seg = r.read_sum(), r.stats()["path_segments"]
    want, segments = oracle.render(sc, gui, 1, 2)
    assert seg == segments
    assert bits_equal(img, want)
    # the default batch: same paths, float addition order only
    r.reset()
    gui.batch = 0
    r.set_options(gui)
    r.render(1, 2)
    assert r.stats()["path_segments"] == segments
    assert np.allclose(r.read_sum(), want, rtol=2e-6, atol=1e-6)
    r.close()


def test_live_reference_cuda_build_if_present(pt):
    """When the reference's own CUDA build travelled with the snapshot (oracle/_ref/ref_pathtracer,
    built in the build container), run it on a case that has no committed golden."""
    exe = os.path.join(os.path.dirname(GOLD), "..", "oracle", "_ref", "ref_pathtracer")
    scene_json = os.path.join(os.path.dirname(GOLD), "..", "scenes", "cornell_tiny.json")
    if not os.path.exists(exe):
        pytest.skip("reference CUDA build not present on this box")
    out = os.path.join("/tmp", "ref_tiny.bin")
    subprocess.run([exe, scene_json, "--iters", "24", "--out", out, "--aperture", "0.3", "--focal", "8"],
                   check=True, stdout=subprocess.DEVNULL)
    sc = load_scene("cornell_tiny")
    r = pt.Renderer(sc)
    g = gui_from_flags([], batch=1)
    g.aperture, g.focal_len = 0.3, 8.0
    r.set_options(g)
    r.render(1, 24)
    img = r.read_sum()
    assert bits_equal(img, np.fromfile(out, dtype=np.float32).reshape(img.shape))
    r.close()
