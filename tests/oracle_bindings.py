"""ctypes access to the CPU oracles under oracle/ -- for tests ONLY (and bench.py's cpu_baseline).

oracle/_build/*.so are built by `make oracle` (gcc/g++, no GPU needed).  oracle/_ref/*.so are the
reference's own sources compiled from /root/reference in the build container; they may be absent
elsewhere, in which case the committed golden vectors are the pin.
"""
import ctypes
import os
import subprocess

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = os.path.join(REPO, "oracle")


def _ensure_built(rel):
    path = os.path.join(ORACLE, rel)
    if not os.path.exists(path) and rel.startswith("_build/"):
        subprocess.run(["make", "-s", "-C", ORACLE, "own"], check=True)
    return path


def _iptr(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


class ScOracle:
    """oracle/sc_oracle.c (restatement of reference cpu.cu) or the reference's cpu.cu itself."""

    def __init__(self, reference=False):
        if reference:
            path = os.path.join(ORACLE, "_ref", "libref_sc_cpu.so")
            if not os.path.exists(path):
                raise FileNotFoundError(path)
            self.lib = ctypes.CDLL(path)
            self._scan = self.lib.ref_cpu_scan
            self._cwo = self.lib.ref_cpu_compact_without_scan
            self._cws = self.lib.ref_cpu_compact_with_scan
            self.lib.ref_cpu_last_ms.restype = ctypes.c_float
        else:
            self.lib = ctypes.CDLL(_ensure_built("_build/liboracle_sc.so"))
            self._scan = self.lib.oracle_scan
            self._cwo = self.lib.oracle_compact_without_scan
            self._cws = self.lib.oracle_compact_with_scan
        self.reference = reference

    def scan(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        o = np.zeros_like(a)
        self._scan(ctypes.c_int(a.size), _iptr(o), _iptr(a))
        return o

    def compact_without_scan(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        o = np.zeros_like(a)
        n = self._cwo(ctypes.c_int(a.size), _iptr(o), _iptr(a))
        return o[:n].copy()

    def compact_with_scan(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        o = np.zeros_like(a)
        n = self._cws(ctypes.c_int(a.size), _iptr(o), _iptr(a))
        return o[:n].copy()

    def last_ms(self):
        return float(self.lib.ref_cpu_last_ms()) if self.reference else None


def gen_array(n_pot, maxval, seed=565):
    """The harness's input (stream_compaction/src/main.cpp:22-23,97-98 with testing_helpers.hpp:52-58):
    n_pot-1 values glibc rand() % maxval after srand(seed), last element forced to 0."""
    lib = ctypes.CDLL(_ensure_built("_build/liboracle_sc.so"))
    a = np.zeros(n_pot, dtype=np.int32)
    if n_pot > 1:
        lib.oracle_gen_array(ctypes.c_int(n_pot - 1), _iptr(a), ctypes.c_int(maxval), ctypes.c_uint(seed))
    return a


class PtOracle:
    """oracle/pt_oracle.cpp: CPU restatement of the reference's path-tracing loop (device arithmetic
    order, bit-identical to the reference's CUDA build).  `scene` is a
    cuda_pathtracer_b200.pathtracer.Scene (POD arrays in the reference's layouts)."""

    def __init__(self):
        self.lib = ctypes.CDLL(_ensure_built("_build/liboracle_pt.so"))
        self.lib.pt_oracle_render.restype = ctypes.c_ulonglong

    def _set_mesh(self, scene, gui=None):
        """hand the scene's triangles / BVH / textures to the oracle (kept alive on self); gui.useBVHtree /
        useBBox select the BVH walk (default) or meshIntersectionTest with / without its bounding-box cull"""
        brute = 0
        if gui is not None and not getattr(gui, "useBVHtree", True):
            brute = 1 if getattr(gui, "useBBox", True) else 2
        self.lib.pt_oracle_set_mesh_mode(brute)
        tris, bvh, texs = getattr(scene, "triangles", ()), getattr(scene, "bvh", ()), getattr(scene, "textures", [])
        dims = (ctypes.c_int * max(1, 3 * len(texs)))(*[v for t in texs for v in (t.width, t.height, t.components)])
        ptrs = (ctypes.c_void_p * max(1, len(texs)))(*[t.pixels.ctypes.data for t in texs])
        self._keep = (scene, dims, ptrs)
        self.lib.pt_oracle_set_mesh(tris if len(tris) else None, len(tris), bvh if len(bvh) else None, len(bvh),
                                    len(texs), dims, ptrs)

    @staticmethod
    def _opts(gui):
        o5 = (ctypes.c_int * 5)(int(gui.russianRoulette), int(gui.sortbyMaterial), int(gui.SSAA), int(gui.DoF), 0)
        o2 = (ctypes.c_float * 2)(float(gui.aperture), float(gui.focal_len))
        return o5, o2

    def rng(self, iteration, index, depth, n):
        out = np.zeros(n, dtype=np.float32)
        self.lib.pt_oracle_rng(iteration, index, depth, n, out.ctypes.data_as(ctypes.c_void_p))
        return out

    def sincos(self, a):
        a = np.ascontiguousarray(a, dtype=np.float32)
        s, c = np.zeros_like(a), np.zeros_like(a)
        self.lib.pt_oracle_sincos(a.size, a.ctypes.data_as(ctypes.c_void_p), s.ctypes.data_as(ctypes.c_void_p),
                                  c.ctypes.data_as(ctypes.c_void_p))
        return s, c

    def raygen(self, scene, gui, iteration):
        o5, o2 = self._opts(gui)
        seg = np.zeros((scene.width * scene.height, 12), dtype=np.float32)
        self.lib.pt_oracle_raygen(ctypes.byref(scene.camera), o5, o2, int(iteration), scene.traceDepth,
                                  seg.ctypes.data_as(ctypes.c_void_p))
        return seg

    def intersect(self, scene, segments, gui=None):
        segments = np.ascontiguousarray(segments, dtype=np.float32)
        out = np.zeros((len(segments), 7), dtype=np.float32)
        self._set_mesh(scene, gui)
        self.lib.pt_oracle_intersect(scene.geoms, len(scene.geoms), len(segments),
                                     segments.ctypes.data_as(ctypes.c_void_p), out.ctypes.data_as(ctypes.c_void_p))
        return out

    def shade(self, scene, gui, iteration, isects, segments):
        seg = np.ascontiguousarray(segments, dtype=np.float32).copy()
        isects = np.ascontiguousarray(isects, dtype=np.float32)
        self._set_mesh(scene)
        self.lib.pt_oracle_shade(scene.materials, int(gui.russianRoulette), int(iteration), len(seg),
                                 isects.ctypes.data_as(ctypes.c_void_p), seg.ctypes.data_as(ctypes.c_void_p))
        return seg

    def render(self, scene, gui, first_iter, n_iters, stride=1, threads=None, out=None):
        """returns (sum image H x W x 3 float32, path segments)"""
        o5, o2 = self._opts(gui)
        if out is None:
            out = np.zeros((scene.height, scene.width, 3), dtype=np.float32)
        if threads is None:
            threads = os.cpu_count() or 1
        self._set_mesh(scene, gui)
        seg = self.lib.pt_oracle_render(ctypes.byref(scene.camera), scene.traceDepth, scene.geoms, len(scene.geoms),
                                        scene.materials, o5, o2, int(first_iter), int(n_iters), int(stride),
                                        out.ctypes.data_as(ctypes.c_void_p), int(threads))
        return out, int(seg)
