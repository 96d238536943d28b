"""Python mirror of the reference's render-driver interface (path_tracer/src/pathtrace.h:6-9:
``InitDataContainer``, ``pathtraceInit``, ``pathtraceFree``, ``pathtrace``) over the C ABI of
include/b200_pt.h.  Same names, argument meaning and conventions (1-based ``iteration``; the
float SUM image is left in ``scene.state_image``; toggles are read from the ``GuiDataContainer``
on every call).  Used by the tests and bench.py; the product is the CUDA code behind the ABI.
Nothing here imports anything under ``oracle/``; there is no CPU fallback.
"""
import ctypes
import struct

import numpy as np

from ._lib import B200Error, load


class Geom(ctypes.Structure):            # == struct Geom, sceneStructs.h:24-40 (272 B)
    _fields_ = [("type", ctypes.c_int32), ("materialid", ctypes.c_int32),
                ("translation", ctypes.c_float * 3), ("rotation", ctypes.c_float * 3),
                ("scale", ctypes.c_float * 3), ("transform", ctypes.c_float * 16),
                ("inverseTransform", ctypes.c_float * 16), ("invTranspose", ctypes.c_float * 16),
                ("tri_start_idx", ctypes.c_int32), ("tri_end_idx", ctypes.c_int32),
                ("boundingbox_idx", ctypes.c_int32), ("min_bound", ctypes.c_float * 3),
                ("max_bound", ctypes.c_float * 3)]


class Material(ctypes.Structure):        # == struct Material, sceneStructs.h:42-56 (48 B)
    _fields_ = [("color", ctypes.c_float * 3), ("specular_exponent", ctypes.c_float),
                ("specular_color", ctypes.c_float * 3), ("hasReflective", ctypes.c_float),
                ("hasRefractive", ctypes.c_float), ("indexOfRefraction", ctypes.c_float),
                ("emittance", ctypes.c_float), ("texture_id", ctypes.c_int32)]


class Camera(ctypes.Structure):          # == struct Camera, sceneStructs.h:58-67 (84 B)
    _fields_ = [("resolution", ctypes.c_int32 * 2), ("position", ctypes.c_float * 3),
                ("lookAt", ctypes.c_float * 3), ("view", ctypes.c_float * 3),
                ("up", ctypes.c_float * 3), ("right", ctypes.c_float * 3),
                ("fov", ctypes.c_float * 2), ("pixelLength", ctypes.c_float * 2)]


class Triangle(ctypes.Structure):        # == class Triangle, sceneStructs.h:103-160 (124 B)
    _fields_ = [("id", ctypes.c_int32), ("vertices", (ctypes.c_float * 3) * 3), ("uv", (ctypes.c_float * 2) * 3),
                ("normals", (ctypes.c_float * 3) * 3), ("bbox_min", ctypes.c_float * 3), ("bbox_max", ctypes.c_float * 3)]


class BVHNode(ctypes.Structure):         # == class BVHTree, BVH_tree.h:57-64 (40 B)
    _fields_ = [("min_corner", ctypes.c_float * 3), ("max_corner", ctypes.c_float * 3), ("sub_areas", ctypes.c_int32),
                ("axis", ctypes.c_int32), ("first_area_idx", ctypes.c_int32), ("rchild_idx", ctypes.c_int32)]


class _TextureC(ctypes.Structure):
    _fields_ = [("width", ctypes.c_int32), ("height", ctypes.c_int32), ("components", ctypes.c_int32),
                ("pixels", ctypes.c_void_p)]


class _SceneExtC(ctypes.Structure):
    _fields_ = [("n_triangles", ctypes.c_int32), ("triangles", ctypes.POINTER(Triangle)),
                ("n_bvh_nodes", ctypes.c_int32), ("bvh", ctypes.POINTER(BVHNode)),
                ("n_textures", ctypes.c_int32), ("textures", ctypes.POINTER(_TextureC))]


class Texture:
    """class Texture (sceneStructs.h:162-189): width x height x components bytes, top row first"""

    def __init__(self, width, height, components, pixels):
        self.width, self.height, self.components = int(width), int(height), int(components)
        self.pixels = np.ascontiguousarray(pixels, dtype=np.uint8).reshape(-1)
        assert self.pixels.size == self.width * self.height * self.components


class _SceneC(ctypes.Structure):
    _fields_ = [("camera", Camera), ("trace_depth", ctypes.c_int32), ("n_geoms", ctypes.c_int32),
                ("geoms", ctypes.POINTER(Geom)), ("n_materials", ctypes.c_int32),
                ("materials", ctypes.POINTER(Material))]


class _OptionsC(ctypes.Structure):
    _fields_ = [("russian_roulette", ctypes.c_int32), ("sort_by_material", ctypes.c_int32),
                ("ssaa", ctypes.c_int32), ("dof", ctypes.c_int32), ("aperture", ctypes.c_float),
                ("focal_len", ctypes.c_float), ("first_bounce_cache", ctypes.c_int32),
                ("batch", ctypes.c_int32), ("geom_bvh", ctypes.c_int32),
                ("mesh_brute_force", ctypes.c_int32), ("mesh_no_bbox", ctypes.c_int32),
                ("use_thrust_partition", ctypes.c_int32)]


class _StatsC(ctypes.Structure):
    _fields_ = [("path_segments", ctypes.c_uint64), ("iterations", ctypes.c_uint64),
                ("kernel_launches", ctypes.c_uint64), ("last_render_ms", ctypes.c_float),
                ("traced_depth", ctypes.c_int32), ("bounce_kernel_ms", ctypes.c_double),
                ("bounce_kernel_launches", ctypes.c_uint64)]


assert ctypes.sizeof(Geom) == 272 and ctypes.sizeof(Material) == 48 and ctypes.sizeof(Camera) == 84
assert ctypes.sizeof(Triangle) == 124 and ctypes.sizeof(BVHNode) == 40


class GuiDataContainer:
    """utilities.h:17-34, same field names and defaults (+ the new knobs)."""

    def __init__(self):
        self.TracedDepth = 0
        self.russianRoulette = True
        self.useBVHtree = True
        self.useBBox = True
        self.sortbyMaterial = False
        self.useThrustPartition = False     # pathtrace.cu:499-503: same stable partition, same image (b200_pt.h)
        self.SSAA = True
        self.DoF = True
        self.aperture = 0.1
        self.focal_len = 10.0
        self.firstBounceCache = False       # new
        self.batch = 0                      # new: iterations in flight (0 = default, 1 = exact order)
        self.geomBVH = 0                    # new: BVH over spheres / cubes (0 = auto, 1 = always, -1 = linear scan)


class Scene:
    """The POD part of the reference's ``Scene`` (scene.h:17-36): geoms, materials, state."""

    def __init__(self, camera, trace_depth, iterations, geoms, materials, triangles=None, bvh=None, textures=None):
        self.camera = camera
        self.traceDepth = int(trace_depth)
        self.iterations = int(iterations)
        self.geoms = geoms
        self.materials = materials
        self.triangles = triangles if triangles is not None else (Triangle * 0)()
        self.bvh = bvh if bvh is not None else (BVHNode * 0)()
        self.textures = list(textures or [])
        w, h = camera.resolution[0], camera.resolution[1]
        self.state_image = np.zeros((h, w, 3), dtype=np.float32)   # RenderState::image (SUM)

    @property
    def width(self):
        return self.camera.resolution[0]

    @property
    def height(self):
        return self.camera.resolution[1]

    @staticmethod
    def from_pod_file(path):
        """Read a "B2PTSCN1" dump: header, Camera, Geom[], Material[] (see oracle/ref_scene_dump.cpp
        and the repo's own loader, which writes the same layout)."""
        raw = open(path, "rb").read()
        if raw[:8] not in (b"B2PTSCN1", b"B2PTSCN2"):
            raise ValueError(f"{path}: not a B2PTSCN1 / B2PTSCN2 scene dump")
        n_geoms, n_mats, depth, iters = struct.unpack_from("<4i", raw, 8)
        off = 24
        cam = Camera.from_buffer_copy(raw, off)
        off += ctypes.sizeof(Camera)
        geoms = (Geom * n_geoms).from_buffer_copy(raw, off)
        off += ctypes.sizeof(Geom) * n_geoms
        mats = (Material * n_mats).from_buffer_copy(raw, off)
        off += ctypes.sizeof(Material) * n_mats
        if raw[:8] == b"B2PTSCN1":
            return Scene(cam, depth, iters, geoms, mats)
        n_tris, n_nodes, n_tex, _ = struct.unpack_from("<4i", raw, off)
        off += 16
        tris = (Triangle * n_tris).from_buffer_copy(raw, off)
        off += ctypes.sizeof(Triangle) * n_tris
        bvh = (BVHNode * n_nodes).from_buffer_copy(raw, off)
        off += ctypes.sizeof(BVHNode) * n_nodes
        textures = []
        for _ in range(n_tex):
            w, h, comp, _pad = struct.unpack_from("<4i", raw, off)
            off += 16
            textures.append(Texture(w, h, comp, np.frombuffer(raw, dtype=np.uint8, count=w * h * comp, offset=off)))
            off += w * h * comp
        return Scene(cam, depth, iters, geoms, mats, tris, bvh, textures)


def _lib():
    lib = load("libb200pt.so")
    if not getattr(lib, "_b200pt_typed", False):
        vp, ci = ctypes.c_void_p, ctypes.c_int
        lib.b200pt_default_options.argtypes = [ctypes.POINTER(_OptionsC)]
        lib.b200pt_default_options.restype = None
        lib.b200pt_create.argtypes = [ctypes.POINTER(_SceneC), ctypes.POINTER(vp)]
        lib.b200pt_create_ext.argtypes = [ctypes.POINTER(_SceneC), ctypes.POINTER(_SceneExtC), ctypes.POINTER(vp)]
        lib.b200pt_destroy.argtypes = [vp]
        lib.b200pt_set_options.argtypes = [vp, ctypes.POINTER(_OptionsC)]
        lib.b200pt_reset.argtypes = [vp]
        lib.b200pt_update_scene.argtypes = [vp, ctypes.POINTER(_SceneC), vp]
        lib.b200pt_set_profiling.argtypes = [vp, ci]
        lib.b200pt_render.argtypes = [vp, ci, ci, ci, vp, ci]
        lib.b200pt_tonemap_u8.argtypes = [vp, ci, vp, vp]
        lib.b200pt_read_sum.argtypes = [vp, vp]
        lib.b200pt_write_sum.argtypes = [vp, vp]
        lib.b200pt_sum_device_ptr.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
        lib.b200pt_get_stats.argtypes = [vp, ctypes.POINTER(_StatsC)]
        lib.b200pt_debug_raygen.argtypes = [vp, ci, vp, vp]
        lib.b200pt_debug_intersect.argtypes = [vp, ci, vp, vp, vp]
        lib.b200pt_debug_shade.argtypes = [vp, ci, ci, vp, vp, vp]
        lib.b200pt_error_string.argtypes = [ci]
        lib.b200pt_error_string.restype = ctypes.c_char_p
        lib._b200pt_typed = True
    return lib


def check(code):
    if code != 0:
        raise B200Error(code, _lib().b200pt_error_string(code).decode())


class Renderer:
    """One b200pt handle.  ``render`` is the batched entry point; ``pathtrace`` below is the
    reference-shaped one-iteration call."""

    def __init__(self, scene):
        self.scene = scene
        sc = _SceneC(scene.camera, scene.traceDepth, len(scene.geoms), scene.geoms,
                     len(scene.materials), scene.materials)
        self._h = ctypes.c_void_p()
        tex = (_TextureC * len(scene.textures))(*[_TextureC(t.width, t.height, t.components, t.pixels.ctypes.data)
                                                  for t in scene.textures])
        ext = _SceneExtC(len(scene.triangles), scene.triangles, len(scene.bvh), scene.bvh, len(scene.textures), tex)
        check(_lib().b200pt_create_ext(ctypes.byref(sc), ctypes.byref(ext), ctypes.byref(self._h)))

    def close(self):
        if self._h:
            _lib().b200pt_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_options(self, gui):
        o = _OptionsC(int(gui.russianRoulette), int(gui.sortbyMaterial), int(gui.SSAA), int(gui.DoF),
                      float(gui.aperture), float(gui.focal_len), int(gui.firstBounceCache), int(gui.batch),
                      int(getattr(gui, "geomBVH", 0)), int(not gui.useBVHtree), int(not gui.useBBox),
                      int(getattr(gui, "useThrustPartition", False)))
        check(_lib().b200pt_set_options(self._h, ctypes.byref(o)))

    def reset(self):
        check(_lib().b200pt_reset(self._h))

    def update_scene(self, scene, stream=0):
        """re-upload geoms / materials / camera (same counts and resolution)"""
        sc = _SceneC(scene.camera, scene.traceDepth, len(scene.geoms), scene.geoms,
                     len(scene.materials), scene.materials)
        check(_lib().b200pt_update_scene(self._h, ctypes.byref(sc), stream))
        self.scene = scene

    def set_profiling(self, enable):
        check(_lib().b200pt_set_profiling(self._h, int(bool(enable))))

    def render(self, first_iter, n_iters, stride=1, stream=0, sync=True):
        check(_lib().b200pt_render(self._h, int(first_iter), int(n_iters), int(stride), stream, int(sync)))

    def tonemap_u8(self, iter_divisor, dev_rgba8_ptr, stream=0):
        check(_lib().b200pt_tonemap_u8(self._h, int(iter_divisor), dev_rgba8_ptr, stream))

    def read_sum(self, out=None):
        if out is None:
            out = np.empty((self.scene.height, self.scene.width, 3), dtype=np.float32)
        assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.size == 3 * self.scene.width * self.scene.height
        check(_lib().b200pt_read_sum(self._h, out.ctypes.data))
        return out

    def write_sum(self, img):
        """replace the accumulation image (restartable rendering)"""
        img = np.ascontiguousarray(img, dtype=np.float32)
        assert img.size == 3 * self.scene.width * self.scene.height
        check(_lib().b200pt_write_sum(self._h, img.ctypes.data))

    def sum_device_ptr(self):
        p, n = ctypes.c_void_p(), ctypes.c_size_t()
        check(_lib().b200pt_sum_device_ptr(self._h, ctypes.byref(p), ctypes.byref(n)))
        return p.value, n.value

    def stats(self):
        s = _StatsC()
        check(_lib().b200pt_get_stats(self._h, ctypes.byref(s)))
        return {"path_segments": s.path_segments, "iterations": s.iterations,
                "kernel_launches": s.kernel_launches, "last_render_ms": s.last_render_ms,
                "traced_depth": s.traced_depth, "bounce_kernel_ms": s.bounce_kernel_ms,
                "bounce_kernel_launches": s.bounce_kernel_launches}

    # per-stage probes (device pointers, reference AoS layouts; see include/b200_pt.h)
    def debug_raygen(self, iteration, d_segments, stream=0):
        check(_lib().b200pt_debug_raygen(self._h, int(iteration), d_segments, stream))

    def debug_intersect(self, n, d_segments, d_isects, stream=0):
        check(_lib().b200pt_debug_intersect(self._h, int(n), d_segments, d_isects, stream))

    def debug_shade(self, iteration, n, d_isects, d_segments, stream=0):
        check(_lib().b200pt_debug_shade(self._h, int(iteration), int(n), d_isects, d_segments, stream))


# ---- the reference's four free functions, on one global renderer (pathtrace.cu:88-111) ----------
_state = {"renderer": None, "gui": None}


def InitDataContainer(imGuiData):
    _state["gui"] = imGuiData


def pathtraceInit(scene):
    pathtraceFree()
    _state["renderer"] = Renderer(scene)


def pathtraceFree():
    r = _state["renderer"]
    if r is not None:                     # tolerates being called before the first Init (main.cpp:144)
        r.close()
    _state["renderer"] = None


def pathtrace(pbo, frame, iteration):
    """One sample per pixel.  ``pbo``: device pointer (int) of a W*H uchar4 buffer or None;
    ``frame`` is unused, as in the reference."""
    r = _state["renderer"]
    if r is None:
        raise RuntimeError("pathtrace() before pathtraceInit()")
    gui = _state["gui"] or GuiDataContainer()
    r.set_options(gui)
    r.render(iteration, 1, 1, sync=False)
    if pbo:
        r.tonemap_u8(iteration, pbo)
    r.read_sum(r.scene.state_image)       # pathtrace.cu:524: the SUM image goes back every call
    gui.TracedDepth = r.stats()["traced_depth"]
