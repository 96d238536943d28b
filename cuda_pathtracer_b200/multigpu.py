"""ctypes mirror of include/b200_mg.h: the two hot paths over the K GPUs of one box in ONE process (one host thread
per device, NCCL over NVLink).  Used by the tests; `cuda_pathtracer_b200.distributed` is the one-process-per-GPU
(torch.distributed) form of the same partitioning.  Nothing here imports anything under ``oracle/``."""
import ctypes

import numpy as np

from ._lib import B200Error, load
from .pathtracer import _OptionsC, _SceneC, _SceneExtC, _StatsC, _TextureC


def _lib():
    load("libb200sc.so")
    load("libb200pt.so")
    lib = load("libb200mg.so")
    if not getattr(lib, "_b200mg_typed", False):
        vp, ci, fp = ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_float)
        lib.b200mg_create.argtypes = [ci, ctypes.POINTER(ci), ctypes.POINTER(vp)]
        lib.b200mg_destroy.argtypes = [vp]
        lib.b200mg_device_count.argtypes = [vp]
        lib.b200mg_pt_create.argtypes = [vp, ctypes.POINTER(_SceneC), ctypes.POINTER(_SceneExtC), ctypes.POINTER(vp)]
        lib.b200mg_pt_destroy.argtypes = [vp]
        lib.b200mg_pt_set_options.argtypes = [vp, ctypes.POINTER(_OptionsC)]
        lib.b200mg_pt_reset.argtypes = [vp]
        lib.b200mg_pt_write_sum.argtypes = [vp, vp]
        lib.b200mg_pt_render.argtypes = [vp, ci, ci]
        lib.b200mg_pt_reduce_read_sum.argtypes = [vp, vp, fp]
        lib.b200mg_pt_preview_u8.argtypes = [vp, ci, vp]
        lib.b200mg_pt_get_stats.argtypes = [vp, ctypes.POINTER(_StatsC)]
        lib.b200mg_scan_i32.argtypes = [vp, ci, vp, vp, fp, fp]
        lib.b200mg_compact_i32.argtypes = [vp, ci, vp, vp, ctypes.POINTER(ci), fp, fp]
        lib.b200mg_error_string.argtypes = [ci]
        lib.b200mg_error_string.restype = ctypes.c_char_p
        lib._b200mg_typed = True
    return lib


def check(code):
    if code != 0:
        raise B200Error(code, _lib().b200mg_error_string(code).decode())


class DeviceGroup:
    """b200mg: NCCL communicators + one stream and one scan context per device"""

    def __init__(self, n_devices=0, devices=None):
        # libb200mg.so binds NCCL at run time (dlopen "libnccl.so.2").  In a process that will also use torch, torch's own
        # NCCL must be the one that gets loaded: import it first, so that the dlopen finds it by its SONAME.
        try:
            import torch  # noqa: F401
        except ImportError:
            pass
        self._h = ctypes.c_void_p()
        arr = (ctypes.c_int * len(devices))(*devices) if devices else None
        check(_lib().b200mg_create(int(len(devices) if devices else n_devices), arr, ctypes.byref(self._h)))

    @property
    def device_count(self):
        return _lib().b200mg_device_count(self._h)

    def close(self):
        if self._h:
            _lib().b200mg_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def scan(self, n, odata, idata):
        """exclusive scan of n host ints over the group's devices; returns (kernel_ms, all_gather_ms)"""
        ms, ex = ctypes.c_float(0), ctypes.c_float(0)
        check(_lib().b200mg_scan_i32(self._h, int(n), odata.ctypes.data, idata.ctypes.data, ctypes.byref(ms), ctypes.byref(ex)))
        return ms.value, ex.value

    def compact(self, n, odata, idata):
        """order-preserving removal of zeros; returns (count, kernel_ms, all_gather_ms)"""
        ms, ex, cnt = ctypes.c_float(0), ctypes.c_float(0), ctypes.c_int(0)
        check(_lib().b200mg_compact_i32(self._h, int(n), odata.ctypes.data, idata.ctypes.data, ctypes.byref(cnt),
                                        ctypes.byref(ms), ctypes.byref(ex)))
        return cnt.value, ms.value, ex.value


class GroupRenderer:
    """b200mg_pt: a renderer replicated over the group, iterations split with stride K"""

    def __init__(self, group, scene):
        self.group, self.scene = group, scene
        sc = _SceneC(scene.camera, scene.traceDepth, len(scene.geoms), scene.geoms, len(scene.materials), scene.materials)
        tex = (_TextureC * len(scene.textures))(*[_TextureC(t.width, t.height, t.components, t.pixels.ctypes.data)
                                                  for t in scene.textures])
        ext = _SceneExtC(len(scene.triangles), scene.triangles, len(scene.bvh), scene.bvh, len(scene.textures), tex)
        self._h = ctypes.c_void_p()
        check(_lib().b200mg_pt_create(group._h, ctypes.byref(sc), ctypes.byref(ext), ctypes.byref(self._h)))

    def close(self):
        if self._h:
            _lib().b200mg_pt_destroy(self._h)
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
        check(_lib().b200mg_pt_set_options(self._h, ctypes.byref(o)))

    def reset(self):
        check(_lib().b200mg_pt_reset(self._h))

    def render(self, first_iter, n_iters):
        check(_lib().b200mg_pt_render(self._h, int(first_iter), int(n_iters)))

    def reduce_read_sum(self):
        out = np.empty((self.scene.height, self.scene.width, 3), dtype=np.float32)
        ms = ctypes.c_float(0)
        check(_lib().b200mg_pt_reduce_read_sum(self._h, out.ctypes.data, ctypes.byref(ms)))
        return out, ms.value

    def preview_u8(self, iter_divisor):
        out = np.empty((self.scene.height, self.scene.width, 4), dtype=np.uint8)
        check(_lib().b200mg_pt_preview_u8(self._h, int(iter_divisor), out.ctypes.data))
        return out

    def stats(self):
        s = _StatsC()
        check(_lib().b200mg_pt_get_stats(self._h, ctypes.byref(s)))
        return {"path_segments": s.path_segments, "iterations": s.iterations, "kernel_launches": s.kernel_launches,
                "traced_depth": s.traced_depth}
