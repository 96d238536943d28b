"""Python mirror of the reference's ``StreamCompaction::{Naive,Efficient,Thrust}`` interface
(reference: stream_compaction/stream_compaction/{naive,efficient,thrust}.h) over the C ABI of
include/b200_sc.h.  Same names, argument meaning and timing convention (``timer()`` reports the
kernels only).  ``StreamCompaction::CPU`` is the reference's host oracle and is deliberately NOT
here (it lives under oracle/, test infrastructure).
"""
import ctypes
import numpy as np

from ._lib import B200Error, load

LOOKBACK, EFFICIENT, NAIVE = 0, 1, 2
LOOKBACK_TILE, LOOKBACK_RING, LOOKBACK_WAVE = 3, 4, 5     # the kernels LOOKBACK picks between, by name (include/b200_sc.h)

_c_int_p = ctypes.POINTER(ctypes.c_int)


def _lib():
    lib = load("libb200sc.so")
    if not getattr(lib, "_b200_typed", False):
        lib.b200_scan_i32.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.POINTER(ctypes.c_float)]
        lib.b200_compact_i32.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, _c_int_p,
                                         ctypes.POINTER(ctypes.c_float)]
        lib.b200_sc_ctx_create.argtypes = [ctypes.POINTER(ctypes.c_void_p)]
        lib.b200_sc_ctx_destroy.argtypes = [ctypes.c_void_p]
        lib.b200_scan_i32_dev.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p]
        lib.b200_compact_i32_dev.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.b200_reduce_i32_dev.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p]
        lib.b200_sc_last_launches.argtypes = [ctypes.c_void_p]
        lib.b200_sc_ctx_reserve.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.b200_sc_ctx_set_ring_min_n.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.b200_scan_i32_noalloc.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        lib.b200_error_string.argtypes = [ctypes.c_int]
        lib.b200_error_string.restype = ctypes.c_char_p
        lib._b200_typed = True
    return lib


def check(code):
    if code != 0:
        raise B200Error(code, _lib().b200_error_string(code).decode())


def _ptr(a):
    """address of a numpy array (host) or torch tensor (host or device)"""
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()


class _Timer:
    """PerformanceTimer look-alike (common.h:46-130): only the GPU reading exists here."""

    def __init__(self):
        self._ms = 0.0

    def getGpuElapsedTimeForPreviousOperation(self):
        return self._ms


class _ScanNamespace:
    def __init__(self, variant):
        self._variant = variant
        self._timer = _Timer()

    def timer(self):
        return self._timer

    def scan(self, n, odata, idata):
        """void scan(int n, int* odata, const int* idata): exclusive prefix sum, int32 wrap."""
        ms = ctypes.c_float(0)
        check(_lib().b200_scan_i32(int(n), _ptr(odata), _ptr(idata), self._variant, ctypes.byref(ms)))
        self._timer._ms = ms.value


class _EfficientNamespace(_ScanNamespace):
    def compact(self, n, odata, idata):
        """int compact(int n, int* odata, const int* idata): drop zeros, keep order, return count."""
        ms = ctypes.c_float(0)
        count = ctypes.c_int(0)
        check(_lib().b200_compact_i32(int(n), _ptr(odata), _ptr(idata), ctypes.byref(count),
                                      ctypes.byref(ms)))
        self._timer._ms = ms.value
        return count.value


Naive = _ScanNamespace(NAIVE)
Efficient = _EfficientNamespace(EFFICIENT)
Thrust = _ScanNamespace(LOOKBACK)      # the single-pass look-back scan; no Thrust inside
SinglePass = _EfficientNamespace(LOOKBACK)
SinglePassTile = _ScanNamespace(LOOKBACK_TILE)   # one tile per CTA
SinglePassWave = _ScanNamespace(LOOKBACK_WAVE)   # one wave of CTAs (small arrays)
SinglePassRing = _ScanNamespace(LOOKBACK_RING)   # persistent TMA ring (what LOOKBACK runs on large aligned arrays)


NOALLOC_MAX_N = 1 << 26
E_TOOLARGE = 10005


def scan_noalloc(n, d_out, d_in, stream=0):
    """b200_scan_i32_noalloc: the look-back scan on device tensors with its scratch in a static device array of the
    library -- allocates nothing; n <= NOALLOC_MAX_N (no counterpart in the reference, see include/b200_sc.h)"""
    check(_lib().b200_scan_i32_noalloc(int(n), _ptr(d_out), _ptr(d_in), stream))


class Context:
    """Device-resident, stream-ordered entry points (b200_*_dev) for torch CUDA tensors."""

    def __init__(self):
        self._h = ctypes.c_void_p()
        check(_lib().b200_sc_ctx_create(ctypes.byref(self._h)))

    def close(self):
        if self._h:
            _lib().b200_sc_ctx_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def scan(self, n, d_out, d_in, variant=LOOKBACK, d_seed=None, d_total=None, stream=0):
        check(_lib().b200_scan_i32_dev(self._h, int(n), _ptr(d_out), _ptr(d_in), variant,
                                       _ptr(d_seed) if d_seed is not None else None,
                                       _ptr(d_total) if d_total is not None else None, stream))

    def compact(self, n, d_out, d_in, d_count, stream=0):
        check(_lib().b200_compact_i32_dev(self._h, int(n), _ptr(d_out), _ptr(d_in), _ptr(d_count), stream))

    def reduce(self, n, d_in, d_total, stream=0):
        check(_lib().b200_reduce_i32_dev(self._h, int(n), _ptr(d_in), _ptr(d_total), stream))

    def reserve(self, n):
        check(_lib().b200_sc_ctx_reserve(self._h, int(n)))

    def set_ring_min_n(self, min_n):
        """from how many elements on the TMA-ring kernels run (0: library default, 1: always, 2**31 - 1: never)"""
        check(_lib().b200_sc_ctx_set_ring_min_n(self._h, int(min_n)))

    def last_launches(self):
        return _lib().b200_sc_last_launches(self._h)
