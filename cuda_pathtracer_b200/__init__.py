"""cuda_pathtracer_b200 -- B200-native (sm_100a) scan/stream-compaction library and wavefront
path tracer behind the C ABI in ``include/``.  This package is only the thin Python binding used
by the tests and bench.py; the product is the CUDA code in ``csrc/`` and the C++ host layer.
"""
from ._lib import B200Error, load, lib_path, declared_symbols  # noqa: F401

__all__ = ["B200Error", "load", "lib_path", "declared_symbols", "stream_compaction"]
