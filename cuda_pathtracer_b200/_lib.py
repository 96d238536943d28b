"""ctypes loader for the in-tree C-ABI libraries (include/*.h).

The product has no CPU fallback: if a library is missing or no CUDA device is usable, calls
raise.  Nothing here imports anything under ``oracle/``.
"""
import ctypes
import os
import re

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(_PKG_DIR)
_cache = {}


class B200Error(RuntimeError):
    """A C-ABI entry point returned a non-zero code."""

    def __init__(self, code, text):
        super().__init__(f"b200 error {code}: {text}")
        self.code = code


def lib_path(name):
    return os.path.join(_PKG_DIR, name)


def load(name):
    """dlopen ``cuda_pathtracer_b200/<name>``; raises if it has not been built (make lib)."""
    if name in _cache:
        return _cache[name]
    path = lib_path(name)
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `make lib` (or __graft_entry__.build()). "
            "There is no CPU or PyTorch fallback for this path.")
    lib = ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL)
    _cache[name] = lib
    return lib


def declared_symbols(header):
    """Function names declared in include/<header> (every ``name(`` at the start of a
    declaration that begins with a C type); used by the ABI test."""
    text = open(os.path.join(_REPO, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    names = re.findall(r"^\s*(?:const\s+)?(?:int|void|char|float|double|unsigned|size_t)\s*\**\s*"
                       r"(b200\w*)\s*\(", text, flags=re.M)
    return sorted(set(names))
