"""ctypes binding of the C ABI declared in include/tomo_tv_b200.h.

There is deliberately no fallback: if the CUDA library is missing or no GPU is usable, importing
is fine (so CPU-only tooling can inspect the package) but every engine call raises.
"""
import ctypes as C
import os

from . import _build

_lib = None


class NativeError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        path = _build.LIB
        if not os.path.exists(path):
            raise NativeError(
                f"{path} not found: build it with `python -m tomo_tv_b200._build` "
                "(tomo_tv_b200 has no CPU fallback)")
        L = C.CDLL(path)
        L.ttv_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise NativeError(lib().ttv_last_error().decode("utf-8", "replace"))


def exported_symbols():
    """Names declared in include/tomo_tv_b200.h (used by the CPU-side ABI test)."""
    import re
    hdr = os.path.join(os.path.dirname(_build.HERE), "include", "tomo_tv_b200.h")
    txt = open(hdr).read()
    return sorted(set(re.findall(r"\b(ttv_[a-zA-Z0-9_]+)\s*\(", txt)))
