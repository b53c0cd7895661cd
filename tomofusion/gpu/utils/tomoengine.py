"""Module `tomoengine` (the reference's pybind11 extension, gpu/utils/tomoengine.cpp:487-534): class tomoengine."""
import sys as _sys
import types as _types

from tomo_tv_b200.engine import tomoengine  # noqa: F401


class _CallableModule(_types.ModuleType):
    def __call__(self, *a, **k):
        return tomoengine(*a, **k)


_sys.modules[__name__].__class__ = _CallableModule
