"""Module `multigpuengine` (gpu/utils/multigpuengine.cpp:385-421): class multigpuengine, one rank per GPU under torchrun."""
import sys as _sys
import types as _types

from tomo_tv_b200.sharded import multigpuengine  # noqa: F401


class _CallableModule(_types.ModuleType):
    def __call__(self, *a, **k):
        return multigpuengine(*a, **k)


_sys.modules[__name__].__class__ = _CallableModule
