"""tomofusion.gpu.utils: the pybind11 extension modules of the reference (`from .tomoengine import *`,
gpu/utils/__init__.py:2) as modules and classes."""
from . import tomoengine as _te, multigpuengine as _mg
from .tomoengine import *  # noqa: F401,F403
from tomo_tv_b200 import pytvlib  # noqa: F401   (tomofusion/gpu/simulator.py:1 imports it from here)

# `from tomofusion.gpu.utils import tomoengine, multigpuengine` yields modules in the reference (tomoengine.tomoengine is
# the class); the same spelling works here, and the modules are also callable as the class itself for convenience
tomoengine = _te
multigpuengine = _mg
