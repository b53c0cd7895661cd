"""Import-path drop-in for the reference package: `tomofusion.*` names resolve to tomo_tv_b200.

The reference is imported as `from tomofusion.gpu.reconstructor import TomoGPU`,
`from tomofusion.gpu.utils import tomoengine, multigpuengine`,
`from tomofusion.chemistry.reconstructor import ChemicalTomo`,
`from tomofusion.chemistry.utils import multimodal, multigpufusion` and `from tomofusion import pytvlib`
(/root/reference/tomofusion/gpu/reconstructor.py:1-3, chemistry/reconstructor.py:1-2, gpu/utils/__init__.py:2,
chemistry/utils/__init__.py:1).  This package re-exports the B200-native classes under exactly those paths, so
user scripts written against the reference run unchanged with this repository on PYTHONPATH.  No compute lives here.
"""
from tomo_tv_b200 import __version__, determine_gpu_config, pytvlib  # noqa: F401
from tomo_tv_b200 import device_count as _device_count

__author__ = "tomo_tv_b200"


def device_count():
    """List of visible GPU ordinals, like the reference (tomofusion/__init__.py:10-18: `list(range(n))`)."""
    return list(range(_device_count()))
