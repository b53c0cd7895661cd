"""tomo_tv_b200 -- B200-native (sm_100a) implementation of tomo_TV's iterative-reconstruction hot path.

Public surface (mirrors the reference, /root/reference/tomofusion):
    TomoGPU                      tomofusion.gpu.reconstructor.TomoGPU
    tomoengine                   tomofusion.gpu.utils.tomoengine (pybind11 class)
    ChemicalTomo                 tomofusion.chemistry.reconstructor.ChemicalTomo
    multimodal                   tomofusion.chemistry.utils.multimodal (pybind11 class)
    multigpuengine               tomofusion.gpu.utils.multigpuengine (one rank per GPU under torchrun)
    multigpufusion               tomofusion.chemistry.utils.multigpufusion (likewise)
    pytvlib.initialize_algorithm / run / wbp_filters / sart_orders
    determine_gpu_config, device_count      tomofusion/__init__.py

The CUDA library (tomo_tv_b200/_lib/libtomo_tv_b200.so, built by `python -m tomo_tv_b200._build`)
is the only compute path: there is no CPU fallback and no ASTRA dependency.
"""
from .engine import tomoengine, device_count, NativeError  # noqa: F401
from . import pytvlib  # noqa: F401
from .reconstructor import TomoGPU  # noqa: F401
from .fusion import multimodal, multigpufusion  # noqa: F401
from .chemistry import ChemicalTomo  # noqa: F401
from .sharded import multigpuengine  # noqa: F401

__version__ = "0.1.0"


def determine_gpu_config(gpu_id: int = -1):
    """Same decision table as the reference (tomofusion/__init__.py:21-34)."""
    ngpus = device_count()
    if ngpus == 0:
        raise ValueError('An Nvidia GPU is Needed for this package!')
    if gpu_id < 0 and ngpus > 1:
        return 'multigpu'
    return 'singleconfig'
