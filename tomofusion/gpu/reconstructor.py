"""tomofusion.gpu.reconstructor.TomoGPU (/root/reference/tomofusion/gpu/reconstructor.py:12) -> tomo_tv_b200.reconstructor."""
from tomo_tv_b200.reconstructor import TomoGPU, pack_tilt_series  # noqa: F401

reconstructor = TomoGPU        # the name tomofusion/gpu/simulator.py:2 imports
