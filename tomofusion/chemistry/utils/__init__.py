"""tomofusion.chemistry.utils: `multimodal`, `multigpufusion` (pybind11 extension modules in the reference,
chemistry/utils/__init__.py:1) and `fusion_helper`."""
from . import fusion_helper, multigpufusion as _mgf  # noqa: F401
from .multimodal import *  # noqa: F401,F403

multigpufusion = _mgf           # reference spelling: multigpufusion.multigpufusion(...) (chemistry/reconstructor.py:59)
