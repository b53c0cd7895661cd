"""Module `multimodal` (chemistry/utils/multimodal.cpp:520-564): class multimodal."""
from tomo_tv_b200.fusion import multimodal  # noqa: F401
