"""Module `multigpufusion` (chemistry/utils/multigpufusion.cpp:459-475): class multigpufusion."""
from tomo_tv_b200.fusion import multigpufusion  # noqa: F401
