"""chemistry/utils/fusion_helper.py of the reference: periodic table and summation-matrix helpers."""
from tomo_tv_b200.chemistry import get_periodic_table, summation_weights, create_weighted_summation_matrix  # noqa: F401
