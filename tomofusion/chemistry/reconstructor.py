"""tomofusion.chemistry.reconstructor.ChemicalTomo (/root/reference/tomofusion/chemistry/reconstructor.py:13)."""
from tomo_tv_b200.chemistry import ChemicalTomo  # noqa: F401
