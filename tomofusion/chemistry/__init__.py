"""tomofusion.chemistry -> tomo_tv_b200 (see tomofusion/__init__.py)."""
