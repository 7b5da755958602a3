from lbm_ferrofluid_b200.macro_compute import LBMMacroCompute2d, LBMMacroCompute3d  # noqa: F401
