from lbm_ferrofluid_b200.magnetic import LBMMagnetic2d, LBMMagnetic3d  # noqa: F401
