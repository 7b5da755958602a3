from lbm_ferrofluid_b200.propagation import LBMPropagation2d, LBMPropagation3d  # noqa: F401
