"""Import-path shim: the reference's demos import ``src.LBM.*`` (demo_3d_LBM_Rosensweig_instability.py:11-21).
Everything here re-exports the B200-native implementation in ``lbm_ferrofluid_b200``."""
