"""lbm_ferrofluid_b200 -- B200-native (sm_100a) LBM-Ferrofluid solver path.

Host-side mirror of the reference's ``src/LBM`` operator classes on top of the
hand-written CUDA kernels in ``csrc/`` (C ABI: ``include/lbmff.h``).  PyTorch is
used for device memory, streams and ``torch.distributed`` only.  CUDA tensors
only: there is no CPU fallback.
"""
from .types import CellType, KBCType  # noqa: F401
from .grid import get_staggered, get_staggered_x, get_staggered_y, get_staggered_z  # noqa: F401
from .propagation import LBMPropagation2d, LBMPropagation3d  # noqa: F401
from .macro_compute import LBMMacroCompute2d, LBMMacroCompute3d  # noqa: F401
from .collision import (  # noqa: F401
    LBMCollision2d, LBMCollision3d, LBMCollisionMRT2d, LBMCollisionSC2d, LBMCollisionHCZ2d, LBMCollisionHCZ3d)
from .magnetic import LBMMagnetic2d, LBMMagnetic3d  # noqa: F401
from .simulation import SimulationParameters, SimulationRunner  # noqa: F401

__version__ = "0.1.0"
