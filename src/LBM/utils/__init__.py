from lbm_ferrofluid_b200.types import CellType, KBCType  # noqa: F401
from lbm_ferrofluid_b200.grid import (  # noqa: F401
    get_staggered, get_staggered_x, get_staggered_y, get_staggered_z)
from lbm_ferrofluid_b200.io import *  # noqa: F401,F403
