from lbm_ferrofluid_b200.collision import (  # noqa: F401
    LBMCollision2d, LBMCollision3d, LBMCollisionMRT2d, LBMCollisionSC2d, LBMCollisionHCZ2d, LBMCollisionHCZ3d)
