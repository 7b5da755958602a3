"""`renderutils.SoftRenderer` for the reference's demos: the host mirror of the reference class on top of this repo's
tiled CUDA soft rasteriser (lbm_ferrofluid_b200/render.py, csrc/soft_raster.cu).  Forward (rendering) only."""
from lbm_ferrofluid_b200.render import SoftRenderer, soft_rasterize  # noqa: F401
