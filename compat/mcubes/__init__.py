"""`mcubes.marching_cubes` for the reference's demos when PyMCubes is absent: backed by the CUDA iso-surface
extraction of this repo (lbm_ferrofluid_b200/isosurface.py, csrc/iso_surface.cu).  Same call shape as PyMCubes
(numpy volume in, (vertices float64 [V,3], triangles int64 [F,3]) out, array-index coordinates); vertices are not
welded and the triangulation of a cube is this repo's derived case table, not PyMCubes' (see compat/README.md)."""
import numpy as np


def marching_cubes(volume, isovalue):
    import torch
    from lbm_ferrofluid_b200.isosurface import marching_cubes as _mc

    vol = torch.as_tensor(np.ascontiguousarray(volume, dtype=np.float32))
    if not torch.cuda.is_available():
        raise RuntimeError("mcubes (compat): lbm_ferrofluid_b200 runs hand-written CUDA kernels only (no CPU fallback)")
    verts, faces = _mc(vol.cuda(), float(isovalue))
    return verts.double().cpu().numpy(), faces.long().cpu().numpy()
