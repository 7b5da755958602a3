"""the vectorised initialisers reproduce the demos' O(N) python loops EXACTLY.

tests/golden/init_demos.npz holds the fields produced by executing the reference demos' own ``main`` up to their
first ``get_feq_`` call (oracle/make_golden.py::demo_initial_fields: demo_3d_LBM_two_droplets.py:130-151,
demo_2d_LBM_two_droplets.py:117-137, demo_2d_LBM.py:85-98, demo_2d_LBM_KBC.py) on grids with odd extents and extents
not divisible by 8, where `<` vs `<=` and float vs integer centres differ.  The initialisers are plain torch
expressions (initialisation, not the solver path), so the comparison runs on the CPU here and on the device in the
`gpu` variant."""
import numpy as np
import pytest
import torch

from lbm_ferrofluid_b200 import initializers as I
from tests.util import load_golden

G = load_golden("init_demos")


def _check(device):
    for tag in ("drop3d_a", "drop3d_b", "drop2d_a", "drop2d_b"):
        res = tuple(int(v) for v in G[tag + "_res"])
        flags = I.shell_flags(res, device=device)
        rho = I.two_droplets_rho(res, flags)
        assert torch.equal(rho.cpu(), G[tag + "_rho"]), tag
        assert not G[tag + "_vel"].any()
    for tag in ("shear_a", "shear_b"):
        res = tuple(int(v) for v in G[tag + "_res"])
        vel = I.double_shear_layer(res, device=device)
        assert torch.equal(vel.cpu(), G[tag + "_vel"]), tag
        assert torch.equal(G[tag + "_rho"], torch.full_like(G[tag + "_rho"], 0.265))


def test_initializers_equal_the_demo_loops_cpu():
    _check("cpu")


@pytest.mark.gpu
def test_initializers_equal_the_demo_loops_gpu():
    _check("cuda")


def test_two_droplets_slab_pieces_equal_the_global_field():
    """z-slab placement (global_depth / z_offset, periodic wrap of the halo planes) cuts the same global field."""
    res = (16, 10, 26)
    flags = I.shell_flags(res, device="cpu")
    full = I.two_droplets_rho(res, flags)
    for zoff, n in ((-2, 12), (6, 12)):
        planes = (torch.arange(n) + zoff) % res[0]
        part = I.two_droplets_rho((n, res[1], res[2]), flags[:, :, planes], global_depth=res[0], z_offset=zoff)
        assert torch.equal(part, full[:, :, planes])
