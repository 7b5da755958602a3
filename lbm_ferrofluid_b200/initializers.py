"""Vectorised on-device initial conditions for the demo configurations.

The reference's demos fill their initial fields with Python loops over every cell
(demo_3d_LBM_two_droplets.py:134-149, demo_2d_LBM.py:86-98) -- hours at 1024^3 / 4096^2.  These build the
same fields with a handful of device-side tensor expressions.  Initialisation only: not part of the timed path.
"""
from __future__ import annotations

import math
from typing import Sequence

import torch

from .types import CellType

FLUID, OBSTACLE = int(CellType.FLUID), int(CellType.OBSTACLE)


def shell_flags(res: Sequence[int], open_axes: Sequence[int] = (), device="cuda") -> torch.Tensor:
    """OBSTACLE outer shell, FLUID inside; axes in ``open_axes`` (0 = x, 1 = y, 2 = z) have no wall
    (demo_3d_LBM_Rosensweig_instability.py:117-121)."""
    dim = len(res)
    fl = torch.full((1, 1, *res), OBSTACLE, dtype=torch.uint8, device=device)
    sl = [slice(None), slice(None)] + [slice(1, -1)] * dim
    for a in open_axes:
        sl[len(sl) - 1 - a] = slice(None)
    fl[tuple(sl)] = FLUID
    return fl


def _coords(res, device, z_offset=0):
    grids = torch.meshgrid(*[torch.arange(n, dtype=torch.float32, device=device) for n in res], indexing="ij")
    if len(res) == 3:
        z, y, x = grids
        return x, y, z + float(z_offset)
    y, x = grids
    return x, y, torch.zeros_like(x)


def rosensweig_rho(res, flags, rho_liquid=0.2508, rho_gas=0.02381, frac=0.5, perturb=2.0, noise=0.0, seed=0,
                   global_depth=None, z_offset=0):
    """Liquid below y = frac*H (demo_3d...Rosensweig...py:128-132), interface height perturbed by
    ``perturb`` * sin(4 pi x/W) cos(4 pi z/D) cells so x/z stencil terms are exercised,
    walls at the liquid value.  ``global_depth`` / ``z_offset`` place a z-slab inside a larger domain."""
    dev = flags.device
    x, y, z = _coords(res, dev, z_offset)
    H, W = res[-2], res[-1]
    D = global_depth if global_depth is not None else (res[0] if len(res) == 3 else 1)
    height = frac * H + perturb * torch.sin(4 * math.pi * x / W) * (
        torch.cos(4 * math.pi * z / D) if len(res) == 3 else 1.0)
    rho = torch.where(y < height, rho_liquid, rho_gas).to(torch.float32).reshape(1, 1, *res)
    if noise:
        gen = torch.Generator(device=dev).manual_seed(seed)
        rho = rho * (1.0 + noise * (torch.rand(rho.shape, device=dev, generator=gen) - 0.5))
    rho[flags == OBSTACLE] = rho_liquid
    return rho.contiguous()


def two_droplets_rho(res, flags, rho_liquid=0.2508, rho_gas=0.02381, global_depth=None, z_offset=0):
    """The initial density of demo_3d_LBM_two_droplets.py:130-149 (2-D: demo_2d_LBM_two_droplets.py:117-134), cell
    for cell: two spheres (discs) of radius min(res)//4 around the INTEGER centres (res[0]//2, res[1]//2,
    3*res[2]//8) and (.., 5*res[2]//8), a cell being liquid when its squared integer distance is <= radius^2; walls
    at the liquid value (:151).  Integer arithmetic throughout, so the field equals the demo's loop exactly
    (tests/test_initializers.py).  ``global_depth`` / ``z_offset`` place a z-slab inside a larger periodic domain."""
    dev = flags.device
    three = len(res) == 3
    H, W = res[-2], res[-1]
    D = global_depth if global_depth is not None else (res[0] if three else 1)
    full = (D, H, W) if three else (H, W)
    r = min(full) // 4
    ax = [torch.arange(n, dtype=torch.int64, device=dev) for n in res]
    if three:
        z = ((ax[0] + int(z_offset)) % D).reshape(-1, 1, 1)
        y, x = ax[1].reshape(1, -1, 1), ax[2].reshape(1, 1, -1)
    else:
        y, x = ax[0].reshape(-1, 1), ax[1].reshape(1, -1)
    liquid = torch.zeros(tuple(res), dtype=torch.bool, device=dev)
    for cx in (3 * W // 8, 5 * W // 8):
        d2 = (x - cx) ** 2 + (y - H // 2) ** 2
        if three:
            d2 = d2 + (z - D // 2) ** 2
        liquid |= d2 <= r * r
    rho = torch.where(liquid, rho_liquid, rho_gas).to(torch.float32).reshape(1, 1, *res)
    rho[flags == OBSTACLE] = rho_liquid
    return rho.contiguous()


def double_shear_layer(res, device="cuda", u0=0.2, sigma=0.05, kappa=80.0):
    """Velocity field of demo_2d_LBM.py:85-98 (same in demo_2d_LBM_KBC.py).  The demo evaluates each value with
    ``math`` in double precision and stores it into the fp32 tensor; here the same expressions, in the same order,
    are evaluated in float64 on the device and rounded once."""
    H, W = res
    j = torch.arange(H, dtype=torch.float64, device=device).reshape(H, 1).expand(H, W)
    i = torch.arange(W, dtype=torch.float64, device=device).reshape(1, W).expand(H, W)
    ux = torch.where(j <= H / 2.0, u0 * torch.tanh(kappa * (1.0 * j / H - 0.25)),
                     u0 * torch.tanh(kappa * (-1.0 * j / H + 0.75)))
    uy = sigma * u0 * torch.sin(2.0 * math.pi * (1.0 * i / W + 0.25))
    return torch.stack([ux, uy]).to(torch.float32).reshape(1, 2, H, W).contiguous()
