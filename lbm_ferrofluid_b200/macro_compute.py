"""Macroscopic moments and EOS (reference: src/LBM/LBM_macro_compute/LBM_macro_compute_{2,3}d.py)."""
from __future__ import annotations

from typing import List

import torch

from . import _host as H
from .lattice import lattice_tensors


class _MacroCompute:
    _dim = 0

    def __init__(self, Q=None, tau: float = 1.0, density_liquid: float = 0.265, density_gas: float = 0.038,
                 rho_liquid: float = 0.265, rho_gas: float = 0.038, device=torch.device("cpu"),
                 dtype=torch.float32, *args, **kwargs):
        self._Q = Q if Q is not None else (9 if self._dim == 2 else 19)
        self._tau = tau
        self._density_liquid, self._density_gas = density_liquid, density_gas
        self._rho_liquid, self._rho_gas = rho_liquid, rho_gas
        self.device, self.dtype = device, dtype
        self._weight, self._e = lattice_tensors(self._dim, device, dtype)

    def get_pressure(self, dx: float, dt: float, density: torch.Tensor) -> torch.Tensor:
        d = H.need_cuda(density, "density")
        out = torch.empty_like(d)
        H.launch("lbmff_equation_of_states", d.device, H.ptr(d), H.ptr(out), d.numel(), float(dx), float(dt), 0)
        return out

    def macro_compute(self, dx: float, dt: float, f: torch.Tensor, rho: torch.Tensor, vel: torch.Tensor,
                      flags: torch.Tensor, g: torch.Tensor = None, pressure: torch.Tensor = None,
                      density: torch.Tensor = None) -> List[torch.Tensor]:
        """rho = sum f, vel = sum f e c / rho, old values kept on OBSTACLE cells; with ``density`` also
        returns the affine rho -> density map (LBM_macro_compute_3d.py:73-120)."""
        f = H.need_cuda(f, "f")
        rho, vel = H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        dim, B, Q, D, Hh, W = H.dims(f)
        rho_new, vel_new = torch.empty_like(rho), torch.empty_like(vel)
        dens = torch.empty_like(rho) if density is not None else None
        p = H.fluid_params(self, dx, dt)
        H.launch("lbmff_macro_compute_" + H.suffix(dim), f.device, H.ptr(f), H.ptr(rho), H.ptr(vel), H.ptr(flags),
                 H.ptr(rho_new), H.ptr(vel_new), H.ptr(dens), p, *H.sizes(dim, B, D, Hh, W))
        return [rho_new, vel_new, dens] if density is not None else [rho_new, vel_new]


class LBMMacroCompute2d(_MacroCompute):
    rank = 2
    _dim = 2

    def get_vort(self, vel: torch.Tensor, dx: float) -> torch.Tensor:
        v = H.need_cuda(vel, "vel")
        dim, B, C, D, Hh, W = H.dims(v)
        out = torch.empty((B, 1, Hh, W), dtype=v.dtype, device=v.device)
        H.launch("lbmff_get_vort_2d", v.device, H.ptr(v), H.ptr(out), float(dx), B, Hh, W)
        return out


class LBMMacroCompute3d(_MacroCompute):
    rank = 3
    _dim = 3
