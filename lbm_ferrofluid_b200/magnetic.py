"""Magnetic Poisson-lattice operator (reference: src/LBM/LBM_magnetic/LBM_magnetic_{2,3}d.py)."""
from __future__ import annotations

import math
from typing import List

import torch

from . import _host as H
from .lattice import lattice_tensors


class _Magnetic:
    _dim = 0

    def __init__(self, Q=None, tau: float = 1.0, density_liquid: float = 0.265, density_gas: float = 0.038,
                 rho_liquid: float = 0.265, rho_gas: float = 0.038, kappa: float = 0.08, tau_f: float = 0.7,
                 tau_g: float = 0.7, contact_angle=math.pi / 2.0, device=torch.device("cpu"),
                 dtype=torch.float32, *args, **kwargs):
        self._Q = Q if Q is not None else (9 if self._dim == 2 else 19)
        self._tau = tau
        self._density_liquid, self._density_gas = density_liquid, density_gas
        self._rho_liquid, self._rho_gas = rho_liquid, rho_gas
        self._kappa, self._tau_f, self._tau_g = kappa, tau_f, tau_g
        self._contact_angle = contact_angle
        self.device, self.dtype = device, dtype
        self._weight, self._e = lattice_tensors(self._dim, device, dtype)

    def get_heq_(self, psi: torch.Tensor) -> torch.Tensor:
        """heq_i = w_i psi, heq_0 -= psi (Chai & Shi; LBM_magnetic_3d.py:108-117).  Helper kept for API
        compatibility; get_H_int evaluates it inside its kernel."""
        heq = psi * self._weight
        heq[:, 0:1, ...] = heq[:, 0:1, ...] - psi
        return heq

    def smooth_phi(self, phi: torch.Tensor, eps: float) -> torch.Tensor:
        x = H.need_cuda(phi, "phi")
        out = torch.empty_like(x)
        H.launch("lbmff_smooth_phi", x.device, H.ptr(x), H.ptr(out), x.numel(), float(eps))
        return out

    def get_H_int(self, dt: float, dx: float, phi: torch.Tensor, flags: torch.Tensor,
                  H_ext_mac: List[torch.Tensor], h: torch.Tensor) -> List[torch.Tensor]:
        """One collision sweep of the Poisson lattice h and H_int = -grad(psi)
        (LBM_magnetic_3d.py:125-188).  Returns [H_int, new_h]."""
        h, phi = H.need_cuda(h, "h"), H.need_cuda(phi, "phi")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        faces = [H.need_cuda(t, "H_ext_mac") for t in H_ext_mac]
        dim, B, Q, D, Hh, W = H.dims(h)
        if len(faces) != dim:
            raise RuntimeError(f"H_ext_mac must hold {dim} face tensors")
        expect = [(B, 1, *([D] if dim == 3 else []), Hh, W + 1), (B, 1, *([D] if dim == 3 else []), Hh + 1, W)]
        if dim == 3:
            expect.append((B, 1, D + 1, Hh, W))
        for t, shp in zip(faces, expect):
            if tuple(t.shape) != shp:
                raise RuntimeError(f"H_ext_mac face has shape {tuple(t.shape)}, expected {shp}")
        H_int = torch.empty((B, dim, *h.shape[2:]), dtype=h.dtype, device=h.device)
        h_new = torch.empty_like(h)
        psi = torch.empty((B, 1, *h.shape[2:]), dtype=h.dtype, device=h.device)
        H.launch("lbmff_get_H_int_" + H.suffix(dim), h.device, H.ptr(phi), H.ptr(flags),
                 *[H.ptr(t) for t in faces], H.ptr(h), float(self._tau), float(dx), float(dt), H.ptr(H_int),
                 H.ptr(h_new), H.ptr(psi), *H.sizes(dim, B, D, Hh, W))
        return [H_int, h_new]


class LBMMagnetic2d(_Magnetic):
    rank = 2
    _dim = 2


class LBMMagnetic3d(_Magnetic):
    rank = 3
    _dim = 3
