"""Collision operators (reference: src/LBM/LBM_collision/*.py).

Same class names, constructor kwargs, method names and argument meaning as the
reference; every method allocates fresh outputs and launches hand-written
sm_100a kernels through the C ABI (include/lbmff.h).  ``capillary_process``
modifies ``rho`` in place on the shell, exactly like the reference.
"""
from __future__ import annotations

import math
from typing import List

import torch

from . import _host as H
from .lattice import lattice_tensors
from .types import KBCType


class _Collision:
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

    # ---- pointwise -------------------------------------------------------
    def equation_of_states(self, dx: float, dt: float, rho: torch.Tensor) -> torch.Tensor:
        """Carnahan-Starling EOS (LBM_collision_3d.py:105-122)."""
        r = H.need_cuda(rho, "rho")
        out = torch.empty_like(r)
        H.launch("lbmff_equation_of_states", r.device, H.ptr(r), H.ptr(out), r.numel(), float(dx), float(dt), 1)
        return out

    def set_gravity(self, gravity: float):
        """Gravity acts along -y (LBM_collision_3d.py:124-131)."""
        vec = [0.0] * self._dim
        vec[1] = -gravity
        self._gravity_scalar = float(gravity)
        self._gravity = torch.tensor(vec, dtype=torch.float32).reshape(1, self._dim, *([1] * self._dim)).to(
            self.device).to(self.dtype)

    def get_feq_(self, dx: float, dt: float, rho: torch.Tensor, vel: torch.Tensor,
                 force: torch.Tensor = None) -> torch.Tensor:
        """Entropic product-form equilibrium (LBM_collision_3d.py:133-187)."""
        rho, vel = H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        force = H.need_cuda(force, "force") if force is not None else None
        dim, B, C, D, Hh, W = H.dims(vel)
        out = torch.empty((B, self._Q, *vel.shape[2:]), dtype=vel.dtype, device=vel.device)
        H.launch("lbmff_get_feq_" + H.suffix(dim), vel.device, H.ptr(rho), H.ptr(vel), H.ptr(force),
                 float(self._tau), float(dx), float(dt), H.ptr(out), *H.sizes(dim, B, D, Hh, W))
        return out

    def _geq(self, dx, dt, rho, vel, density, pressure, force, feq):
        rho, vel = H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        density, pressure = H.need_cuda(density, "density"), H.need_cuda(pressure, "pressure")
        feq = H.need_cuda(feq, "feq") if feq is not None else None
        force = H.need_cuda(force, "force") if (force is not None and feq is None) else None
        dim, B, C, D, Hh, W = H.dims(vel)
        out = torch.empty((B, self._Q, *vel.shape[2:]), dtype=vel.dtype, device=vel.device)
        H.launch("lbmff_get_geq_" + H.suffix(dim), vel.device, H.ptr(rho), H.ptr(vel), H.ptr(density),
                 H.ptr(pressure), H.ptr(force), H.ptr(feq), float(self._tau), float(dx), float(dt), H.ptr(out),
                 *H.sizes(dim, B, D, Hh, W))
        return out

    # ---- stencils --------------------------------------------------------
    @staticmethod
    def get_grad(input_: torch.Tensor, dx: float, flags: torch.Tensor) -> torch.Tensor:
        """Isotropic gradient with the reference's shell rules (LBM_collision_3d.py:209-279)."""
        if input_.shape[1] != 1:
            raise RuntimeError("To get your grad operation, channel dim has to be 1")
        x = H.need_cuda(input_, "input_")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        dim, B, C, D, Hh, W = H.dims(x)
        out = torch.empty((B, dim, *x.shape[2:]), dtype=x.dtype, device=x.device)
        H.launch(f"lbmff_get_grad_{dim}d", x.device, H.ptr(x), H.ptr(flags), H.ptr(out), float(dx),
                 *H.sizes(dim, B, D, Hh, W))
        return out

    def get_laplacian(self, input_: torch.Tensor, dx: float, flags: torch.Tensor = None) -> torch.Tensor:
        """Isotropic Laplacian, zero on the shell, flags ignored (LBM_collision_3d.py:281-318)."""
        x = H.need_cuda(input_, "input_")
        dim, B, C, D, Hh, W = H.dims(x)
        if C != 1:
            raise RuntimeError("get_laplacian expects a single-channel field")
        out = torch.empty_like(x)
        H.launch(f"lbmff_get_laplacian_{dim}d", x.device, H.ptr(x), H.ptr(out), float(dx),
                 *H.sizes(dim, B, D, Hh, W))
        return out

    # ---- BGK -------------------------------------------------------------
    def _bgk(self, dx, dt, f, rho, vel, flags, force):
        f, rho, vel = H.need_cuda(f, "f"), H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        force = H.need_cuda(force, "force") if force is not None else None
        dim, B, Q, D, Hh, W = H.dims(f)
        out = torch.empty_like(f)
        H.launch("lbmff_collision_bgk_" + H.suffix(dim), f.device, H.ptr(f), H.ptr(rho), H.ptr(vel), H.ptr(flags),
                 H.ptr(force), float(self._tau), float(dx), float(dt), H.ptr(out), *H.sizes(dim, B, D, Hh, W))
        return out


class LBMCollision2d(_Collision):
    rank = 2
    _dim = 2

    def get_geq_(self, dx, dt, rho, density, vel, pressure, force, feq=None):
        """2-D argument order (LBM_collision_2d.py:163-181)."""
        return self._geq(dx, dt, rho, vel, density, pressure, force, feq)

    def collision(self, dx, dt, f, rho, vel, flags, force, KBC_type: int = None) -> torch.Tensor:
        """BGK with velocity-shift forcing; KBC_type accepted and ignored (LBM_collision_2d.py:258-287)."""
        return self._bgk(dx, dt, f, rho, vel, flags, force)


class LBMCollision3d(_Collision):
    rank = 3
    _dim = 3

    def get_geq_(self, dx, dt, rho, vel, density, pressure, force, feq=None):
        """3-D argument order (LBM_collision_3d.py:189-207)."""
        return self._geq(dx, dt, rho, vel, density, pressure, force, feq)

    def collision(self, dx, dt, f, rho, vel, flags, force) -> torch.Tensor:
        """BGK with velocity-shift forcing (LBM_collision_3d.py:320-346)."""
        return self._bgk(dx, dt, f, rho, vel, flags, force)


class LBMCollisionMRT2d(LBMCollision2d):
    """D2Q9 entropic multi-relaxation (KBC) collision (LBM_collision_MRT_2d.py)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.C_mat = torch.zeros((self._Q, 3, 3), dtype=self.dtype, device=self.device)

    def preset_KBC(self, dx: float, dt: float):
        """Raw-moment matrix c^(p+q) ex^p ey^q (LBM_collision_MRT_2d.py:23-60); the kernels build the
        same constants on the fly, the attribute is kept for API compatibility."""
        c = dx / dt
        e = self._e.reshape(self._Q, 2).to(self.dtype)
        C = torch.zeros((self._Q, 3, 3), dtype=self.dtype, device=self.device)
        for p in range(3):
            for q in range(3):
                if (p, q) in ((0, 0), (1, 0), (0, 1), (2, 0), (0, 2), (1, 1), (2, 2), (2, 1), (1, 2)):
                    C[:, p, q] = (c ** (p + q)) * e[:, 0] ** p * e[:, 1] ** q
        self.C_mat = C.reshape(1, self._Q, 3, 3, 1, 1)

    def get_s_by_KBC(self, dx, dt, f, rho, vel, KBC_type: int = None) -> torch.Tensor:
        assert self._Q == 9
        f, rho, vel = H.need_cuda(f, "f"), H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        dim, B, Q, D, Hh, W = H.dims(f)
        out = torch.empty_like(f)
        H.launch("lbmff_get_s_by_kbc_d2q9", f.device, H.ptr(f), H.ptr(rho), H.ptr(vel),
                 1 if KBCType.is_KBC(KBC_type) else 0, float(dx), float(dt), H.ptr(out), B, Hh, W)
        return out

    def _kbc(self, dx, dt, f, rho, vel, flags, force, KBC_type):
        f, rho, vel = H.need_cuda(f, "f"), H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        force = H.need_cuda(force, "force") if force is not None else None
        dim, B, Q, D, Hh, W = H.dims(f)
        out = torch.empty_like(f)
        H.launch("lbmff_collision_kbc_d2q9", f.device, H.ptr(f), H.ptr(rho), H.ptr(vel), H.ptr(flags),
                 H.ptr(force), float(self._tau), 1 if KBCType.is_KBC(KBC_type) else 0, float(dx), float(dt),
                 H.ptr(out), B, Hh, W)
        return out

    def collision(self, dx, dt, f, rho, vel, flags, force, g=None, pressure=None, dfai=None, dprho=None,
                  KBC_type: int = None):
        """Entropic KBC collision on non-OBSTACLE cells (LBM_collision_MRT_2d.py:184-226)."""
        return self._kbc(dx, dt, f, rho, vel, flags, force, KBC_type)


class LBMCollisionSC2d(LBMCollision2d):
    """Shan-Chen pseudopotential multiphase, 2-D (LBM_collision_SC_2d.py)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._density_wall = 0.5 * (self._density_gas + self._density_liquid)

    def calculate_force(self, dx, dt, density, flags, force) -> torch.Tensor:
        d = H.need_cuda(density, "density")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        dim, B, C, D, Hh, W = H.dims(d)
        out = torch.empty((B, 2, Hh, W), dtype=d.dtype, device=d.device)
        H.launch("lbmff_sc_force_d2q9", d.device, H.ptr(d), H.ptr(flags), float(self._density_liquid),
                 float(self._density_gas), float(dx), float(dt), H.ptr(out), B, Hh, W)
        return out

    def collision(self, dx, dt, f, rho, density, vel, flags, force, KBC_type: int = None):
        """SC force then BGK (LBM_collision_SC_2d.py:99-137)."""
        force = self.calculate_force(dx=dx, dt=dt, density=density, flags=flags, force=force)
        return self._bgk(dx, dt, f, rho, vel, flags, force)


class _HCZMixin:
    """He-Chen-Zhang two-distribution multiphase model (LBM_collision_HCZ_{2,3}d.py)."""

    def _cot(self) -> float:
        ca = self._contact_angle
        key = (id(ca), ca._version) if torch.is_tensor(ca) else ("f", float(ca))
        cached = getattr(self, "_cot_cache", None)
        if cached is None or cached[0] != key:
            self._cot_cache = (key, H.cot_of_contact_angle(ca))   # one device sync, then cached
        return self._cot_cache[1]

    def capillary_process(self, rho, vel, flags, force, dt: float = 1.0, dx: float = 1.0, g=None, density=None,
                          pressure=None, H2=None, phi=None):
        """Wetting BC (in place on rho), surface tension, gravity, Kelvin force, g-moments.
        Returns [rho, vel, density, pressure, force, dfai, dprho] (HCZ_3d.py:21-263)."""
        if not hasattr(self, "_gravity_scalar"):
            raise AttributeError(f"'{type(self).__name__}' object has no attribute '_gravity' (call set_gravity)")
        if not rho.is_contiguous():
            raise RuntimeError("capillary_process modifies rho in place: rho must be contiguous")
        rho = H.need_cuda(rho, "rho")
        vel, g = H.need_cuda(vel, "vel"), H.need_cuda(g, "g")
        density, pressure = H.need_cuda(density, "density"), H.need_cuda(pressure, "pressure")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        if (H2 is None) != (phi is None):
            raise RuntimeError("H2 and phi must be given together")
        H2 = H.need_cuda(H2, "H2") if H2 is not None else None
        phi = H.need_cuda(phi, "phi") if phi is not None else None
        dim, B, C, D, Hh, W = H.dims(rho)
        vel_o, dens_o, pres_o = torch.empty_like(vel), torch.empty_like(rho), torch.empty_like(rho)
        force_o, dfai, dprho = torch.empty_like(vel), torch.empty_like(vel), torch.empty_like(vel)
        ws = torch.empty((B, 4, *rho.shape[2:]), dtype=rho.dtype, device=rho.device)
        p = H.fluid_params(self, dx, dt, cot=self._cot(), gravity=self._gravity_scalar)
        H.launch("lbmff_capillary_process_" + H.suffix(dim), rho.device, H.ptr(rho), H.ptr(vel), H.ptr(flags),
                 H.ptr(g), H.ptr(density), H.ptr(pressure), H.ptr(H2), H.ptr(phi), p, H.ptr(vel_o), H.ptr(dens_o),
                 H.ptr(pres_o), H.ptr(force_o), H.ptr(dfai), H.ptr(dprho), H.ptr(ws), *H.sizes(dim, B, D, Hh, W))
        return [rho, vel_o, dens_o, pres_o, force_o, dfai, dprho]

    def smooth_phi(self, phi: torch.Tensor, eps: float) -> torch.Tensor:
        x = H.need_cuda(phi, "phi")
        out = torch.empty_like(x)
        H.launch("lbmff_smooth_phi", x.device, H.ptr(x), H.ptr(out), x.numel(), float(eps))
        return out

    def compute_Gamma(self, dx: float, dt: float, vel: torch.Tensor) -> torch.Tensor:
        v = H.need_cuda(vel, "vel")
        dim, B, C, D, Hh, W = H.dims(v)
        out = torch.empty((B, self._Q, *v.shape[2:]), dtype=v.dtype, device=v.device)
        H.launch("lbmff_compute_gamma_" + H.suffix(dim), v.device, H.ptr(v), float(dx), float(dt), H.ptr(out),
                 *H.sizes(dim, B, D, Hh, W))
        return out

    def _hcz(self, dx, dt, f, rho, vel, density, flags, force, g, pressure, dfai, dprho, kbc_on_g) -> List:
        f, g = H.need_cuda(f, "f"), H.need_cuda(g, "g")
        rho, vel, density = H.need_cuda(rho, "rho"), H.need_cuda(vel, "vel"), H.need_cuda(density, "density")
        force, pressure = H.need_cuda(force, "force"), H.need_cuda(pressure, "pressure")
        dfai, dprho = H.need_cuda(dfai, "dfai"), H.need_cuda(dprho, "dprho")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        dim, B, Q, D, Hh, W = H.dims(f)
        f_o, g_o = torch.empty_like(f), torch.empty_like(g)
        p = H.fluid_params(self, dx, dt)
        args = [H.ptr(f), H.ptr(g), H.ptr(rho), H.ptr(vel), H.ptr(density), H.ptr(flags), H.ptr(force),
                H.ptr(pressure), H.ptr(dfai), H.ptr(dprho), p]
        if dim == 2:
            args.append(int(kbc_on_g))
        H.launch("lbmff_collision_hcz_" + H.suffix(dim), f.device, *args, H.ptr(f_o), H.ptr(g_o),
                 *H.sizes(dim, B, D, Hh, W))
        return [f_o, g_o]


class LBMCollisionHCZ2d(_HCZMixin, LBMCollisionMRT2d):
    rank = 2

    def collision(self, dx, dt, f, rho, vel, density, flags, force, g=None, pressure=None, dfai=None, dprho=None,
                  KBC_type: int = None):
        """f, g HCZ collisions on FLUID; entropic KBC relaxation of g when KBC_type is not None
        (LBM_collision_HCZ_2d.py:193-286)."""
        mode = 0 if KBC_type is None else (1 if KBCType.is_KBC(KBC_type) else 2)
        return self._hcz(dx, dt, f, rho, vel, density, flags, force, g, pressure, dfai, dprho, mode)


class LBMCollisionHCZ3d(_HCZMixin, LBMCollision3d):
    rank = 3

    def collision(self, dx, dt, f, rho, vel, density, flags, force, g=None, pressure=None, dfai=None, dprho=None,
                  KBC_type: int = None):
        """f, g HCZ collisions on FLUID; KBC_type ignored in 3-D (LBM_collision_HCZ_3d.py:283-361)."""
        return self._hcz(dx, dt, f, rho, vel, density, flags, force, g, pressure, dfai, dprho, False)
