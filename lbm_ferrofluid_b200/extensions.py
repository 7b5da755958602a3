"""SURVEY 8(f) N4 -- extensions the north star names and the reference does not contain.

``ExtSinglePhaseSolver``: single-phase flow on **D3Q19 or D3Q27**, **BGK, three-rate MRT or (D3Q27) entropic KBC** collision
(the reference's D2Q9 KBC scheme, ``LBM_collision_MRT_2d.py:113-226``, carried to 3-D), with **no / velocity-shift /
Guo** forcing, streamed **A/B pull or esoteric-pull in place** (one population buffer instead of two).  One kernel
per step (``csrc/ext_lattice.cu``), the reference's tensor layout ``[B, Q, D, H, W]``, its entropic product
equilibrium (``LBM_collision_3d.py:133-187``) and its full-way bounce-back on ``OBSTACLE`` cells
(``LBM_propagation_3d.py:113-142``).  With ``lattice="D3Q19", forcing="shift", streaming="ab"`` it is the same
step as ``SinglePhaseSolver`` (the reference's ``demo_2d_LBM.py:101-121`` loop in 3-D).

There is no reference oracle for these rows; ``tests/test_gpu_ext.py`` pins them to the D3Q19 path, to
``oracle/ext_oracle.py`` and to mass / momentum balances.  No CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence, Union

import torch

from . import _host as H
from . import _lib

_FORCING = {None: 0, "none": 0, "shift": 1, "guo": 2}
_STREAMING = {"ab": 0, "esoteric": 1}
_COLLISION = {"bgk": 0, "kbc": 1, "mrt": 2}
_LATTICES = {"D3Q19": 19, "D3Q27": 27}


def lattice_tables(lattice: str):
    """(e [Q][3] ints, opp [Q] ints, w [Q] floats) of a lattice, straight from the library's tables (host call)."""
    Q = _LATTICES[lattice]
    e = (ctypes.c_int * (3 * Q))()
    opp = (ctypes.c_int * Q)()
    w = (ctypes.c_double * Q)()
    lib = _lib.load()
    got = lib.lbmff_ext_lattice_info(Q, ctypes.cast(e, ctypes.c_void_p), ctypes.cast(opp, ctypes.c_void_p),
                                     ctypes.cast(w, ctypes.c_void_p))
    if got != Q:
        raise RuntimeError(f"lbmff_ext_lattice_info({Q}) returned {got}")
    return [tuple(e[3 * q:3 * q + 3]) for q in range(Q)], list(opp), list(w)


class ExtSinglePhaseSolver:
    """Single-phase BGK / MRT / KBC on D3Q19 / D3Q27 with selectable forcing and streaming (3-D only).

    collision="mrt": three relaxation times -- ``tau`` for the deviatoric second-order moments (shear viscosity
    cs2 (tau - 1/2) dt, as in BGK), ``tau_bulk`` for their trace, ``tau_high`` for every higher-order moment.

    flags: uint8 ``[B, 1, D, H, W]`` (FLUID = 1, OBSTACLE = 2).  force: ``None``, a 3-sequence (uniform body
    force x, y, z) or a ``[B, 3, D, H, W]`` tensor.  ``f`` always returns the populations in the standard layout
    (post-collision population q at its own cell), whatever the streaming mode holds internally.
    storage="fp16": populations are STORED as fp16 deviations from the rest equilibrium of density ``rho0``
    (arithmetic stays fp32); ``rho0`` defaults to the mean of the density given to ``initialize``.  BGK only.
    """

    def __init__(self, flags: torch.Tensor, *, tau: float, lattice: str = "D3Q27", collision: str = "bgk",
                 forcing: Optional[str] = None, streaming: str = "ab", dx: float = 1.0, dt: float = 1.0,
                 force: Union[None, Sequence[float], torch.Tensor] = None, storage: str = "fp32",
                 rho0: Optional[float] = None, tau_bulk: Optional[float] = None, tau_high: Optional[float] = None):
        if lattice not in _LATTICES:
            raise ValueError(f"lattice must be one of {sorted(_LATTICES)}")
        if collision not in _COLLISION:
            raise ValueError("collision must be 'bgk', 'kbc' or 'mrt'")
        if collision == "kbc" and lattice != "D3Q27":
            raise ValueError("the 3-D KBC collision needs lattice='D3Q27'")
        if forcing not in _FORCING:
            raise ValueError("forcing must be None, 'shift' or 'guo'")
        if streaming not in _STREAMING:
            raise ValueError("streaming must be 'ab' or 'esoteric'")
        if storage not in ("fp32", "fp16"):
            raise ValueError("storage must be 'fp32' or 'fp16'")
        if storage == "fp16" and collision != "bgk":
            raise ValueError("fp16 population storage is built for the BGK collision only")
        self.fp16, self.rho0 = storage == "fp16", (0.0 if rho0 is None else float(rho0))
        self._auto_rho0 = self.fp16 and rho0 is None
        self.flags = H.need_cuda(flags, "flags", torch.uint8)
        if self.flags.dim() != 5:
            raise RuntimeError("ExtSinglePhaseSolver is 3-D: flags must be [B, 1, D, H, W]")
        self.device = self.flags.device
        _, self.B, _, self.D, self.Hh, self.W = H.dims(self.flags)
        self.res = tuple(self.flags.shape[2:])
        self.lattice, self.Q = lattice, _LATTICES[lattice]
        self.forcing, self.streaming = _FORCING[forcing], _STREAMING[streaming]
        self.collision = _COLLISION[collision]
        self.tau, self.dx, self.dt = float(tau), float(dx), float(dt)
        if collision != "mrt" and (tau_bulk is not None or tau_high is not None):
            raise ValueError("tau_bulk / tau_high belong to collision='mrt'")
        # MRT: relaxation times of the trace and of the higher-order moments (default: tau, i.e. BGK)
        self.tau_bulk = self.tau if tau_bulk is None else float(tau_bulk)
        self.tau_high = self.tau if tau_high is None else float(tau_high)
        self._force_t, self._force_u = None, None
        if self.forcing:
            if force is None:
                raise ValueError("forcing requested without a force")
            if torch.is_tensor(force):
                self._force_t = H.need_cuda(force, "force")
                if tuple(self._force_t.shape) != (self.B, 3, *self.res):
                    raise RuntimeError("force must be [B, 3, D, H, W]")
            else:
                fx, fy, fz = (float(v) for v in force)
                self._force_u = (ctypes.c_double * 3)(fx, fy, fz)
        opts = dict(dtype=torch.float32, device=self.device)
        self.rho = torch.zeros((self.B, 1, *self.res), **opts)
        self.vel = torch.zeros((self.B, 3, *self.res), **opts)
        nbuf = 1 if self.streaming else 2
        self._f = [torch.zeros((self.B, self.Q, *self.res), device=self.device,
                               dtype=torch.float16 if self.fp16 else torch.float32) for _ in range(nbuf)]
        self._cur = 0
        self.steps_done = 0

    # ---- state -------------------------------------------------------------------------------------------
    def _sizes(self):
        return (self.B, self.D, self.Hh, self.W)

    def _suffix(self):
        return "d3q19" if self.Q == 19 else "d3q27"

    @property
    def population_bytes(self) -> int:
        """Bytes of population storage held (esoteric pull: one buffer; A/B: two)."""
        return sum(t.numel() * t.element_size() for t in self._f)

    def _layout(self) -> int:
        """0 standard slots (A/B pull); 1 / 2 esoteric pull before an even / odd step."""
        return 1 + (self.steps_done & 1) if self.streaming else 0

    @property
    def f(self) -> torch.Tensor:
        if self.fp16:
            out = torch.empty(self._f[0].shape, dtype=torch.float32, device=self.device)
            H.launch("lbmff_ext_unpack_fp16_" + self._suffix(), self.device, H.ptr(self._f[self._cur]), H.ptr(out),
                     self._layout(), self.rho0, *self._sizes())
            return out
        if not self.streaming:
            return self._f[self._cur]
        out = torch.empty_like(self._f[0])
        H.launch("lbmff_ext_esoteric_convert_" + self._suffix(), self.device, H.ptr(self._f[0]), H.ptr(out),
                 self.steps_done & 1, 0, *self._sizes())
        return out

    def set_f(self, f: torch.Tensor):
        """Load populations given in the standard layout."""
        f = H.need_cuda(f, "f")
        if tuple(f.shape) != tuple(self._f[0].shape):  # fp32 in the standard layout, whatever the storage
            raise RuntimeError(f"f must be {tuple(self._f[0].shape)}")
        if self.fp16:
            H.launch("lbmff_ext_pack_fp16_" + self._suffix(), self.device, H.ptr(f), H.ptr(self._f[self._cur]),
                     self._layout(), self.rho0, *self._sizes())
        elif not self.streaming:
            self._f[self._cur].copy_(f)
        else:
            if f.data_ptr() == self._f[0].data_ptr():
                f = f.clone()
            H.launch("lbmff_ext_esoteric_convert_" + self._suffix(), self.device, H.ptr(f), H.ptr(self._f[0]),
                     self.steps_done & 1, 1, *self._sizes())
        return self

    def initialize(self, rho: torch.Tensor, vel: torch.Tensor):
        """f = feq(rho, vel, force = 0), the way every demo starts (demo_2d_LBM.py:99)."""
        rho = H.need_cuda(rho, "rho")
        vel = H.need_cuda(vel, "vel")
        self.steps_done = 0
        self._cur = 0
        if self._auto_rho0:
            self.rho0 = float(rho.mean())          # one sync, at initialisation only
        self.set_f(_feq_torch(rho, vel, self.lattice, self.dx, self.dt))
        self.rho.copy_(rho)
        self.vel.copy_(vel)
        return self

    # ---- stepping ----------------------------------------------------------------------------------------
    def step(self, n: int = 1):
        mrt = self.collision == _COLLISION["mrt"]
        name = ("lbmff_ext_stream_collide_fp16_" if self.fp16 else
                ("lbmff_ext_stream_collide_mrt_" if mrt else "lbmff_ext_stream_collide_")) + self._suffix()
        fu = ctypes.cast(self._force_u, ctypes.c_void_p) if self._force_u is not None else ctypes.c_void_p(0)
        for _ in range(int(n)):
            if self.streaming:
                src = dst = self._f[0]
            else:
                src, dst = self._f[self._cur], self._f[1 - self._cur]
            mode = (self.forcing, self.streaming, self.steps_done & 1)
            if self.fp16:
                mode = (*mode, self.rho0)
            elif mrt:
                mode = (self.tau_bulk, self.tau_high, *mode)
            else:
                mode = (self.collision, *mode)
            H.launch(name, self.device, H.ptr(src), H.ptr(dst), H.ptr(self.flags), H.ptr(self._force_t), fu,
                     H.ptr(self.rho), H.ptr(self.vel), self.tau, *mode, self.dx, self.dt, *self._sizes())
            if not self.streaming:
                self._cur = 1 - self._cur
            self.steps_done += 1
        return self


def _feq_torch(rho: torch.Tensor, vel: torch.Tensor, lattice: str, dx: float, dt: float) -> torch.Tensor:
    """Initial condition only (outside the per-step path): the product equilibrium of LBM_collision_3d.py:133-187
    on either lattice, a few pointwise torch ops on the device the inputs live on."""
    e, _, w = lattice_tables(lattice)
    c = dx / dt
    s = torch.sqrt(1.0 + 3.0 * vel * vel / c / c)
    base = (2.0 - s).prod(dim=1, keepdim=True)
    ratio = (2.0 * vel / c + s) / (1.0 - vel / c)
    out = torch.empty((rho.shape[0], len(e), *rho.shape[2:]), dtype=rho.dtype, device=rho.device)
    for q, eq in enumerate(e):
        v = base
        for a in range(3):
            if eq[a] == 1:
                v = v * ratio[:, a:a + 1]
            elif eq[a] == -1:
                v = v / ratio[:, a:a + 1]
        out[:, q:q + 1] = rho * float(torch.tensor(w[q], dtype=torch.float32)) * v
    return out


from .collision import LBMCollision3d  # noqa: E402


class LBMCollisionSC3d(LBMCollision3d):
    """Shan-Chen pseudopotential multiphase on D3Q19: ``LBMCollisionSC2d`` (LBM_collision_SC_2d.py) carried to
    3-D with the same method names and argument meaning.  The reference's Shan-Chen model is 2-D only and its
    factory returns ``None`` in 3-D (simulation_runner.py ``create_collision_SC``); this class is constructed
    directly, the factory keeps the reference behaviour."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._density_wall = 0.5 * (self._density_gas + self._density_liquid)

    def calculate_force(self, dx, dt, density, flags, force=None) -> torch.Tensor:
        d = H.need_cuda(density, "density")
        flags = H.need_cuda(flags, "flags", torch.uint8)
        dim, B, C, D, Hh, W = H.dims(d)
        if dim != 3:
            raise RuntimeError("LBMCollisionSC3d needs [B, 1, D, H, W] tensors")
        out = torch.empty((B, 3, D, Hh, W), dtype=d.dtype, device=d.device)
        H.launch("lbmff_sc_force_d3q19", d.device, H.ptr(d), H.ptr(flags), float(self._density_liquid),
                 float(self._density_gas), float(dx), float(dt), H.ptr(out), B, D, Hh, W)
        return out

    def collision(self, dx, dt, f, rho, density, vel, flags, force=None, KBC_type: int = None):
        """SC force then BGK, like LBM_collision_SC_2d.py:99-137."""
        force = self.calculate_force(dx=dx, dt=dt, density=density, flags=flags, force=force)
        return self._bgk(dx, dt, f, rho, vel, flags, force)
