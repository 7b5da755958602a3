"""CPU oracle for the LBM-Ferrofluid per-timestep solver path.

TEST INFRASTRUCTURE ONLY.  This module is a dimension-generic torch-CPU
restatement of the reference algorithm (meng-yijie1996/LBM-Ferrofluid,
``src/LBM``).  It is imported only by ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s CPU-baseline / ``--impl reference`` legs, and only as the
checker or as the timed CPU baseline -- never by the product path
(``lbm_ferrofluid_b200``), which calls hand-written sm_100a kernels and has no
CPU fallback.

Parity is PINNED: ``oracle/make_golden.py`` imports the unmodified reference
from ``/root/reference`` (which has no tests or golden vectors of its own,
SURVEY.md section 4) and stores its outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every function below against them.

Conventions (reference: ``LBM_propagation_3d.py:67-109``): tensors are
``[B, C, (D,) H, W]``, last axis = x, then y, then z.  ``lat.e[q] = (ex, ey[, ez])``.
Every function takes/returns fresh tensors and works in the dtype of its
inputs (fp32 for parity with the reference, fp64 for error floors).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

FLUID = 1      # CellType.FLUID     (reference: src/LBM/utils/types.py:9)
OBSTACLE = 2   # CellType.OBSTACLE  (reference: src/LBM/utils/types.py:10)

# KBCType values (reference: src/LBM/utils/types.py:61-66)
LBGK, KBC_A, KBC_B, KBC_C, KBC_D = 0, 0b10000101, 0b10000110, 0b10001001, 0b10001010

MAG_K = 0.33                  # hard-coded susceptibility (LBM_magnetic_3d.py:153, HCZ_3d.py:233)
MU0 = 4 * math.pi * 1e-7      # HCZ_3d.py:232


# --------------------------------------------------------------------------
# lattices
# --------------------------------------------------------------------------
@dataclass(frozen=True)
class Lattice:
    dim: int
    Q: int
    e: Tuple[Tuple[int, ...], ...]
    w: Tuple[float, ...]
    opp: Tuple[int, ...]


def _make_lattice(e, w):
    e = tuple(tuple(v) for v in e)
    opp = tuple(e.index(tuple(-c for c in v)) for v in e)
    return Lattice(dim=len(e[0]), Q=len(e), e=e, w=tuple(w), opp=opp)


# reference ordering: LBM_collision_2d.py:66-83 / weights :44-63
D2Q9 = _make_lattice(
    [(0, 0), (1, 0), (0, 1), (-1, 0), (0, -1), (1, 1), (-1, 1), (-1, -1), (1, -1)],
    [4 / 9] + [1 / 9] * 4 + [1 / 36] * 4,
)
# reference ordering: LBM_collision_3d.py:76-103 / weights :47-73
D3Q19 = _make_lattice(
    [(0, 0, 0),
     (1, 0, 0), (0, 1, 0), (-1, 0, 0), (0, -1, 0),
     (1, 1, 0), (-1, 1, 0), (-1, -1, 0), (1, -1, 0),
     (0, 0, 1),
     (1, 0, 1), (0, 1, 1), (-1, 0, 1), (0, -1, 1),
     (0, 0, -1),
     (1, 0, -1), (0, 1, -1), (-1, 0, -1), (0, -1, -1)],
    [1 / 3] + [1 / 18] * 4 + [1 / 36] * 4 + [1 / 18] + [1 / 36] * 4 + [1 / 18] + [1 / 36] * 4,
)


def lattice_for(t: torch.Tensor) -> Lattice:
    """Pick the lattice from the rank of a ``[B,C,...]`` tensor."""
    return D2Q9 if t.dim() == 4 else D3Q19


def _wt(lat: Lattice, like: torch.Tensor) -> torch.Tensor:
    # the reference builds weights as fp32 ``torch.Tensor`` then casts
    # (LBM_collision_3d.py:47-74): keep that rounding.
    w = torch.tensor(lat.w, dtype=torch.float32).to(like.dtype)
    return w.reshape(1, lat.Q, *([1] * lat.dim))


def _axis(lat: Lattice, a: int) -> int:
    """torch dim index (negative) of spatial axis a (0=x,1=y,2=z)."""
    return -1 - a


def _interior(t: torch.Tensor, dim: int) -> torch.Tensor:
    sl = [Ellipsis] + [slice(1, -1)] * dim
    return t[tuple(sl)]


def _shift(t: torch.Tensor, dim: int, off: Sequence[int]) -> torch.Tensor:
    """View of the interior-sized window displaced by ``off`` (x,y[,z]) cells."""
    sl = [Ellipsis] + [None] * dim
    for a in range(dim):
        o = off[a]
        n = t.shape[-1 - a]
        sl[len(sl) - 1 - a] = slice(1 + o, n - 1 + o)
    return t[tuple(sl)]


# --------------------------------------------------------------------------
# a1/a2 streaming and bounce-back
# --------------------------------------------------------------------------
def propagation(f: torch.Tensor) -> torch.Tensor:
    """Periodic pull streaming f'_q(x) = f_q(x - e_q).

    Reference: LBM_propagation_3d.py:18-111, LBM_propagation_2d.py:18-68 (the
    pad / wrap / shifted-copy sequence there is exactly a periodic roll).
    """
    lat = lattice_for(f)
    out = torch.empty_like(f)
    for q in range(lat.Q):
        axes = [a for a in range(lat.dim) if lat.e[q][a] != 0]
        if axes:
            out[:, q] = torch.roll(f[:, q], shifts=[lat.e[q][a] for a in axes], dims=[-1 - a for a in axes])
        else:
            out[:, q] = f[:, q]
    return out


def rebounce_obstacle(f: torch.Tensor, flags: torch.Tensor) -> torch.Tensor:
    """Full-way bounce-back: swap opposite populations on OBSTACLE cells.

    Reference: LBM_propagation_3d.py:113-142, LBM_propagation_2d.py:70-86.
    """
    lat = lattice_for(f)
    return torch.where(flags == OBSTACLE, f[:, list(lat.opp)], f)


# --------------------------------------------------------------------------
# a3/a4 macroscopic moments and EOS
# --------------------------------------------------------------------------
def get_pressure(dx: float, dt: float, density: torch.Tensor) -> torch.Tensor:
    """Carnahan-Starling pressure, operation order of LBM_macro_compute_3d.py:59-71."""
    c = dx / dt
    RT = c * c / 3.0
    a = 12.0 * RT
    b = 4.0
    t = b * density / 4.0
    return density * RT * t * (4.0 - 2.0 * t) / torch.pow(1 - t, 3) - a * density * density + density * RT


def equation_of_states(dx: float, dt: float, rho: torch.Tensor) -> torch.Tensor:
    """Carnahan-Starling pressure, operation order of LBM_collision_3d.py:105-122."""
    c = dx / dt
    RT = c * c / 3.0
    a = 12.0 * RT
    b = 4.0
    t = b * rho / 4.0
    return rho * RT * (4.0 * t - 2.0 * t * t) / torch.pow(1.0 - t, 3) + rho * RT - a * rho * rho


def affine_density(rho, density_liquid, density_gas, rho_liquid, rho_gas):
    """rho -> density map (LBM_macro_compute_3d.py:107-114, HCZ_3d.py:213-215)."""
    return density_gas + (density_liquid - density_gas) * ((rho - rho_gas) / (rho_liquid - rho_gas))


def macro_compute(dx, dt, f, rho, vel, flags, density=None,
                  density_liquid=0.265, density_gas=0.038, rho_liquid=0.265, rho_gas=0.038):
    """rho = sum f, u = sum f e c / rho; old values kept on OBSTACLE.

    Reference: LBM_macro_compute_3d.py:73-120, LBM_macro_compute_2d.py:65-107.
    """
    lat = lattice_for(f)
    c = dx / dt
    obst = flags == OBSTACLE
    rho_new = torch.where(obst, rho, f.sum(dim=1, keepdim=True))
    comps = []
    for a in range(lat.dim):
        acc = torch.zeros_like(rho)
        for q in range(lat.Q):
            if lat.e[q][a] != 0:
                acc = acc + f[:, q:q + 1] * lat.e[q][a]
        comps.append(acc)
    vel_new = torch.where(obst, vel, torch.cat(comps, dim=1) * (c / rho_new))
    if density is not None:
        return [rho_new, vel_new, affine_density(rho_new, density_liquid, density_gas, rho_liquid, rho_gas)]
    return [rho_new, vel_new]


def get_vort(vel: torch.Tensor, dx: float) -> torch.Tensor:
    """2-D vorticity, LBM_macro_compute_2d.py:109-117."""
    v = ((vel[..., 0:1, 2:, 1:-1] - vel[..., 0:1, :-2, 1:-1])
         - (vel[..., 1:2, 1:-1, 2:] - vel[..., 1:2, 1:-1, :-2])) / (2.0 * dx)
    return F.pad(v, pad=(1, 1, 1, 1), mode="replicate")


# --------------------------------------------------------------------------
# a5/a6/a14 equilibria
# --------------------------------------------------------------------------
def get_feq(dx, dt, rho, vel, force=None, tau=1.0, lat: Optional[Lattice] = None):
    """Entropic product-form equilibrium (LBM_collision_3d.py:133-187, _2d.py:113-161)."""
    lat = lat or (D2Q9 if vel.shape[1] == 2 else D3Q19)
    if force is not None:
        vel = vel + tau * force / rho
    c = dx / dt
    s = torch.sqrt(1.0 + 3.0 * vel * vel / c / c)
    base = None
    ratio = []
    for a in range(lat.dim):
        ua, sa = vel[:, a:a + 1], s[:, a:a + 1]
        base = (2.0 - sa) if base is None else base * (2.0 - sa)
        ratio.append((2.0 * ua / c + sa) / (1.0 - ua / c))
    w = _wt(lat, rho)
    out = torch.empty((rho.shape[0], lat.Q, *rho.shape[2:]), dtype=rho.dtype)
    for q in range(lat.Q):
        val = base
        for a in range(lat.dim):
            if lat.e[q][a] == 1:
                val = val * ratio[a]
            elif lat.e[q][a] == -1:
                val = val * (1.0 / ratio[a])
        out[:, q:q + 1] = rho * w[:, q:q + 1] * val
    return out


def get_geq(dx, dt, rho, vel, density, pressure, force=None, feq=None, tau=1.0):
    """geq = w (p - cs2 density) + cs2 density / rho * feq (LBM_collision_3d.py:189-207)."""
    c = dx / dt
    cs2 = c * c / 3.0
    if feq is None:
        feq = get_feq(dx, dt, rho, vel, force, tau)
    lat = D2Q9 if vel.shape[1] == 2 else D3Q19
    return _wt(lat, rho) * (pressure - cs2 * density) + cs2 * density / rho * feq


def compute_Gamma(dx, dt, vel):
    """Second-order polynomial Gamma(u) (LBM_collision_HCZ_3d.py:271-281)."""
    lat = D2Q9 if vel.shape[1] == 2 else D3Q19
    c = dx / dt
    cs2 = c * c / 3.0
    uv = (vel * vel).sum(dim=1, keepdim=True)
    w = _wt(lat, vel)
    out = torch.empty((vel.shape[0], lat.Q, *vel.shape[2:]), dtype=vel.dtype)
    for q in range(lat.Q):
        eu = torch.zeros_like(uv)
        for a in range(lat.dim):
            if lat.e[q][a] != 0:
                eu = eu + vel[:, a:a + 1] * (lat.e[q][a] * c)
        out[:, q:q + 1] = w[:, q:q + 1] * (1.0 + eu / cs2 + 0.5 * eu * eu / cs2 / cs2 - 0.5 * uv / cs2)
    return out


def smooth_phi(phi: torch.Tensor, eps: float) -> torch.Tensor:
    """Smoothed Heaviside (LBM_collision_HCZ_3d.py:265-269, LBM_magnetic_3d.py:119-123)."""
    return (phi > eps) * 1.0 + (torch.abs(phi) <= eps) * (
        0.5 + (0.5 / eps) * phi + (0.5 / math.pi) * torch.sin((math.pi / eps) * phi))


def susceptibility(phi: torch.Tensor, dx: float) -> torch.Tensor:
    """chi = k (1 - H_eps(phi)), eps = 0.1 dx (LBM_magnetic_3d.py:162, HCZ_3d.py:233-234)."""
    return MAG_K * (1.0 - smooth_phi(phi, 0.1 * dx))


# --------------------------------------------------------------------------
# a7/a8 stencils, a17 staggered faces
# --------------------------------------------------------------------------
def get_grad(input_: torch.Tensor, dx: float, flags: torch.Tensor) -> torch.Tensor:
    """Isotropic central gradient with the reference's shell rules.

    Reference: LBM_collision_3d.py:209-279 (3-D: (2*axis + 8 diagonals)/12dx),
    LBM_collision_2d.py:183-227 (2-D: (4*axis + 4 diagonals)/12dx).  OBSTACLE
    shell cells take the nearest interior value first; the output is the
    interior gradient replicate-padded.
    """
    if input_.shape[1] != 1:
        raise RuntimeError("To get your grad operation, channel dim has to be 1")
    dim = input_.dim() - 2
    pad = (1, 1) * dim
    x = torch.where(flags == OBSTACLE, F.pad(_interior(input_, dim), pad=pad, mode="replicate"), input_)
    axis_w = 2.0 if dim == 3 else 4.0
    comps = []
    for a in range(dim):
        ea = [0] * dim
        ea[a] = 1
        ma = [-v for v in ea]
        acc = None
        for b in range(dim):
            if b == a:
                continue
            for sgn in (1, -1):
                p = list(ea); p[b] = sgn
                m = list(ma); m[b] = sgn
                term = _shift(x, dim, p) - _shift(x, dim, m)
                acc = term if acc is None else acc + term
        comps.append((axis_w * (_shift(x, dim, ea) - _shift(x, dim, ma)) + acc) / 12.0 / dx)
    return F.pad(torch.cat(comps, dim=1), pad=pad, mode="replicate")


def get_laplacian(input_: torch.Tensor, dx: float, flags: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Isotropic Laplacian, zero on the shell, flags ignored.

    Reference: LBM_collision_3d.py:281-318 ((2*S6 + S12 - 24c)/6dx^2),
    LBM_collision_2d.py:229-256 ((4*S4 + S4diag - 20c)/5dx^2).
    """
    dim = input_.dim() - 2
    near, far = None, None
    offs = [[-1, 0, 1]] * dim
    import itertools
    for off in itertools.product(*offs):
        n = sum(abs(v) for v in off)
        if n == 1:
            t = _shift(input_, dim, off)
            near = t if near is None else near + t
        elif n == 2:
            t = _shift(input_, dim, off)
            far = t if far is None else far + t
    ctr = _interior(input_, dim)
    if dim == 3:
        val = (2.0 * near + far - 24 * ctr) / 6.0 / (dx * dx)
    else:
        val = (4.0 * near + far - 20 * ctr) / 5.0 / (dx * dx)
    return F.pad(val, pad=(1, 1) * dim, mode="constant", value=0)


def get_staggered_axis(input_: torch.Tensor, a: int) -> torch.Tensor:
    """Face average along axis a, replicate-padded to N+1 faces (grid.py:7-49)."""
    if input_.dim() not in (4, 5) or (a == 2 and input_.dim() != 5):
        raise RuntimeError("A grid has to be 2D(3D) [B, C, (D), H, W] to be staggered")
    d = input_.dim() - 1 - a
    n = input_.shape[d]
    mid = (input_.narrow(d, 1, n - 1) + input_.narrow(d, 0, n - 1)) * 0.5
    first = mid.narrow(d, 0, 1)
    last = mid.narrow(d, n - 2, 1)
    return torch.cat([first, mid, last], dim=d)


def get_staggered(input_: torch.Tensor) -> List[torch.Tensor]:
    """grid.py:52-64."""
    dim = input_.shape[1]
    if dim < 2 or dim > 3:
        raise RuntimeError("Only 2D or 3D scene supported")
    return [get_staggered_axis(input_[:, a:a + 1], a) for a in range(dim)]


# --------------------------------------------------------------------------
# a9 BGK, a10 KBC, a11 Shan-Chen
# --------------------------------------------------------------------------
def collision_bgk(dx, dt, f, rho, vel, flags, force, tau):
    """BGK with velocity-shift forcing on non-OBSTACLE cells (LBM_collision_3d.py:320-346)."""
    feq = get_feq(dx, dt, rho, vel, force, tau)
    return torch.where(flags == OBSTACLE, f, (1.0 - 1.0 / tau) * f + feq / tau)


def kbc_s_part(dx, dt, f, rho, vel, KBC_type):
    """The 's' (shear) part of the KBC moment split for D2Q9.

    Reference: LBM_collision_MRT_2d.py:62-182.  Every KBCType predicate tests
    the shared high bit (types.py:72-107), so any KBC_* value selects centred
    moments with s = {T, N, Pi_xy}; ``None``/LBGK selects raw moments with
    s = {T, Pi_xy, Q_xxy, Q_xyy, A}.
    """
    lat = D2Q9
    c = dx / dt
    kbc = KBC_type is not None and (KBC_type & 0b10000000) > 0
    fr = f / rho
    M = {}
    for p in range(3):
        for q in range(3):
            if (p, q) in ((0, 0), (1, 0), (0, 1)):
                continue
            acc = torch.zeros_like(rho)
            for i in range(lat.Q):
                ex, ey = lat.e[i]
                if kbc:
                    vx = c * torch.pow(ex - vel[:, 0:1], p)
                    vy = c * torch.pow(ey - vel[:, 1:2], q)
                    acc = acc + vx * vy * fr[:, i:i + 1]
                else:
                    coef = (c ** (p + q)) * (ex ** p) * (ey ** q)
                    if coef != 0:
                        acc = acc + coef * fr[:, i:i + 1]
            M[(p, q)] = acc
    T = M[(2, 0)] + M[(0, 2)]
    N = M[(2, 0)] - M[(0, 2)]
    Pxy = M[(1, 1)]
    Qxxy, Qxyy, A = M[(2, 1)], M[(1, 2)], M[(2, 2)]
    s = [torch.zeros_like(rho) for _ in range(9)]
    s[0] = s[0] + rho * (1 - T)
    for i in (1, 3, 2, 4):
        s[i] = s[i] + 0.5 * rho * (0.5 * T)
    for i, sl in ((5, 1), (6, -1), (7, 1), (8, -1)):
        s[i] = s[i] + 0.25 * rho * (sl * Pxy)
    if kbc:
        for i in (1, 3):
            s[i] = s[i] + 0.5 * rho * (0.5 * N)
        for i in (2, 4):
            s[i] = s[i] + 0.5 * rho * (0.5 * (-N))
    else:
        s[0] = s[0] + rho * A
        for i, sg in ((1, 1), (3, -1)):
            s[i] = s[i] + 0.5 * rho * (sg * (vel[:, 0:1] - Qxyy) - A)
        for i, lm in ((2, 1), (4, -1)):
            s[i] = s[i] + 0.5 * rho * (lm * (vel[:, 1:2] - Qxxy) - A)
        for i, sg, lm in ((5, 1, 1), (6, -1, 1), (7, -1, -1), (8, 1, -1)):
            s[i] = s[i] + 0.25 * rho * (A + sg * Qxyy + lm * Qxxy)
    return torch.cat(s, dim=1)


def _kbc_relax(dx, dt, f, feq, rho, vel, tau, KBC_type):
    ds = kbc_s_part(dx, dt, f, rho, vel, KBC_type) - kbc_s_part(dx, dt, feq, rho, vel, KBC_type)
    dh = (f - feq) - ds
    beta = 0.5 / tau
    gamma = 1.0 / beta - (2.0 - 1.0 / beta) * (ds * dh / feq).sum(dim=1, keepdim=True) / (
        dh * dh / feq).sum(dim=1, keepdim=True)
    return f + beta * (-2.0 * ds - gamma * dh)


def collision_kbc(dx, dt, f, rho, vel, flags, force, tau, KBC_type):
    """Entropic KBC collision on non-OBSTACLE cells (LBM_collision_MRT_2d.py:184-226)."""
    feq = get_feq(dx, dt, rho, vel, force, tau)
    return torch.where(flags == OBSTACLE, f, _kbc_relax(dx, dt, f, feq, rho, vel, tau, KBC_type))


def sc_force(dx, dt, density, flags, density_liquid, density_gas):
    """Shan-Chen interaction force with the C-S EOS pseudopotential.

    Reference: LBM_collision_SC_2d.py:22-97 (G = -1/3, wall psi from the mean of
    the liquid and gas densities, zero outside the array, force on FLUID only).
    """
    lat = D2Q9
    c = dx / dt
    cs2 = c * c / 3.0
    RT, a, b, G1 = cs2, 12.0 * cs2, 4.0, -1.0 / 3.0

    def psi_of(d, sqrt, power):
        t = d * b / 4.0
        v = RT * (1.0 + (4.0 * t - 2.0 * t * t) / power(1.0 - t, 3)) - a * d - cs2
        return sqrt(2.0 * d * v / G1 / cs2)

    psx = psi_of(density, torch.sqrt, torch.pow)
    wall = psi_of(0.5 * (density_gas + density_liquid), math.sqrt, math.pow)
    pad = F.pad(torch.where(flags == OBSTACLE, wall * torch.ones_like(psx), psx), pad=(1, 1, 1, 1))
    w = _wt(lat, density)
    comps = []
    for ax in range(2):
        acc = torch.zeros_like(psx)
        for q in range(1, 9):
            ex, ey = lat.e[q]
            if lat.e[q][ax] == 0:
                continue
            nb = pad[..., 1 + ey:pad.shape[-2] - 1 + ey, 1 + ex:pad.shape[-1] - 1 + ex]
            acc = acc + w[:, q:q + 1] * lat.e[q][ax] * nb
        comps.append(acc)
    force = -G1 * psx * c * torch.cat(comps, dim=1)
    return torch.where(flags == FLUID, force, torch.zeros_like(force))


def collision_sc(dx, dt, f, rho, density, vel, flags, tau, density_liquid, density_gas):
    """Shan-Chen collision = SC force + BGK (LBM_collision_SC_2d.py:99-137)."""
    force = sc_force(dx, dt, density, flags, density_liquid, density_gas)
    return collision_bgk(dx, dt, f, rho, vel, flags, force, tau)


# --------------------------------------------------------------------------
# a12 capillary process, a15 HCZ collision
# --------------------------------------------------------------------------
def _wetting_3d(rho, flags, cot):
    """Ordered in-place wetting boundary on the rho shell (HCZ_3d.py:84-211)."""
    eps = 1e-6
    X, Y, Z = -1, -2, -3

    def face(axis, wall, inner1, inner2, use_tan):
        # tangential axes
        t1, t2 = [a for a in (X, Y, Z) if a != axis]
        # central differences on the first interior plane along both tangential axes
        def tdiff(t):
            full = rho
            hi = [slice(None)] * rho.dim()
            lo = [slice(None)] * rho.dim()
            ot = t2 if t == t1 else t1
            hi[axis] = inner1; lo[axis] = inner1
            hi[t] = slice(2, None); lo[t] = slice(None, -2)
            hi[ot] = slice(1, -1); lo[ot] = slice(1, -1)
            return full[tuple(hi)] - full[tuple(lo)]
        d1, d2 = tdiff(t1), tdiff(t2)
        hlp = torch.sqrt(eps + d1 * d1 + d2 * d2)
        tgt = [slice(None)] * rho.dim()
        src = [slice(None)] * rho.dim()
        tgt[axis] = wall; src[axis] = inner2
        for t in (t1, t2):
            tgt[t] = slice(1, -1); src[t] = slice(1, -1)
        new = rho[tuple(src)] + cot * hlp if use_tan else rho[tuple(src)]
        rho[tuple(tgt)] = torch.where(flags[tuple(tgt)] == OBSTACLE, new, rho[tuple(tgt)])

    face(X, 0, 1, 2, True)
    face(X, -1, -2, -3, True)
    face(Y, 0, 1, 2, True)
    face(Y, -1, -2, -3, True)
    face(Z, 0, 1, 2, False)
    face(Z, -1, -2, -3, False)

    m = slice(1, -1)
    # 12 edge lines, in reference order (HCZ_3d.py:171-184); index order [z, y, x]
    for (z, y, x), (az, ay, ax), (bz, by, bx) in (
        ((m, 0, 0), (m, 0, 1), (m, 1, 0)), ((m, 0, -1), (m, 0, -2), (m, 1, -1)),
        ((m, -1, 0), (m, -1, 1), (m, -2, 0)), ((m, -1, -1), (m, -1, -2), (m, -2, -1)),
        ((0, m, 0), (0, m, 1), (1, m, 0)), ((0, m, -1), (0, m, -2), (1, m, -1)),
        ((-1, m, 0), (-1, m, 1), (-2, m, 0)), ((-1, m, -1), (-1, m, -2), (-2, m, -1)),
        ((0, 0, m), (0, 1, m), (1, 0, m)), ((0, -1, m), (0, -2, m), (1, -1, m)),
        ((-1, 0, m), (-1, 1, m), (-2, 0, m)), ((-1, -1, m), (-1, -2, m), (-2, -1, m)),
    ):
        rho[..., z, y, x] = 0.5 * (rho[..., az, ay, ax] + rho[..., bz, by, bx])
    # 8 corners (HCZ_3d.py:187-211)
    for z in (0, -1):
        for y in (0, -1):
            for x in (0, -1):
                zi, yi, xi = (1 if z == 0 else -2), (1 if y == 0 else -2), (1 if x == 0 else -2)
                rho[..., z, y, x] = (rho[..., z, y, xi] + rho[..., z, yi, x] + rho[..., zi, y, x]) / 3.0
    return rho


def _wetting_2d(rho, flags, cot):
    """Ordered in-place wetting boundary on the rho shell (HCZ_2d.py:84-121)."""
    def col(wall, inner1, inner2):
        hlp = torch.abs(rho[..., 2:, inner1] - rho[..., :-2, inner1])
        rho[..., 1:-1, wall] = torch.where(flags[..., 1:-1, wall] == OBSTACLE,
                                           rho[..., 1:-1, inner2] + cot * hlp, rho[..., 1:-1, wall])

    def row(wall, inner1, inner2):
        hlp = torch.abs(rho[..., inner1, 2:] - rho[..., inner1, :-2])
        rho[..., wall, 1:-1] = torch.where(flags[..., wall, 1:-1] == OBSTACLE,
                                           rho[..., inner2, 1:-1] + cot * hlp, rho[..., wall, 1:-1])

    col(0, 1, 2); col(-1, -2, -3); row(0, 1, 2); row(-1, -2, -3)
    rho[..., 0, 0] = 0.5 * (rho[..., 1, 0] + rho[..., 0, 1])
    rho[..., -1, 0] = 0.5 * (rho[..., -2, 0] + rho[..., -1, 1])
    rho[..., 0, -1] = 0.5 * (rho[..., 0, -2] + rho[..., 1, -1])
    rho[..., -1, -1] = 0.5 * (rho[..., -2, -2] + rho[..., -2, -2])   # sic: HCZ_2d.py:121
    return rho


def capillary_process(rho, vel, flags, dt, dx, g, density, pressure, H2=None, phi=None, *,
                      kappa, gravity, contact_angle, density_liquid, density_gas, rho_liquid, rho_gas):
    """Wetting BC, surface tension, gravity, Kelvin force, g-moments.

    Reference: LBM_collision_HCZ_3d.py:21-263, LBM_collision_HCZ_2d.py:21-173.
    ``rho`` is modified IN PLACE (shell only) and returned, like the reference.
    ``gravity`` is the scalar passed to ``set_gravity`` (acts along -y).
    Returns [rho, vel, density, pressure, force, dfai, dprho].
    """
    dim = rho.dim() - 2
    lat = D2Q9 if dim == 2 else D3Q19
    pad = (1, 1) * dim
    c = dx / dt
    RT = c * c / 3.0
    prho = F.pad(_interior(pressure - RT * density, dim), pad=pad, mode="replicate")
    fai = F.pad(_interior(equation_of_states(dx, dt, rho) - rho * RT, dim), pad=pad, mode="replicate")
    ca = contact_angle if torch.is_tensor(contact_angle) else torch.tensor([contact_angle], dtype=rho.dtype)
    cot = torch.tan(math.pi / 2.0 - ca.to(rho.dtype))
    rho = _wetting_3d(rho, flags, cot) if dim == 3 else _wetting_2d(rho, flags, cot)
    density = affine_density(rho, density_liquid, density_gas, rho_liquid, rho_gas)
    lap = get_laplacian(density, dx)
    force = kappa * density * get_grad(lap, dx, flags)
    gvec = torch.zeros(dim, dtype=torch.float32)
    gvec[1] = -gravity
    force = force + gvec.to(rho.dtype).reshape(1, dim, *([1] * dim)) * density
    if H2 is not None:
        chi = susceptibility(phi, dx)
        force = force + (-0.5 * MU0 * H2 * get_grad(chi, dx, flags))
    dfai = get_grad(fai, dx, flags)
    dprho = get_grad(prho, dx, flags)
    fluid = flags == FLUID
    mom = []
    for a in range(dim):
        acc = torch.zeros_like(rho)
        for q in range(lat.Q):
            if lat.e[q][a] != 0:
                acc = acc + g[:, q:q + 1] * lat.e[q][a]
        mom.append(acc)
    macro_vel = (torch.cat(mom, dim=1) * c + 0.5 * dt * RT * force) / RT / density
    vel = torch.where(fluid, macro_vel, vel)
    macro_p = g.sum(dim=1, keepdim=True) - 0.5 * dt * (vel * dprho).sum(dim=1, keepdim=True)
    pressure = torch.where(fluid, macro_p, pressure)
    return [rho, vel, density, pressure, force, dfai, dprho]


def collision_hcz(dx, dt, f, rho, vel, density, flags, force, g, pressure, dfai, dprho,
                  tau_f, tau_g, KBC_type=None):
    """He-Chen-Zhang f and g collisions on FLUID cells.

    Reference: LBM_collision_HCZ_3d.py:283-361 (3-D ignores KBC_type),
    LBM_collision_HCZ_2d.py:193-286 (2-D: KBC relaxation on g when KBC_type is not None).
    """
    dim = rho.dim() - 2
    lat = D2Q9 if dim == 2 else D3Q19
    c = dx / dt
    RT = c * c / 3.0
    feq = get_feq(dx, dt, rho, vel, None)
    geq = get_geq(dx, dt, rho, vel, density, pressure, None, feq)
    Gam = compute_Gamma(dx, dt, vel)
    w = _wt(lat, rho)

    def edot(vec):  # sum_a (e_qa c - u_a) vec_a  -> [B,Q,...]
        out = torch.empty_like(f)
        for q in range(lat.Q):
            acc = torch.zeros_like(rho)
            for a in range(dim):
                acc = acc + (lat.e[q][a] * c - vel[:, a:a + 1]) * vec[:, a:a + 1]
            out[:, q:q + 1] = acc
        return out

    f_force = dt * (1.0 - 0.5 / tau_f) * Gam / RT * edot(-dfai) * dt
    g_force = (1.0 - 0.5 / tau_g) * (Gam * edot(force) + (Gam - w) * edot(-dprho)) * dt
    if dim == 3:
        cf = f + (f_force - f / tau_f + feq / tau_f)
        cg = g + (g_force - g / tau_g + geq / tau_g)
    else:
        cg = g + 1.0 / tau_g * (geq - g)
        if KBC_type is not None:
            cg = _kbc_relax(dx, dt, g, geq, rho, vel, tau_g, KBC_type)
        cf = f + 1.0 / tau_f * (feq - f) + f_force
        cg = cg + g_force
    fluid = flags == FLUID
    return [torch.where(fluid, cf, f), torch.where(fluid, cg, g)]


# --------------------------------------------------------------------------
# a16 magnetic Poisson-LBM sweep
# --------------------------------------------------------------------------
def magnetic_rhs(dx, phi, flags, H_ext_mac):
    """Sweep-invariant source div(chi_face H_face) dx / (1 + chi), FLUID only.

    Reference: LBM_magnetic_3d.py:161-179, LBM_magnetic_2d.py:139-152.
    """
    dim = phi.dim() - 2
    chi = susceptibility(phi, dx)
    div = None
    for a in range(dim):
        face = get_staggered_axis(chi, a) * H_ext_mac[a]
        d = phi.dim() - 1 - a
        n = phi.shape[d]
        t = face.narrow(d, 1, n) - face.narrow(d, 0, n)
        div = t if div is None else div + t
    rhs = div * dx / (1.0 + chi)
    return torch.where(flags == FLUID, rhs, torch.zeros_like(rhs))


def psi_of_h(h):
    """psi = sum_{i>=1} h_i / (1 - w0) (LBM_magnetic_3d.py:155-157)."""
    lat = lattice_for(h)
    w0 = _wt(lat, h)[:, 0:1]
    return h[:, 1:].sum(dim=1, keepdim=True) / (1.0 - w0)


def get_H_int(dt, dx, phi, flags, H_ext_mac, h, tau=1.0):
    """One collision sweep of the Poisson lattice + H_int = -grad psi.

    Reference: LBM_magnetic_3d.py:125-188, LBM_magnetic_2d.py:105-164.
    """
    lat = lattice_for(h)
    c = dx / dt
    cs2 = c * c / 3.0
    w = _wt(lat, h)
    psi = psi_of_h(h)
    heq = psi * w
    heq[:, 0:1] = heq[:, 0:1] - psi
    rhs = magnetic_rhs(dx, phi, flags, H_ext_mac)
    add_h = dt * w * rhs * (cs2 * (0.5 - tau) * dt)
    coll = (1.0 - 1.0 / tau) * h + (1.0 / tau) * heq + add_h
    new_h = torch.where(flags == OBSTACLE, h, coll)
    return [-get_grad(psi, dx, flags), new_h]


# --------------------------------------------------------------------------
# demo-level step loops (the time loop lives in the reference's demo scripts)
# --------------------------------------------------------------------------
@dataclass
class HCZParams:
    """Parameters of the HCZ (+magnetic) demos; defaults = 3-D Rosensweig demo
    (demo_3d_LBM_Rosensweig_instability.py:38-77)."""
    dx: float = 1.0
    dt: float = 1.0
    density_gas: float = 0.02381
    density_liquid: float = 0.2508
    rho_gas: float = 0.02381
    rho_liquid: float = 0.2508
    kappa: float = 0.01
    tau_f: float = 0.68
    tau_g: float = 0.68
    tau_h: float = 1.0
    contact_angle: float = 0.5 * math.pi
    gravity: float = 1e-4
    n_sweeps: int = 30
    KBC_type: Optional[int] = None
    magnetic: bool = True


def phase_field(density, p: HCZParams):
    """phi = -(2 (density - rho_g)/(rho_l - rho_g) - 1)  (demo_3d...Rosensweig...py:162)."""
    return -(2.0 * (density - p.density_gas) / (p.density_liquid - p.density_gas) - 1.0)


def hcz_init(rho, density, vel, p: HCZParams):
    """Initial pressure, f, g from rho/density/vel (demo_3d...Rosensweig...py:133-145).

    Note the reference initialises f from ``density`` and g with ``feq=f``."""
    pressure = get_pressure(p.dx, p.dt, density)
    f = get_feq(p.dx, p.dt, density, vel, torch.zeros_like(vel), 1.0)
    g = get_geq(p.dx, p.dt, rho, vel, density, pressure, None, f)
    return pressure, f, g


def hcz_step(state: dict, p: HCZParams) -> dict:
    """One outer step of the HCZ(+magnetic) demos.

    Reference loop: demo_3d_LBM_Rosensweig_instability.py:151-209,
    demo_2d_LBM_Rosensweig_instability.py:130-188, demo_3d_LBM_multiphase.py:145-182.
    ``state`` keys: f,g,h,rho,vel,density,pressure,flags,mflags,H_ext (cell-centred
    [B,dim,...]),H_ext_mac (list of faces).  Returns a new dict (adds phi, psi, H2, force).
    """
    s = dict(state)
    s["f"] = propagation(s["f"])
    s["g"] = propagation(s["g"])
    s["rho"], s["vel"], s["density"] = macro_compute(
        p.dx, p.dt, s["f"], s["rho"], s["vel"], s["flags"], s["density"],
        p.density_liquid, p.density_gas, p.rho_liquid, p.rho_gas)
    s["f"] = rebounce_obstacle(s["f"], s["flags"])
    s["g"] = rebounce_obstacle(s["g"], s["flags"])
    H2 = None
    phi = None
    if p.magnetic:
        phi = phase_field(s["density"], p)
        h = s["h"]
        H_int = None
        for _ in range(p.n_sweeps):
            h = propagation(h)
            h = rebounce_obstacle(h, s["mflags"])
            if _ == p.n_sweeps - 1:
                s["psi"] = psi_of_h(h)
            H_int, h = get_H_int(p.dt, p.dx, phi, s["mflags"], s["H_ext_mac"], h, p.tau_h)
        s["h"] = h
        Ht = s["H_ext"] + H_int
        H2 = (Ht * Ht).sum(dim=1, keepdim=True)
        s["phi"], s["H2"], s["H_int"] = phi, H2, H_int
    rho_in = s["rho"].clone()
    s["rho"], s["vel"], s["density"], s["pressure"], s["force"], dfai, dprho = capillary_process(
        rho_in, s["vel"], s["flags"], p.dt, p.dx, s["g"], s["density"], s["pressure"], H2, phi,
        kappa=p.kappa, gravity=p.gravity, contact_angle=p.contact_angle,
        density_liquid=p.density_liquid, density_gas=p.density_gas,
        rho_liquid=p.rho_liquid, rho_gas=p.rho_gas)
    s["f"], s["g"] = collision_hcz(p.dx, p.dt, s["f"], s["rho"], s["vel"], s["density"], s["flags"],
                                   s["force"], s["g"], s["pressure"], dfai, dprho,
                                   p.tau_f, p.tau_g, p.KBC_type)
    return s


def single_phase_step(f, rho, vel, flags, force, dx, dt, tau, KBC_type=None):
    """One step of demo_2d_LBM.py:101-121 / demo_2d_LBM_KBC.py (stream, macro, bounce, collide)."""
    f = propagation(f)
    rho, vel = macro_compute(dx, dt, f, rho, vel, flags)
    f = rebounce_obstacle(f, flags)
    if KBC_type is None:
        f = collision_bgk(dx, dt, f, rho, vel, flags, force, tau)
    else:
        f = collision_kbc(dx, dt, f, rho, vel, flags, force, tau, KBC_type)
    return f, rho, vel
