"""CPU oracle for the SURVEY 8(f) N4 extensions (D3Q27, Guo forcing, esoteric-pull streaming).

TEST INFRASTRUCTURE ONLY -- imported by ``tests/`` alone, never by the product path.

**Parity unpinned against the reference**: the reference has none of these (its only 3-D lattice is D3Q19,
``LBM_collision_3d.py:76-103``; its only forcing is the velocity shift of ``LBM_collision_3d.py:141-143``; its
streaming is out of place, ``LBM_propagation_3d.py:18-111``).  What this module is pinned to instead:

* every building block is the pinned oracle's (``oracle/lbm_oracle.py``): streaming = periodic roll, full-way
  bounce-back, the entropic product equilibrium ``get_feq`` (which takes the lattice as an argument);
* with ``lat = D3Q19`` and ``forcing = "shift"``, :func:`ext_step` IS ``lbm_oracle.single_phase_step``
  (``tests/test_ext_oracle.py`` checks equality), so the D3Q19 leg inherits the reference pin;
* the Guo source term is the published one (Guo, Zheng & Shi 2002):
  ``S_q = (1 - 1/2tau) w_q [ (e_q - u) / cs2 + (e_q . u) e_q / cs2^2 ] . F`` with ``u = (sum f e + F dt/2) / rho``;
  its zeroth and first moments (0 and ``(1 - 1/2tau) F``) are checked in the CPU tests.
"""
from __future__ import annotations

import torch

from . import lbm_oracle as O

# D3Q27: the reference's D3Q19 order first, then the eight corners as opposite pairs
# (mirrors lbm_ferrofluid_b200/csrc/ext_lattice.cu::Q27; tests compare both against lbmff_ext_lattice_info).
_CORNERS = [(1, 1, 1), (-1, -1, -1), (1, 1, -1), (-1, -1, 1), (1, -1, 1), (-1, 1, -1), (-1, 1, 1), (1, -1, -1)]


def _w27(e):
    n = sum(abs(c) for c in e)
    return (8 / 27, 2 / 27, 1 / 54, 1 / 216)[n]


D3Q27 = O._make_lattice(list(O.D3Q19.e) + _CORNERS, [_w27(e) for e in list(O.D3Q19.e) + _CORNERS])
D3Q19 = O.D3Q19


def lattice(Q: int) -> O.Lattice:
    return {19: D3Q19, 27: D3Q27}[Q]


def propagation(f, lat):
    """Periodic pull streaming on any lattice (lbm_oracle.propagation restated with an explicit lattice)."""
    out = torch.empty_like(f)
    for q in range(lat.Q):
        axes = [a for a in range(lat.dim) if lat.e[q][a] != 0]
        out[:, q] = torch.roll(f[:, q], shifts=[lat.e[q][a] for a in axes], dims=[-1 - a for a in axes]) if axes \
            else f[:, q]
    return out


def _e_tensor(lat, like):
    return torch.tensor(lat.e, dtype=like.dtype).reshape(1, lat.Q, lat.dim, *([1] * lat.dim))


def kbc3_relax(f, feq, rho, uc, tau, lat, gamma_override=None):
    """Entropic multi-relaxation on D3Q27: the reference's 2-D KBC (LBM_collision_MRT_2d.py:113-151, 184-226:
    s = {T, N, Pi} of the centred second moments, h = the rest, gamma from the entropic estimate with weights
    1/feq) carried to 3-D through the D3Q27 moment representation (Karlin, Boesch & Chikatamarla 2014):
    rest -rho T; axes rho/6 (2 N_a - N_b + T) ...; planar diagonals rho/4 s l Pi; corners 0.
    uc: centring velocity in lattice units.  gamma = 2 (BGK) where dh vanishes (the reference's 2-D code
    returns 0/0 there; this extension does not keep that quirk)."""
    assert lat.Q == 27
    e = _e_tensor(lat, f)
    d = ((f - feq) / rho).unsqueeze(2)                       # [B,Q,1,...]
    cen = e - uc.unsqueeze(1)                                # [B,Q,3,...]
    M = lambda a, b: (d[:, :, 0] * cen[:, :, a] * cen[:, :, b]).sum(dim=1, keepdim=True)
    xx, yy, zz, xy, xz, yz = M(0, 0), M(1, 1), M(2, 2), M(0, 1), M(0, 2), M(1, 2)
    T, Nxz, Nyz = xx + yy + zz, xx - zz, yy - zz
    ds = torch.zeros_like(f)
    for q in range(27):
        ex, ey, ez = lat.e[q]
        n = abs(ex) + abs(ey) + abs(ez)
        if n == 0:
            v = -rho * T
        elif n == 1:
            v = rho / 6.0 * ((2 * Nxz - Nyz + T) if ex else ((2 * Nyz - Nxz + T) if ey else (T - Nxz - Nyz)))
        elif n == 2:
            v = rho / 4.0 * ((ex * ey * xy) if ez == 0 else ((ex * ez * xz) if ey == 0 else (ey * ez * yz)))
        else:
            continue
        ds[:, q:q + 1] = v
    dh = (f - feq) - ds
    beta = 0.5 / tau
    num = (ds * dh / feq).sum(dim=1, keepdim=True)
    den = (dh * dh / feq).sum(dim=1, keepdim=True)
    gamma = torch.where(den > 0, 1.0 / beta - (2.0 - 1.0 / beta) * num / torch.where(den > 0, den, torch.ones_like(den)),
                        torch.full_like(den, 2.0))
    if gamma_override is not None:
        gamma = torch.full_like(den, gamma_override)
    return f - beta * (2.0 * ds + gamma * dh), ds, dh


def mrt3_split(f, feq, rho, uc, lat):
    """Moment split of f - feq used by the three-rate MRT collision (D3Q19 and D3Q27): the same {T, N, Pi}
    representation as kbc3_relax, with the stress part separated into its deviatoric (N, Pi: shear) and trace (T:
    bulk) pieces; dh is everything else (third- and higher-order moments).  Only rest, axis and planar-diagonal
    populations carry ds, so the split exists on both lattices.  Returns (ds_dev, ds_bulk, dh)."""
    e = _e_tensor(lat, f)
    d = ((f - feq) / rho).unsqueeze(2)
    cen = e - uc.unsqueeze(1)
    M = lambda a, b: (d[:, :, 0] * cen[:, :, a] * cen[:, :, b]).sum(dim=1, keepdim=True)
    xx, yy, zz, xy, xz, yz = M(0, 0), M(1, 1), M(2, 2), M(0, 1), M(0, 2), M(1, 2)
    T, Nxz, Nyz = xx + yy + zz, xx - zz, yy - zz
    dev, bulk = torch.zeros_like(f), torch.zeros_like(f)
    for q in range(lat.Q):
        ex, ey, ez = lat.e[q]
        n = abs(ex) + abs(ey) + abs(ez)
        if n == 0:
            bulk[:, q:q + 1] = -rho * T
        elif n == 1:
            bulk[:, q:q + 1] = rho / 6.0 * T
            dev[:, q:q + 1] = rho / 6.0 * ((2 * Nxz - Nyz) if ex else ((2 * Nyz - Nxz) if ey else (-Nxz - Nyz)))
        elif n == 2:
            dev[:, q:q + 1] = rho / 4.0 * ((ex * ey * xy) if ez == 0 else ((ex * ez * xz) if ey == 0 else (ey * ez * yz)))
    return dev, bulk, (f - feq) - dev - bulk


def mrt3_relax(f, feq, rho, uc, tau, tau_bulk, tau_high, lat):
    """Three-rate MRT: f - ds_dev / tau - ds_bulk / tau_bulk - dh / tau_high.  tau sets the shear viscosity
    cs2 (tau - 1/2) dt exactly as in BGK; tau_bulk the bulk viscosity; tau_high damps the non-hydrodynamic
    moments.  All three equal is BGK; tau_high = 1 / (gamma beta) with the entropic gamma is kbc3_relax."""
    dev, bulk, dh = mrt3_split(f, feq, rho, uc, lat)
    return f - dev / tau - bulk / tau_bulk - dh / tau_high


def ext_step(f, rho, vel, flags, force, dx, dt, tau, lat, forcing="shift", collision="bgk", tau_bulk=None,
             tau_high=None):
    """stream -> macro -> bounce -> BGK / KBC / MRT; forcing in {None, "shift", "guo"}.  Returns (f, rho, vel)."""
    def relax(f, feq, rho, uc):
        if collision == "kbc":
            return kbc3_relax(f, feq, rho, uc, tau, lat)[0]
        if collision == "mrt":
            return mrt3_relax(f, feq, rho, uc, tau, tau if tau_bulk is None else tau_bulk,
                              tau if tau_high is None else tau_high, lat)
        return (1.0 - 1.0 / tau) * f + feq / tau

    c = dx / dt
    cs2 = c * c / 3.0
    f = propagation(f, lat)
    obst = flags == O.OBSTACLE
    e = _e_tensor(lat, f)
    rho_new = f.sum(dim=1, keepdim=True)
    mom = (f.unsqueeze(2) * e).sum(dim=1) * c
    rho = torch.where(obst, rho, rho_new)
    f = torch.where(obst, f[:, list(lat.opp)], f)
    w = O._wt(lat, f)
    if forcing == "guo":
        u = (mom + 0.5 * dt * force) / rho_new
        vel = torch.where(obst, vel, u)
        feq = O.get_feq(dx, dt, rho, vel, None, tau, lat=lat)
        ec = e * c
        eF = (ec * force.unsqueeze(1)).sum(dim=2)
        eu = (ec * vel.unsqueeze(1)).sum(dim=2)
        uF = (vel * force).sum(dim=1, keepdim=True)
        S = (1.0 - 0.5 / tau) * w * ((eF - uF) / cs2 + eu * eF / cs2 / cs2)
        post = relax(f, feq, rho, vel / c) + dt * S
    else:
        vel = torch.where(obst, vel, mom / rho_new)
        feq = O.get_feq(dx, dt, rho, vel, force if forcing == "shift" else None, tau, lat=lat)
        post = relax(f, feq, rho, vel / c)
    return torch.where(obst, f, post), rho, vel


def esoteric_layout(f_std, lat, parity):
    """Slot layout of an esoteric-pull buffer before the load of a step of the given parity (Lehmann 2022):
    parity 0: slot p @ x = f_o(x), slot o @ x+e_p = f_p(x);  parity 1: slot o @ x = f_o(x), slot p @ x+e_p = f_p(x)."""
    out = torch.empty_like(f_std)
    out[:, 0] = f_std[:, 0]
    for p in range(1, lat.Q):
        o = lat.opp[p]
        if p > o:
            continue
        axes = [a for a in range(lat.dim) if lat.e[p][a] != 0]
        moved = torch.roll(f_std[:, p], shifts=[lat.e[p][a] for a in axes], dims=[-1 - a for a in axes])
        if parity == 0:
            out[:, p] = f_std[:, o]
            out[:, o] = moved
        else:
            out[:, o] = f_std[:, o]
            out[:, p] = moved
    return out


def sc_force_3d(dx, dt, density, flags, density_liquid, density_gas):
    """lbm_oracle.sc_force (LBM_collision_SC_2d.py:22-97) with the D3Q19 lattice and a 3-D zero pad."""
    import math
    import torch.nn.functional as F
    lat = D3Q19
    c = dx / dt
    cs2 = c * c / 3.0
    RT, a, b, G1 = cs2, 12.0 * cs2, 4.0, -1.0 / 3.0

    def psi_of(d, sqrt, power):
        t = d * b / 4.0
        v = RT * (1.0 + (4.0 * t - 2.0 * t * t) / power(1.0 - t, 3)) - a * d - cs2
        return sqrt(2.0 * d * v / G1 / cs2)

    psx = psi_of(density, torch.sqrt, torch.pow)
    wall = psi_of(0.5 * (density_gas + density_liquid), math.sqrt, math.pow)
    pad = F.pad(torch.where(flags == O.OBSTACLE, wall * torch.ones_like(psx), psx), pad=(1, 1, 1, 1, 1, 1))
    w = O._wt(lat, density)
    D, Hh, W = density.shape[2:]
    comps = []
    for ax in range(3):
        acc = torch.zeros_like(psx)
        for q in range(1, lat.Q):
            ex, ey, ez = lat.e[q]
            if lat.e[q][ax] == 0:
                continue
            nb = pad[..., 1 + ez:1 + ez + D, 1 + ey:1 + ey + Hh, 1 + ex:1 + ex + W]
            acc = acc + w[:, q:q + 1] * lat.e[q][ax] * nb
        comps.append(acc)
    force = -G1 * psx * c * torch.cat(comps, dim=1)
    return torch.where(flags == O.FLUID, force, torch.zeros_like(force))
