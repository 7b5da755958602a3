"""CPU checks of the N4 extension oracle (oracle/ext_oracle.py) and of the library's lattice tables.

The reference has no D3Q27 / Guo / esoteric-pull, so the extension oracle is pinned to the reference-pinned
oracle where they overlap (D3Q19 + velocity-shift forcing) and to the defining identities elsewhere.
"""
import ctypes

import pytest
import torch

from oracle import ext_oracle as X
from oracle import lbm_oracle as O


def _state(lat, res=(6, 7, 9), seed=0, dtype=torch.float64, walls=True):
    g = torch.Generator().manual_seed(seed)
    B = 2
    rho = 0.3 + 0.05 * torch.rand((B, 1, *res), generator=g, dtype=dtype)
    vel = 0.04 * (torch.rand((B, 3, *res), generator=g, dtype=dtype) - 0.5)
    f = O.get_feq(1.0, 1.0, rho, vel, None, 1.0, lat=lat)
    f = f * (1.0 + 0.02 * (torch.rand(f.shape, generator=g, dtype=dtype) - 0.5))
    flags = torch.full((B, 1, *res), O.FLUID, dtype=torch.uint8)
    if walls:
        flags[:, :, 0] = O.OBSTACLE
        flags[:, :, :, -1] = O.OBSTACLE
        flags[:, :, 2:4, 3:5, 4:6] = O.OBSTACLE
    force = 1e-3 * (torch.rand((B, 3, *res), generator=g, dtype=dtype) - 0.5)
    return f, rho, vel, flags, force


def test_lattice_tables_match_library():
    from lbm_ferrofluid_b200.extensions import lattice_tables
    for name, lat in (("D3Q19", X.D3Q19), ("D3Q27", X.D3Q27)):
        e, opp, w = lattice_tables(name)
        assert [tuple(v) for v in e] == [tuple(v) for v in lat.e]
        assert list(opp) == list(lat.opp)
        w32 = torch.tensor(lat.w, dtype=torch.float32).double().tolist()
        assert w == pytest.approx(w32, abs=0)


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
def test_lattice_moments(lat):
    e = torch.tensor(lat.e, dtype=torch.float64)
    w = torch.tensor(lat.w, dtype=torch.float64)
    assert abs(w.sum().item() - 1.0) < 1e-15
    assert torch.allclose((w[:, None] * e).sum(0), torch.zeros(3, dtype=torch.float64), atol=1e-16)
    second = torch.einsum("q,qa,qb->ab", w, e, e)
    assert torch.allclose(second, torch.eye(3, dtype=torch.float64) / 3.0, atol=1e-15)
    fourth = torch.einsum("q,qa,qb,qc,qd->abcd", w, e, e, e, e)
    d = torch.eye(3, dtype=torch.float64)
    iso = (torch.einsum("ab,cd->abcd", d, d) + torch.einsum("ac,bd->abcd", d, d) + torch.einsum("ad,bc->abcd", d, d)) / 9.0
    assert torch.allclose(fourth, iso, atol=1e-15)
    for q in range(lat.Q):
        assert tuple(-c for c in lat.e[q]) == tuple(lat.e[lat.opp[q]])
    if lat.Q == 27:  # the D3Q27 weights are the tensor product of the 1-D weights (2/3, 1/6, 1/6)
        w1 = {0: 2 / 3, 1: 1 / 6, -1: 1 / 6}
        for q in range(27):
            assert abs(lat.w[q] - w1[lat.e[q][0]] * w1[lat.e[q][1]] * w1[lat.e[q][2]]) < 1e-16
        assert [tuple(v) for v in lat.e[:19]] == [tuple(v) for v in O.D3Q19.e]


def test_d3q19_shift_step_is_the_pinned_single_phase_step():
    f, rho, vel, flags, force = _state(X.D3Q19)
    a = X.ext_step(f, rho, vel, flags, force, 1.0, 1.0, 0.8, X.D3Q19, "shift")
    b = O.single_phase_step(f, rho, vel, flags, force, 1.0, 1.0, 0.8)
    for x, y, name in zip(a, b, ("f", "rho", "vel")):
        assert (x - y).abs().max().item() < 1e-14, name


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
def test_feq_moments_on_both_lattices(lat):
    _, rho, vel, _, _ = _state(lat, walls=False)
    feq = O.get_feq(1.0, 1.0, rho, vel, None, 1.0, lat=lat)
    e = X._e_tensor(lat, feq)
    assert (feq.sum(1, keepdim=True) - rho).abs().max().item() < (2e-8 if lat.Q == 27 else 1e-5)  # fp32-rounded weights, like the reference's
    mom = (feq.unsqueeze(2) * e).sum(1)
    assert (mom - rho * vel).abs().max().item() < (2e-8 if lat.Q == 27 else 1e-5)  # fp32-rounded weights, like the reference's


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
def test_guo_step_mass_and_momentum_balance(lat):
    f, rho, vel, flags, force = _state(lat, walls=False)
    tau, dt = 0.7, 1.0
    fo, rho1, vel1 = X.ext_step(f, rho, vel, flags, force, 1.0, dt, tau, lat, "guo")
    e = X._e_tensor(lat, f)
    assert abs((fo.sum() / f.sum()).item() - 1.0) < (5e-8 if lat.Q == 27 else 1e-5)  # weight rounding / D3Q19 feq defect
    dmom = (fo.unsqueeze(2) * e).sum(1).sum(dim=(0, 2, 3, 4)) - (f.unsqueeze(2) * e).sum(1).sum(dim=(0, 2, 3, 4))
    want = dt * force.sum(dim=(0, 2, 3, 4))
    # D3Q19's product equilibrium carries an O(u^3) momentum defect; D3Q27's is exact
    assert (dmom - want).abs().max().item() < (1e-6 if lat.Q == 27 else 1e-4)
    # physical velocity = (sum f e + F dt / 2) / rho
    fs = X.propagation(f, lat)
    u = ((fs.unsqueeze(2) * e).sum(1) + 0.5 * dt * force) / fs.sum(1, keepdim=True)
    assert (vel1 - u).abs().max().item() < 1e-15


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
@pytest.mark.parametrize("parity", [0, 1])
def test_esoteric_layout_semantics(lat, parity):
    """Walking the esoteric-pull load rule over the layout returns exactly the pulled populations."""
    f, *_ = _state(lat, res=(4, 5, 6))
    ep = X.esoteric_layout(f, lat, parity)
    pulled = X.propagation(f, lat)
    got = torch.empty_like(f)
    got[:, 0] = ep[:, 0]
    for p in range(1, lat.Q):
        o = lat.opp[p]
        if p > o:
            continue
        axes = [a for a in range(3) if lat.e[p][a] != 0]
        at_nbr = lambda t: torch.roll(t, shifts=[-lat.e[p][a] for a in axes], dims=[-1 - a for a in axes])  # value at x + e_p
        if parity == 0:
            got[:, p] = ep[:, o]
            got[:, o] = at_nbr(ep[:, p])
        else:
            got[:, p] = ep[:, p]
            got[:, o] = at_nbr(ep[:, o])
    assert torch.equal(got, pulled)


def test_library_exports_extension_symbols():
    from lbm_ferrofluid_b200 import _lib
    lib = _lib.load()
    for name in ("lbmff_ext_stream_collide_d3q19", "lbmff_ext_stream_collide_d3q27",
                 "lbmff_ext_esoteric_convert_d3q19", "lbmff_ext_esoteric_convert_d3q27", "lbmff_ext_lattice_info"):
        assert hasattr(lib, name)
    assert lib.lbmff_ext_lattice_info(13, ctypes.c_void_p(0), ctypes.c_void_p(0), ctypes.c_void_p(0)) == -1


def test_kbc3_split_properties():
    """ds carries exactly the second-order moment difference, dh none of it; both are mass- and momentum-free;
    gamma = 2 is BGK; the entropic gamma stays near 2 close to equilibrium."""
    lat = X.D3Q27
    f, rho, vel, flags, force = _state(lat, walls=False)
    fs = X.propagation(f, lat)
    e = X._e_tensor(lat, fs)
    rho = fs.sum(1, keepdim=True)
    u = (fs.unsqueeze(2) * e).sum(1) / rho
    feq = O.get_feq(1.0, 1.0, rho, u, None, 1.0, lat=lat)
    tau = 0.6
    post, ds, dh = X.kbc3_relax(fs, feq, rho, u, tau, lat)
    for part in (ds, dh):
        assert part.sum(1).abs().max().item() < 1e-8          # fp32-rounded weights in feq: 1e-9, not 1e-16
        assert (part.unsqueeze(2) * e).sum(1).abs().max().item() < 1e-8
    second = lambda t: torch.einsum("bqzyx,qi,qj->bijzyx", t, e.reshape(27, 3), e.reshape(27, 3))
    assert (second(ds) - second(fs - feq)).abs().max().item() < 1e-8
    assert second(dh).abs().max().item() < 1e-8
    bgk = (1.0 - 1.0 / tau) * fs + feq / tau
    assert (X.kbc3_relax(fs, feq, rho, u, tau, lat, gamma_override=2.0)[0] - bgk).abs().max().item() < 1e-15
    assert (post.sum(1, keepdim=True) - rho).abs().max().item() < 1e-8
    assert ((post.unsqueeze(2) * e).sum(1) - rho * u).abs().max().item() < 1e-8
    # centring: the second-moment DIFFERENCE does not depend on the centre when mass and momentum agree
    post0 = X.kbc3_relax(fs, feq, rho, torch.zeros_like(u), tau, lat)[0]
    assert (post0 - post).abs().max().item() < 1e-7


def test_kbc3_step_runs_and_conserves():
    lat = X.D3Q27
    f, rho, vel, flags, force = _state(lat)
    m0 = f.sum().item()
    for _ in range(4):
        f, rho, vel = X.ext_step(f, rho, vel, flags, force, 1.0, 1.0, 0.55, lat, None, "kbc")
    assert torch.isfinite(f).all()
    assert abs(f.sum().item() / m0 - 1.0) < 1e-7


def test_sc_force_3d_reduces_to_the_pinned_2d_force():
    """A field that does not vary along z: the x, y components of the D3Q19 force equal the pinned D2Q9 force
    (the in-plane D3Q19 weights, summed over the z neighbours, are the D2Q9 weights) and F_z vanishes."""
    g = torch.Generator().manual_seed(4)
    d2 = 0.05 + 0.2 * torch.rand((1, 1, 9, 11), generator=g, dtype=torch.float64)
    fl2 = torch.ones((1, 1, 9, 11), dtype=torch.uint8)
    fl2[:, :, 3:5, 4:7] = O.OBSTACLE
    F2 = O.sc_force(1.0, 1.0, d2, fl2, 0.265, 0.038)
    d3 = d2.unsqueeze(2).repeat(1, 1, 7, 1, 1)
    fl3 = fl2.unsqueeze(2).repeat(1, 1, 7, 1, 1)
    F3 = X.sc_force_3d(1.0, 1.0, d3, fl3, 0.265, 0.038)
    mid = F3[:, :, 3]                                     # away from the zero-padded z ends
    assert (mid[:, :2] - F2).abs().max().item() < 1e-7   # fp32-rounded weights
    assert mid[:, 2].abs().max().item() < 1e-12


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
def test_mrt3_split_properties(lat):
    """The stress pieces of the MRT split are mass- and momentum-free; ds_dev carries exactly the traceless part of
    the second-moment difference, ds_bulk exactly its trace (isotropically), dh none of it; equal rates are BGK;
    on D3Q27, tau_high = 1 / (gamma beta) with gamma fixed reproduces kbc3_relax."""
    f, rho, vel, flags, force = _state(lat, walls=False)
    fs = X.propagation(f, lat)
    rho = fs.sum(1, keepdim=True)
    e = X._e_tensor(lat, fs)
    vel = (fs.unsqueeze(2) * e).sum(1) / rho
    feq = O.get_feq(1.0, 1.0, rho, vel, None, 1.0, lat=lat)
    dev, bulk, dh = X.mrt3_split(fs, feq, rho, vel, lat)
    cen = e - vel.unsqueeze(1)   # moments about the velocity, like the split itself (on D3Q19 f - feq carries feq's
    second = lambda t: torch.einsum("bqzyx,bqizyx,bqjzyx->bijzyx", t, cen, cen)   # small mass / momentum defect)
    for t in (dev, bulk):   # (dh keeps whatever mass / momentum defect f - feq itself has: feq's, on D3Q19)
        assert t.sum(1).abs().max().item() < 1e-15
        assert (t.unsqueeze(2) * e).sum(1).abs().max().item() < 1e-15
    eye = torch.eye(3, dtype=f.dtype).view(1, 3, 3, 1, 1, 1)
    P = second(fs - feq)
    tr = (P * eye).sum(dim=(1, 2), keepdim=True)
    assert (second(dev) - (P - eye * tr / 3.0)).abs().max().item() < 1e-15
    assert (second(bulk) - eye * tr / 3.0).abs().max().item() < 1e-15
    assert second(dh).abs().max().item() < 1e-15
    tau = 0.7
    bgk = (1.0 - 1.0 / tau) * fs + feq / tau
    assert (X.mrt3_relax(fs, feq, rho, vel, tau, tau, tau, lat) - bgk).abs().max().item() < 1e-15
    if lat.Q == 27:
        gamma, beta = 1.6, 0.5 / tau
        kbc = X.kbc3_relax(fs, feq, rho, vel, tau, lat, gamma_override=gamma)[0]
        assert (X.mrt3_relax(fs, feq, rho, vel, tau, tau, 1.0 / (gamma * beta), lat) - kbc).abs().max().item() < 1e-15


@pytest.mark.parametrize("lat", [X.D3Q19, X.D3Q27], ids=["D3Q19", "D3Q27"])
@pytest.mark.parametrize("tau_bulk,tau_high", [(0.6, 0.6), (1.0, 1.0), (0.9, 1.4)])
def test_mrt_shear_wave_decays_with_the_viscosity_of_tau_alone(lat, tau_bulk, tau_high):
    """u_x(y) = U sin(2 pi y / L) in a periodic box decays as exp(-nu k^2 t), nu = cs2 (tau - 1/2) dt, whatever the
    bulk and higher-order rates are."""
    import math
    tau, L, U, steps = 0.6, 32, 1e-3, 120
    res = (2, L, 2)
    flags = torch.full((1, 1, *res), O.FLUID, dtype=torch.uint8)
    rho = torch.full((1, 1, *res), 0.25, dtype=torch.float64)
    vel = torch.zeros((1, 3, *res), dtype=torch.float64)
    y = torch.arange(L, dtype=torch.float64).view(1, L, 1)
    vel[0, 0] = U * torch.sin(2 * math.pi * y / L)
    f = O.get_feq(1.0, 1.0, rho, vel, None, 1.0, lat=lat)
    force = torch.zeros_like(vel)
    for _ in range(steps):
        f, rho, vel = X.ext_step(f, rho, vel, flags, force, 1.0, 1.0, tau, lat, None, "mrt", tau_bulk, tau_high)
    amp = (vel[0, 0] * torch.sin(2 * math.pi * y / L)).mean().item() * 2.0
    k2 = (2 * math.pi / L) ** 2
    want = U * math.exp(-(tau - 0.5) / 3.0 * k2 * steps)
    assert abs(amp / want - 1.0) < 5e-3
