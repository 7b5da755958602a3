"""Shared helpers for the parity tests (test infrastructure)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = {}
    for k in z.files:
        a = z[k]
        out[k] = torch.from_numpy(a) if a.ndim > 0 else a.item()
    return out


def rel_l2(a, b):
    a = a.double().cpu()
    b = b.double().cpu()
    den = b.norm().item()
    return (a - b).norm().item() / (den if den > 0 else 1.0)


def rel_linf(a, b):
    """max |a-b| scaled by max |b| (scale-relative, robust to sign changes/zeros)."""
    a = a.double().cpu()
    b = b.double().cpu()
    den = b.abs().max().item()
    return (a - b).abs().max().item() / (den if den > 0 else 1.0)


def assert_close(a, b, tol, what=""):
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    assert torch.isfinite(a.double()).all(), f"{what}: non-finite values"
    e = rel_linf(a, b)
    assert e <= tol, f"{what}: scaled max error {e:.3e} > {tol:.1e}"


# ---------------------------------------------------------------------------------------------------
# Parity bar for floating-point trajectories.
#
# Two fp32 evaluations of the same formulas differ by rounding; how much that matters is field dependent
# (vel is a small difference of O(1e-2) populations).  The reference's own fp32 result is therefore only
# defined up to ITS distance from the exact (fp64) evaluation of the same algorithm.  Bar, per field:
#     err(ours, fp64 truth)  <=  FLOOR_FACTOR * err(reference fp32, fp64 truth) + ABS_FLOOR
# i.e. we must be as accurate as the reference's fp32 path, within a factor 2.  rel-L2 norms.
# ---------------------------------------------------------------------------------------------------
FLOOR_FACTOR = 2.0
ABS_FLOOR = 2e-6


def assert_within_fp32_floor(ours, ref32, truth64, what, mask=None, factor=FLOOR_FACTOR, abs_floor=ABS_FLOOR):
    ours, ref32, truth64 = ours.detach().cpu().double(), ref32.detach().cpu().double(), truth64.detach().cpu().double()
    assert torch.isfinite(ours).all(), f"{what}: non-finite values"
    if mask is not None:
        ours, ref32, truth64 = ours * mask, ref32 * mask, truth64 * mask
    e_ref = rel_l2(ref32, truth64)
    e_our = rel_l2(ours, truth64)
    bar = factor * e_ref + abs_floor
    assert e_our <= bar, f"{what}: rel-L2 to fp64 truth {e_our:.3e} > {bar:.3e} (reference fp32 floor {e_ref:.3e})"
    return e_our, e_ref


def oracle_hcz_trajectory(G, dtype=torch.float32, steps=None):
    """Run the CPU oracle on the configuration stored in a traj_* golden (same init as make_golden.traj_hcz)."""
    from oracle import lbm_oracle as O

    dim = len(G["res"])
    res = tuple(int(v) for v in G["res"])
    p = O.HCZParams(kappa=G["kappa"], tau_f=G["tau_f"], tau_g=G["tau_g"], gravity=G["gravity"],
                    contact_angle=torch.tensor([G["contact_angle"]]).to(dtype), n_sweeps=int(G["n_sweeps"]),
                    KBC_type=O.KBC_A if G["kbc"] else None, magnetic=bool(G["magnetic"]))
    rho = G["rho0"].to(dtype).clone()
    density = rho.clone()
    vel = torch.zeros((1, dim, *res), dtype=dtype)
    pressure, f, g = O.hcz_init(rho, density, vel, p)
    Hext = torch.zeros((1, dim, *res), dtype=dtype)
    Hext[:, 1] = G["mag"]
    s = dict(f=f, g=g, h=torch.zeros_like(f), rho=rho, vel=vel, density=density, pressure=pressure,
             flags=G["flags"], mflags=G["mflags"], H_ext=Hext, H_ext_mac=O.get_staggered(Hext))
    for _ in range(int(G["steps"]) if steps is None else steps):
        s = O.hcz_step(s, p)
    return s


# ---------------------------------------------------------------------------------------------------
# Converged magnetic potential (lbmff_poisson_cg_*): the checker is the reference's own sweep.
# ---------------------------------------------------------------------------------------------------
def poisson_case(res, seed=0, interior_wall=True):
    from oracle.make_golden import shell_flags

    g = torch.Generator().manual_seed(seed)
    dim = len(res)
    mflags = shell_flags(res, open_axes=(1,))       # periodic along y, walls on the other faces
    if interior_wall:
        sl = (0, 0) + tuple(slice(n // 3, n // 3 + 3) for n in res)
        mflags[sl] = 2
    rhs = torch.randn((1, 1, *res), generator=g, dtype=torch.float32) * 0.05
    # a smooth dipole-like part on top of the noise: the physical source is smooth
    ax = [torch.linspace(-1, 1, n) for n in res]
    grid = torch.meshgrid(*ax, indexing="ij")
    rhs[0, 0] += 0.5 * torch.exp(-8 * sum(x * x for x in grid)) * grid[-2]
    rhs[mflags != 1] = 0.0                           # lbmff_magnetic_source_* writes 0 off FLUID cells
    return mflags, rhs, dim


def poisson_b_and_drift(rhs, mflags, dx=1.0, dt=1.0):
    """fp64, by rolls: b = sum_q w_q S(fluid neighbour, else the cell itself) / (1 - w0), the wall-weight fraction
    omega = sum_{q: wall neighbour} w_q / (1 - w0), and the constant c = sum b / sum (1 + omega) that every sweep of
    the stationary state adds to psi (wall-returned populations are one sweep older, hence the 1 + omega)."""
    O = _O()
    rhs = rhs.double()
    lat = O.lattice_for(rhs)
    w = O._wt(lat, rhs)
    S = dt * rhs * ((dx / dt) ** 2 / 3.0 * (0.5 - 1.0) * dt)
    fluid = mflags != 2
    b = torch.zeros_like(S)
    omega = torch.zeros_like(S)
    axes = list(range(lat.dim))
    for q in range(1, lat.Q):
        roll = dict(shifts=[lat.e[q][a] for a in axes], dims=[-1 - a for a in axes])
        nb_fluid = torch.roll(fluid, **roll)
        b += w[:, q:q + 1] * torch.where(nb_fluid, torch.roll(S, **roll), S)
        omega += w[:, q:q + 1] * (~nb_fluid).double()
    inv = 1.0 / (1.0 - w[:, 0:1])
    b, omega = torch.where(fluid, b * inv, 0 * b), torch.where(fluid, omega * inv, 0 * b)
    c = b.flatten(1).sum(1) / (fluid.double() * (1.0 + omega)).flatten(1).sum(1)
    return b, omega, c


def one_oracle_sweep(psi, rhs, mflags, dx=1.0, dt=1.0, drift=0.0):
    """fp64: post-collision populations of the stationary state built from psi, then ONE sweep of the reference
    (propagation, rebounce_obstacle, psi = sum h / (1 - w0)); returns psi after the sweep.  ``drift`` is the
    constant the stationary state gains per sweep: what a wall holds left its cell one sweep earlier."""
    psi, rhs = psi.double(), rhs.double()
    lat = _O().lattice_for(torch.empty((1, 1) + tuple(psi.shape[2:])))
    w = _O()._wt(lat, psi)
    cs2 = (dx / dt) ** 2 / 3.0
    S = dt * rhs * (cs2 * (0.5 - 1.0) * dt)
    A = psi + S
    fluid = mflags != 2
    h = torch.where(fluid, w * A, torch.zeros_like(w * A))
    h[:, 0:1] = torch.where(fluid, h[:, 0:1] - psi, torch.zeros_like(psi))
    # a wall cell s holds, in slot q, what the fluid cell s + e_q sent into it one sweep ago (already swapped by
    # rebounce_obstacle): w_q A(s + e_q); it streams back to s + e_q in this sweep
    for q in range(1, lat.Q):
        axes = list(range(lat.dim))
        src = torch.roll(torch.where(fluid, A - drift, torch.zeros_like(A)), shifts=[-lat.e[q][a] for a in axes],
                         dims=[-1 - a for a in axes])
        h[:, q:q + 1] = torch.where(fluid, h[:, q:q + 1], w[:, q:q + 1] * src)
    # the sweep: stream, bounce; the bounce of THIS sweep only matters for the next one
    hs = _O().propagation(h)
    return _O().psi_of_h(_O().rebounce_obstacle(hs, mflags))


def _O():
    from oracle import lbm_oracle
    return lbm_oracle
