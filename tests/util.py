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
