"""CPU check of the linear system behind lbmff_poisson_cg_* (csrc/poisson_cg.cu header): the solution of
L psi = b, with L the wall-aware weighted graph Laplacian and b the weighted neighbour sum of the source, is a fixed
point (up to the constant drift) of the reference's sweep as the oracle executes it -- and the sweeps converge to it."""
import numpy as np
import pytest
import torch

from oracle import lbm_oracle as O
from tests.util import one_oracle_sweep, poisson_b_and_drift, poisson_case


def dense_system(mflags, rhs, dx=1.0, dt=1.0):
    """fp64 dense L and b exactly as k_cg_apply / k_cg_rhs form them (fluid cells only, periodic indices)."""
    res = tuple(mflags.shape[2:])
    lat = O.D2Q9 if len(res) == 2 else O.D3Q19
    w = [float(np.float32(x)) for x in lat.w]        # the kernels use the fp32-rounded weights
    fl = (mflags[0, 0] != 2).numpy()
    S = (dt * rhs[0, 0].double().numpy() * ((dx / dt) ** 2 / 3.0 * (0.5 - 1.0) * dt))
    idx = -np.ones(res, dtype=np.int64)
    cells = np.argwhere(fl)
    idx[fl] = np.arange(len(cells))
    n = len(cells)
    L = np.zeros((n, n))
    b = np.zeros(n)
    omega = np.zeros(n)
    inv = 1.0 / (1.0 - w[0])
    for r, c in enumerate(cells):
        for q in range(1, lat.Q):
            e = lat.e[q][::-1]                         # lattice vectors are (x, y[, z]); arrays are [(z,) y, x]
            nb = tuple((c[a] - e[a]) % res[a] for a in range(len(res)))
            if fl[nb]:
                L[r, r] += w[q] * inv
                L[r, idx[nb]] -= w[q] * inv
                b[r] += w[q] * inv * S[nb]
            else:
                b[r] += w[q] * inv * S[tuple(c)]
                omega[r] += w[q] * inv
    # the stationary state of the sweeps gains c per sweep; populations returned by a wall are one sweep older
    c = b.sum() / (1.0 + omega).sum()
    return L, b - c * (1.0 + omega), fl, c


@pytest.mark.parametrize("res", [(12, 10), (7, 8, 6)])
def test_fixed_point_of_the_oracle_sweep_solves_the_graph_laplacian_system(res):
    mflags, rhs, dim = poisson_case(res, seed=4)
    L, b, fl, c = dense_system(mflags, rhs)
    assert np.allclose(L, L.T) and np.abs(L.sum(1)).max() < 1e-12          # symmetric, constants in the null space
    assert abs(b.sum()) < 1e-12                                             # the projected b is in the range
    sol = np.linalg.lstsq(L, b, rcond=None)[0]
    assert np.abs(L @ sol - b).max() < 1e-10
    _, _, c_rolls = poisson_b_and_drift(rhs, mflags)
    assert abs(c_rolls.item() - c) < 1e-12
    psi = torch.zeros((1, 1, *res), dtype=torch.float64)
    psi[0, 0][torch.from_numpy(fl)] = torch.from_numpy(sol)
    d = (one_oracle_sweep(psi, rhs, mflags, drift=c) - psi)[0, 0][torch.from_numpy(fl)]
    assert (d - c).abs().max().item() < 1e-7 * np.abs(sol).max()           # every cell gains c (fp32-rounded weights: sum_q w_q = 1 - w0 only to 1e-8)


def test_the_oracle_sweeps_converge_to_that_solution():
    res = (10, 9)
    mflags, rhs, dim = poisson_case(res, seed=5)
    L, b, fl, c = dense_system(mflags, rhs)
    sol = np.linalg.lstsq(L, b, rcond=None)[0]
    # the reference's loop: propagation, rebounce_obstacle, one collision sweep; rhs enters through get_H_int's
    # formula, reproduced here with the oracle's pieces in fp64 (tau = 1)
    lat = O.D2Q9
    h = torch.zeros((1, 9, *res), dtype=torch.float64)
    w = O._wt(lat, h)
    S = rhs.double() * ((1.0 / 3.0) * (0.5 - 1.0))
    for _ in range(2500):
        h = O.rebounce_obstacle(O.propagation(h), mflags)
        psi = O.psi_of_h(h)
        heq = psi * w
        heq[:, 0:1] = heq[:, 0:1] - psi
        h = torch.where(mflags == 2, h, heq + w * S)
    p = psi[0, 0][torch.from_numpy(fl)].numpy()
    assert np.abs((p - p.mean()) - (sol - sol.mean())).max() < 1e-6 * np.abs(sol).max()
