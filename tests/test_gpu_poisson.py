"""Converged magnetic potential: lbmff_poisson_cg_* / PoissonCG / FerroSolver("converged").

The reference has no converged solver, so there is no golden: the checker is the reference's OWN sweep.  A solution
psi* is correct iff it is a fixed point (modulo the additive constant the sweeps themselves drift by) of one
propagation -> rebounce_obstacle -> psi_of_h sweep, which the CPU oracle executes in fp64
(oracle/lbm_oracle.py, pinned against the reference's goldens), and iff the parity-tested GPU sweeps run long enough
arrive at the same potential."""
import math

import pytest
import torch

from tests.util import one_oracle_sweep as _one_oracle_sweep, poisson_b_and_drift, poisson_case as _case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("res", [(40, 36), (18, 20, 22)])
def test_cg_solution_is_a_fixed_point_of_the_reference_sweep(res):
    from lbm_ferrofluid_b200.poisson import PoissonCG

    mflags, rhs, dim = _case(res)
    cg = PoissonCG(mflags.cuda(), tol=1e-6, max_iters=1500, keep_history=True)
    psi = torch.zeros((1, 1, *res), device="cuda")
    cg.solve(rhs.cuda(), psi)
    info = cg.info()[0]
    assert info["converged"], info
    assert info["iters"] < 400, info                  # Jacobi-PCG on n ~ 40: O(n) iterations, not O(n^2) sweeps
    hist = cg.history()[0]
    assert len(hist) == info["iters"] + 1 and hist[-1] <= 1e-6 * info["rhs_norm"] * 1.001
    fluid = (mflags != 2)
    _, _, c = poisson_b_and_drift(rhs, mflags)        # fp64, by rolls: what every sweep adds in the stationary state
    assert abs(info["drift"] - c.item()) <= 1e-5 * abs(c.item()) + 1e-12
    after = _one_oracle_sweep(psi.cpu(), rhs, mflags, drift=c.item())
    d = (after - psi.cpu().double())[fluid]
    scale = (psi.cpu().double()[fluid] - psi.cpu().double()[fluid].mean()).abs().max()
    assert scale > 1e-3
    assert (d - c.item()).abs().max().item() <= 2e-5 * scale.item(), ((d - c.item()).abs().max().item(), scale.item())


@pytest.mark.parametrize("res", [(32, 28), (16, 14, 18)])
def test_cg_equals_the_sweeps_run_to_convergence(res):
    from lbm_ferrofluid_b200 import _host as H
    from lbm_ferrofluid_b200.poisson import PoissonCG

    mflags, rhs, dim = _case(res, seed=1)
    mf, r = mflags.cuda(), rhs.cuda()
    Q = 9 if dim == 2 else 19
    hA = torch.zeros((1, Q, *res), device="cuda")
    hB = torch.zeros_like(hA)
    psi_sw = torch.zeros((1, 1, *res), device="cuda")
    H.launch("lbmff_magnetic_sweeps_" + H.suffix(dim), mf.device, H.ptr(hA), H.ptr(hB), H.ptr(mf), H.ptr(r),
             H.ptr(psi_sw), 6000, 1.0, 1.0, 1.0, *H.sizes(dim, 1, res[0] if dim == 3 else 1, res[-2], res[-1]))
    psi = torch.zeros_like(psi_sw)
    PoissonCG(mf, tol=1e-6, max_iters=1000).solve(r, psi)
    fluid = (mf != 2)
    a = psi_sw[fluid].double()
    b = psi[fluid].double()
    a, b = a - a.mean(), b - b.mean()
    assert (a - b).norm().item() <= 1e-4 * a.norm().item(), (a - b).norm().item() / a.norm().item()


def test_cg_is_deterministic_warm_start_is_free_and_batches_are_independent():
    from lbm_ferrofluid_b200.poisson import PoissonCG

    res = (20, 18, 16)
    m0, r0, _ = _case(res, seed=2)
    m1, r1, _ = _case(res, seed=3, interior_wall=False)
    mf, r = torch.cat([m0, m1]).cuda(), torch.cat([r0, r1]).cuda()
    cg = PoissonCG(mf, tol=1e-5, max_iters=300)
    p1, p2 = torch.zeros_like(r), torch.zeros_like(r)
    cg.solve(r, p1)
    i1 = cg.info()
    cg.solve(r, p2)
    assert torch.equal(p1, p2) and all(i["converged"] for i in i1)
    single = PoissonCG(m1.cuda(), tol=1e-5, max_iters=300)
    q = torch.zeros((1, 1, *res), device="cuda")
    single.solve(r1.cuda(), q)
    assert torch.equal(q[0], p1[1])
    before = p1.clone()
    cg.solve(r, p1)                                   # warm start from the solution: nothing (or one polish step) to do
    assert all(i["iters"] <= 1 and i["converged"] for i in cg.info())
    assert (before - p1).abs().max().item() <= 1e-5 * before.abs().max().item()


def test_graph_while_loop_equals_one_by_one_launches(monkeypatch):
    """The device-side WHILE loop (CUDA graph conditional node) and max_iters plain launches give the same bits,
    the same iteration count and the same residual history."""
    from lbm_ferrofluid_b200.poisson import PoissonCG

    res = (18, 16, 20)
    m, r, _ = _case(res, seed=6)
    out = []
    for no_graph in ("0", "1"):
        monkeypatch.setenv("LBMFF_CG_NO_GRAPH", no_graph)
        cg = PoissonCG(m.cuda(), tol=1e-5, max_iters=150, keep_history=True)
        psi = torch.zeros((1, 1, *res), device="cuda")
        cg.solve(r.cuda(), psi)
        out.append((psi.clone(), cg.info()[0], cg.history()[0]))
    assert out[0][1]["converged"] and 5 < out[0][1]["iters"] < 150
    assert torch.equal(out[0][0], out[1][0]) and out[0][1] == out[1][1] and out[0][2] == out[1][2]
    # the iteration cap also ends the loop
    monkeypatch.setenv("LBMFF_CG_NO_GRAPH", "0")
    cg = PoissonCG(m.cuda(), tol=0.0, max_iters=7)
    cg.solve(r.cuda(), torch.zeros((1, 1, *res), device="cuda"))
    assert cg.info()[0]["iters"] == 7 and not cg.info()[0]["converged"]


def test_cg_argument_errors():
    from lbm_ferrofluid_b200 import _host as H
    from lbm_ferrofluid_b200.poisson import PoissonCG
    from lbm_ferrofluid_b200.solver import FerroSolver

    m, r, _ = _case((12, 12, 12))
    cg = PoissonCG(m.cuda(), max_iters=5)
    psi = torch.zeros_like(r).cuda()
    with pytest.raises(RuntimeError, match="tau == 1"):
        H.launch("lbmff_poisson_cg_3d", psi.device, H.ptr(r.cuda()), H.ptr(cg.mflags), H.ptr(psi), H.ptr(cg._work),
                 H.ptr(cg._scratch), 5, 1e-4, 0.8, 1.0, 1.0, H.ptr(None), H.ptr(None), 1, 12, 12, 12)
    with pytest.raises(RuntimeError, match="null pointer"):
        H.launch("lbmff_poisson_cg_3d", psi.device, H.ptr(None), H.ptr(cg.mflags), H.ptr(psi), H.ptr(cg._work),
                 H.ptr(cg._scratch), 5, 1e-4, 1.0, 1.0, 1.0, H.ptr(None), H.ptr(None), 1, 12, 12, 12)
    with pytest.raises(RuntimeError, match="sized for"):
        cg.solve(r.cuda(), psi, max_iters=6)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        cg.solve(r, psi)
    with pytest.raises(RuntimeError, match="tau == 1"):
        FerroSolver(m.cuda(), magnetic_flags=m.cuda(), H_ext=[0.0, 1.0, 0.0], tau=0.9, magnetic_path="converged")


def test_converged_path_matches_many_sweeps_in_the_full_step():
    """FerroSolver("converged") against the reference algorithm with its sweep count raised until the potential is
    converged every step (n_sweeps = 1500 at 20 x 24 x 20): same physics, so the trajectories agree to the
    tolerance of the two iterative solves -- while n_sweeps = 30 (the demo's value) visibly has not converged."""
    from lbm_ferrofluid_b200.solver import FerroSolver
    from oracle.make_golden import perturbed_interface, shell_flags

    res, steps = (20, 24, 20), 4
    flags = shell_flags(res)
    mflags = shell_flags(res, open_axes=(1,))
    rho0 = perturbed_interface(res, 0.5, 0.2508, 0.02381, torch.float32, False)

    def run(path, n):
        s = FerroSolver(flags.cuda(), magnetic_flags=mflags.cuda(), H_ext=[0.0, 60.0, 0.0], n_sweeps=n,
                        magnetic_path=path, cg_tol=1e-6, cg_max_iters=300)
        s.initialize(rho0.cuda()).step(steps)
        return s

    ref, cg, demo = run("scalar", 1500), run("converged", 30), run("scalar", 30)
    assert all(i["converged"] for i in cg.cg.info())
    fluid = (mflags != 2).cuda()

    def centred(s):
        p = s.psi[fluid].double()
        return p - p.mean()

    e_cg = ((centred(cg) - centred(ref)).norm() / centred(ref).norm()).item()
    e_demo = ((centred(demo) - centred(ref)).norm() / centred(ref).norm()).item()
    assert e_cg <= 2e-4, e_cg
    assert e_demo > 20 * e_cg, (e_demo, e_cg)
    for k in ("rho", "pressure", "vel"):
        a, b = getattr(cg, k).double(), getattr(ref, k).double()
        assert torch.isfinite(a).all()
        assert ((a - b).norm() / b.norm()).item() <= (5e-3 if k == "vel" else 1e-4), k
    assert abs(cg.total_mass() - ref.total_mass()) <= 1e-5 * abs(ref.total_mass())
    assert cg.launches_per_step == 1 + (1 + 5) + 5 + 1 + 6 + 1
    assert math.isfinite(cg.total_mass())


@pytest.mark.parametrize("res", [(40, 36), (90, 70), (18, 20, 22), (40, 36, 44), (33, 47, 29)])
@pytest.mark.parametrize("graph", [True, False])
def test_multigrid_pcg_solves_the_same_system_in_far_fewer_iterations(res, graph, monkeypatch):
    """precondition="mg": one geometric-multigrid V(2,2) cycle per CG iteration (Galerkin coarse operators on cell
    aggregates, walls as zero-weight links; csrc/poisson_mg.cuh).  Same system, so the solution equals the plain CG
    one (up to the constant) and is a fixed point of the reference's sweep; iterations must drop by a large factor and
    stay nearly independent of the grid (odd / ragged extents and an interior wall included)."""
    from lbm_ferrofluid_b200.poisson import PoissonCG

    monkeypatch.setenv("LBMFF_CG_NO_GRAPH", "0" if graph else "1")
    mflags, rhs, dim = _case(res, seed=2)
    mf, r = mflags.cuda(), rhs.cuda()
    plain = PoissonCG(mf, tol=1e-6, max_iters=2000)
    psi0 = torch.zeros((1, 1, *res), device="cuda")
    plain.solve(r, psi0)
    pcg = PoissonCG(mf, tol=1e-6, max_iters=100, keep_history=True, precondition="mg")
    psi = torch.zeros_like(psi0)
    pcg.solve(r, psi)
    i0, i1 = plain.info()[0], pcg.info()[0]
    assert i1["converged"], i1
    assert i1["iters"] <= 20 and i1["iters"] * 4 <= i0["iters"], (i0["iters"], i1["iters"])
    hist = pcg.history()[0]
    assert hist[-1] <= 1e-6 * i1["rhs_norm"] * 1.001 and all(h == h for h in hist)
    fluid = (mf != 2)
    a, b = psi0[fluid].double(), psi[fluid].double()
    a, b = a - a.mean(), b - b.mean()
    assert (a - b).norm().item() <= 2e-5 * a.norm().item(), (a - b).norm().item() / a.norm().item()
    assert torch.equal(psi[~fluid], torch.zeros_like(psi[~fluid]))      # wall cells untouched
    # deterministic
    psi2 = torch.zeros_like(psi0)
    pcg.solve(r, psi2)
    assert torch.equal(psi, psi2)


def test_multigrid_pcg_warm_start_and_batch():
    from lbm_ferrofluid_b200.poisson import PoissonCG

    res = (24, 20, 28)
    m0, r0, _ = _case(res, seed=3)
    m1, r1, _ = _case(res, seed=4, interior_wall=False)
    mf, r = torch.cat([m0, m1]).cuda(), torch.cat([r0, r1]).cuda()
    pcg = PoissonCG(mf, tol=1e-5, max_iters=60, precondition="mg")
    psi = torch.zeros((2, 1, *res), device="cuda")
    pcg.solve(r, psi)
    first = [i["iters"] for i in pcg.info()]
    single = PoissonCG(m1.cuda(), tol=1e-5, max_iters=60, precondition="mg")
    ps = torch.zeros((1, 1, *res), device="cuda")
    single.solve(r1.cuda(), ps)
    assert torch.equal(ps[0], psi[1])                 # batch items are independent
    pcg.solve(r, psi)                                 # warm start: already converged
    assert all(i["iters"] <= 1 for i in pcg.info()) and all(f >= 3 for f in first)
    with pytest.raises(ValueError):
        PoissonCG(mf, precondition="ilu")


@pytest.mark.parametrize("rows", ["2", "4"])
@pytest.mark.parametrize("res", [(18, 20, 22), (33, 47, 29), (70, 24, 72), (64, 40, 96)])
def test_tiled_multigrid_smoother_equals_the_generic_one(res, rows, monkeypatch):
    """k_mg0_sweep_tiled (32 x 8 column marching along z, plane sums out of a shared-memory ring) against the generic
    level-0 smoother k_mg0_sweep (LBMFF_MG_TILED=0): the same preconditioner up to fp32 summation order, so the same
    solution, the same iteration count (+-1) and matching residual histories -- ragged x / y / z extents, z chunks that
    do not divide D, and an interior wall included."""
    from lbm_ferrofluid_b200.poisson import PoissonCG

    mflags, rhs, dim = _case(res, seed=5)
    mf, r = mflags.cuda(), rhs.cuda()
    out = {}
    monkeypatch.setenv("LBMFF_MG_ROWS", rows)          # rows of the tile per thread: 2 (default) or 4
    for tiled in ("1", "0"):
        monkeypatch.setenv("LBMFF_MG_TILED", tiled)
        pcg = PoissonCG(mf, tol=1e-6, max_iters=60, keep_history=True, precondition="mg")
        psi = torch.zeros((1, 1, *res), device="cuda")
        pcg.solve(r, psi)
        out[tiled] = (psi, pcg.info()[0], pcg.history()[0])
    (pa, ia, ha), (pb, ib, hb) = out["1"], out["0"]
    assert ia["converged"] and ib["converged"], (ia, ib)
    assert abs(ia["iters"] - ib["iters"]) <= 1 and ia["iters"] <= 20, (ia["iters"], ib["iters"])
    for k in range(1, min(4, len(ha), len(hb))):      # early history: the two preconditioners are the same operator
        assert abs(ha[k] - hb[k]) <= 1e-3 * hb[0], (k, ha[k], hb[k])
    fluid = (mf != 2)
    a, b = pa[fluid].double(), pb[fluid].double()
    a, b = a - a.mean(), b - b.mean()
    assert (a - b).norm().item() <= 2e-5 * b.norm().item(), (a - b).norm().item() / b.norm().item()
    assert torch.equal(pa[~fluid], torch.zeros_like(pa[~fluid]))


def test_ferrosolver_with_multigrid_preconditioner_follows_the_plain_cg_trajectory():
    """FerroSolver(magnetic_path="converged", cg_precondition="mg"): the same converged step with the V-cycle inside
    the solve (hierarchy built by the first solve, reused afterwards) -- psi and the flow fields follow the plain-CG
    solver to the tolerance of the two iterative solves, in a third of the iterations or fewer on the cold steps."""
    from lbm_ferrofluid_b200.solver import FerroSolver
    from oracle.make_golden import perturbed_interface, shell_flags

    res, steps = (24, 28, 24), 4
    flags = shell_flags(res)
    mflags = shell_flags(res, open_axes=(1,))
    rho0 = perturbed_interface(res, 0.5, 0.2508, 0.02381, torch.float32, False)
    out = {}
    for pre in (None, "mg"):
        s = FerroSolver(flags.cuda(), magnetic_flags=mflags.cuda(), H_ext=[0.0, 60.0, 0.0], magnetic_path="converged",
                        cg_tol=1e-6, cg_max_iters=300, cg_precondition=pre)
        s.initialize(rho0.cuda())
        iters = []
        for _ in range(steps):
            s.step()
            i = s.cg.info()[0]
            assert i["converged"], (pre, i)
            iters.append(i["iters"])
        out[pre] = (s, iters)
    (a, ia), (b, ib) = out[None], out["mg"]
    assert b.cg._hierarchy_built
    assert ib[0] * 3 <= ia[0], (ia, ib)
    fluid = (mflags != 2).cuda()
    pa, pb = a.psi[fluid].double(), b.psi[fluid].double()
    pa, pb = pa - pa.mean(), pb - pb.mean()
    assert ((pa - pb).norm() / pa.norm()).item() <= 2e-5
    for k in ("rho", "pressure", "vel"):
        x, y = getattr(a, k).double(), getattr(b, k).double()
        assert ((x - y).norm() / x.norm()).item() <= (1e-4 if k == "vel" else 1e-6), k


@pytest.mark.parametrize("res", [(18, 20, 22), (33, 47, 29), (70, 24, 72)])
def test_tiled_cg_operator_equals_the_generic_one(res, monkeypatch):
    """k_cg_apply_tiled (3-D, plane ring in shared memory, laplace()'s differences in the same order) against
    k_cg_apply (LBMFF_CG_TILED=0): Ap is the same bit for bit, only the block partition of the p.Ap partial sums
    differs, so the two solves follow each other to rounding -- same iteration count (+-1), same solution -- on
    ragged extents with an interior wall."""
    from lbm_ferrofluid_b200.poisson import PoissonCG

    mflags, rhs, dim = _case(res, seed=7)
    mf, r = mflags.cuda(), rhs.cuda()
    out = {}
    for tiled in ("1", "0"):
        monkeypatch.setenv("LBMFF_CG_TILED", tiled)
        cg = PoissonCG(mf, tol=1e-6, max_iters=1500, keep_history=True)
        psi = torch.zeros((1, 1, *res), device="cuda")
        cg.solve(r, psi)
        out[tiled] = (psi, cg.info()[0], cg.history()[0])
    (pa, ia, ha), (pb, ib, hb) = out["1"], out["0"]
    assert ia["converged"] and ib["converged"], (ia, ib)
    assert abs(ia["iters"] - ib["iters"]) <= 1, (ia["iters"], ib["iters"])
    assert abs(ha[1] - hb[1]) <= 1e-5 * hb[0] and abs(ha[3] - hb[3]) <= 1e-4 * hb[0]
    fluid = (mf != 2)
    a, b = pa[fluid].double(), pb[fluid].double()
    a, b = a - a.mean(), b - b.mean()
    assert (a - b).norm().item() <= 2e-5 * b.norm().item()
    assert torch.equal(pa[~fluid], torch.zeros_like(pa[~fluid]))
