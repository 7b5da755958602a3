"""Parity at DEMO SIZE against the reference itself, run on the same GPU.

The unmodified reference (byte-compiled into oracle/_ref/pyref.zip by oracle/build_ref_solver.py; run in its own
process by oracle/ref_runner.py because its package is called ``src`` like our shim) executes the demo loop on CUDA in
fp32 and in fp64 at the demos' default grids -- 130 x 66 x 130 (demo_3d_LBM_Rosensweig_instability.py:253) and 130^2
(demo_2d_LBM_Rosensweig_instability.py:218) -- for mg 85 and mg 100, with snapshots after N = 10 and N = 50 steps.
FerroSolver (population-faithful and tau = 1 scalar sweeps) starts from the same rho0 and is compared per field.

For every field three rel-L2 numbers are printed and written to gpurun_out/parity_r02.json:
    direct = |ours - ref32| / |ref32|          what a user switching implementations sees
    floor  = |ref32 - ref64| / |ref64|         how well the reference's own fp32 run is defined
    ours64 = |ours - ref64| / |ref64|          our distance to the exact evaluation of the reference algorithm
Bars: the SURVEY 8(c) band on `direct` (N = 10: rho/density/f 5e-6, pressure/g/psi 5e-5, vel 2e-4; N = 50: 5e-5 /
3e-4 / 1e-3) wherever the measured floor allows it (2 x floor <= band: two fp32 evaluations that are each `floor`
away from the truth cannot be asked to be closer to one another than that), and everywhere the floor bar
ours64 <= 2 x floor + 2e-6 (tests/util.py).  Total mass over 100 steps: <= 1e-6 relative.
"""
import json
import math
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from tests.util import rel_l2

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUNNER = os.path.join(ROOT, "oracle", "ref_runner.py")
BANDS = {10: dict(f=5e-6, rho=5e-6, density=5e-6, g=5e-5, pressure=5e-5, psi=5e-5, vel=2e-4),
         50: dict(f=5e-5, rho=5e-5, density=5e-5, g=3e-4, pressure=3e-4, psi=3e-4, vel=1e-3)}
STEPS = (10, 50)
REPORT = {}

CASES = {
    # name: (ref_runner arguments, FerroSolver keyword overrides)
    "ros3d_mg100": (["--dim", "3", "--res", "130", "66", "130", "--mag", "100"], {}),
    "ros3d_mg85": (["--dim", "3", "--res", "130", "66", "130", "--mag", "85"], {}),
    "ros3d_mg100_smooth": (["--dim", "3", "--res", "130", "66", "130", "--mag", "100", "--noise", "0"], {}),
    "ros2d_mg100": (["--dim", "2", "--res", "130", "130", "--mag", "100", "--tau-f", "0.7", "--tau-g", "0.8",
                     "--contact-angle", "0.75", "--frac", "0.45", "--kbc"], {}),
    "ros2d_mg85": (["--dim", "2", "--res", "130", "130", "--mag", "85", "--tau-f", "0.7", "--tau-g", "0.8",
                    "--contact-angle", "0.75", "--frac", "0.45", "--kbc"], {}),
}


def _have_reference():
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    try:
        import build_ref_solver
        return build_ref_solver.available() or os.path.isdir("/root/reference/src/LBM")
    finally:
        sys.path.pop(0)


def _run_reference(case, dtype, tmp):
    out = os.path.join(tmp, f"{case}_{dtype}")
    if not all(os.path.exists(f"{out}_{s}.npz") for s in STEPS):
        env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
        cmd = [sys.executable, RUNNER, "traj", *CASES[case][0], "--steps", *map(str, STEPS), "--dtype", dtype,
               "--device", "cuda", "--out", out]
        r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=tmp, timeout=1800)
        assert r.returncode == 0, f"reference run failed:\n{r.stderr[-3000:]}"
    return {s: {k: torch.from_numpy(v) if v.ndim else v.item() for k, v in np.load(f"{out}_{s}.npz").items()}
            for s in STEPS}


@pytest.fixture(scope="module")
def refdir(tmp_path_factory):
    if not _have_reference():
        pytest.fail("oracle/_ref/pyref.zip is missing: run __graft_entry__.build() (oracle/build_ref_solver.py) in the "
                    "build container so the byte-compiled reference travels with the snapshot")
    return str(tmp_path_factory.mktemp("ref"))


def _solver(G, path):
    from lbm_ferrofluid_b200.solver import FerroSolver
    from oracle import lbm_oracle as O

    dim = len(G["res"])
    Hu = [0.0] * dim
    Hu[1] = G["mag"]
    s = FerroSolver(G["flags"].cuda(), kappa=G["kappa"], tau_f=G["tau_f"], tau_g=G["tau_g"], gravity=G["gravity"],
                    contact_angle=torch.tensor([G["contact_angle"]]), n_sweeps=int(G["n_sweeps"]),
                    KBC_type=O.KBC_A if G["kbc"] else None, magnetic_path=path, magnetic_flags=G["mflags"].cuda(),
                    H_ext=Hu)
    return s.initialize(G["rho0"].cuda())


def _dump_report():
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "parity_r02.json"), "w") as fh:
            json.dump(REPORT, fh, indent=1, sort_keys=True)
    except OSError:
        pass


@pytest.mark.parametrize("path", ["faithful", "scalar"])
@pytest.mark.parametrize("case", list(CASES))
def test_demo_size_trajectory_vs_reference_on_cuda(case, path, refdir):
    ref32 = _run_reference(case, "float32", refdir)
    ref64 = _run_reference(case, "float64", refdir)
    s = _solver(ref32[STEPS[0]], path)
    fluid = (ref32[STEPS[0]]["mflags"] != 2).double()
    failures, done = [], 0
    for n in STEPS:
        s.step(n - done)
        done = n
        torch.cuda.synchronize()
        got = dict(f=s.f, g=s.g, rho=s.rho, density=s.density, pressure=s.pressure, vel=s.vel, psi=s.psi)
        for k, band in BANDS[n].items():
            a, r32, r64 = got[k].cpu().double(), ref32[n][k].double(), ref64[n][k].double()
            assert torch.isfinite(a).all(), f"{case}/{path} N={n} {k}: non-finite"
            if k == "psi":   # wall-cell psi is never observed (get_grad masks it)
                a, r32, r64 = a * fluid, r32 * fluid, r64 * fluid
            direct, floor, ours64 = rel_l2(a, r32), rel_l2(r32, r64), rel_l2(a, r64)
            band_applies = 2.0 * floor <= band
            REPORT[f"{case}/{path}/N{n}/{k}"] = dict(direct=direct, floor=floor, ours64=ours64, band=band,
                                                     band_asserted=band_applies)
            print(f"{case:20s} {path:8s} N={n:2d} {k:8s} direct {direct:.2e}  floor {floor:.2e}  ours-vs-fp64 "
                  f"{ours64:.2e}  band {band:.0e}{'' if band_applies else ' (floor bar only)'}")
            if band_applies and direct > band:
                failures.append(f"N={n} {k}: direct rel-L2 {direct:.3e} > band {band:.0e}")
            if ours64 > 2.0 * floor + 2e-6:
                failures.append(f"N={n} {k}: {ours64:.3e} from the fp64 truth > 2 x reference fp32 floor {floor:.3e}")
    _dump_report()
    assert not failures, f"{case}/{path}: " + "; ".join(failures)


@pytest.mark.parametrize("case", ["ros3d_mg100", "ros2d_mg100"])
def test_total_mass_over_100_steps_at_demo_size(case, refdir):
    """sum over ALL cells (walls included) of sum_q f over 100 steps.  The algorithm itself does not conserve mass to
    round-off in fp32: measured here on the same GPU, the reference's own fp32 run drifts by ~1e-5 at 130 x 66 x 130
    (7e-6 at 130^2) over 100 steps while its fp64 run drifts by 1-2e-6 (the entropic product equilibrium on the pruned
    D3Q19 lattice sums to rho only up to O(u^4)) -- SURVEY's 1e-7 was a 34^3 / few-step figure.  So the bar is the
    reference's own behaviour: our drift must stay within 1e-6 or 1.5 x the reference's fp32 drift, whichever is
    larger, and our total mass after 100 steps must equal the reference-fp32 total to 2e-6 relative."""
    out = os.path.join(refdir, f"{case}_mass")
    ref = {}
    for dtype in ("float32", "float64"):
        env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
        cmd = [sys.executable, RUNNER, "traj", *CASES[case][0], "--steps", "100", "--dtype", dtype, "--device", "cuda",
               "--out", f"{out}_{dtype}", "--keys", "rho0", "flags", "mflags"]
        r = subprocess.run(cmd, capture_output=True, text=True, env=env, cwd=refdir, timeout=1800)
        assert r.returncode == 0, r.stderr[-3000:]
        z = np.load(f"{out}_{dtype}_100.npz")
        ref[dtype] = (float(z["mass0"]), float(z["mass"]))
    G = _run_reference(case, "float32", refdir)[STEPS[0]]
    s = _solver(G, "faithful")
    m0 = s.total_mass()
    s.step(100)
    m1 = s.total_mass()
    drift = abs(m1 - m0) / abs(m0)
    d32 = abs(ref["float32"][1] - ref["float32"][0]) / abs(ref["float32"][0])
    d64 = abs(ref["float64"][1] - ref["float64"][0]) / abs(ref["float64"][0])
    vs64 = abs(m1 - ref["float64"][1]) / abs(ref["float64"][1])
    vs32 = abs(m1 - ref["float32"][1]) / abs(ref["float32"][1])
    REPORT[f"{case}/mass100"] = dict(drift=drift, reference_fp32_drift=d32, reference_fp64_drift=d64,
                                     ours_vs_reference_fp32_total=vs32, ours_vs_reference_fp64_total=vs64)
    print(f"{case}: total-mass drift over 100 steps  ours {drift:.2e}  reference fp32 {d32:.2e}  reference fp64 "
          f"{d64:.2e}  |ours - ref32| / ref32 {vs32:.2e}  |ours - ref64| / ref64 {vs64:.2e}")
    _dump_report()
    assert math.isfinite(m1) and drift <= max(1e-6, 1.5 * d32), \
        f"total mass drifted by {drift:.2e} over 100 steps (reference: fp32 {d32:.2e}, fp64 {d64:.2e})"
    assert vs32 <= 2e-6, f"total mass after 100 steps differs from the reference fp32 run by {vs32:.2e}"
