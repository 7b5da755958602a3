"""The reference's UNMODIFIED demo scripts against this repo (drop-in boundary).

Runs only where the reference checkout exists (the build container).  Each script is copied to a temp directory
(the reference tree is read-only and the scripts write next to themselves), started with this repo ahead of it on
PYTHONPATH, and must get through imports, argparse, SimulationParameters / SimulationRunner factories and its own
initial-condition code.  On a machine without a GPU the scripts select device = cpu, so the run must stop exactly at
the first solver op with our "no CPU fallback" RuntimeError -- any earlier failure is a hole in the boundary."""
import glob
import os
import shutil
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DEMOS = os.environ.get("LBMFF_REFERENCE", "/root/reference") + "/demo"
DEMOS = sorted(os.path.basename(p) for p in glob.glob(REF_DEMOS + "/demo_*.py"))

pytestmark = pytest.mark.skipif(not DEMOS or torch.cuda.is_available(),
                                reason="needs the reference checkout and a machine without a GPU")

# scripts whose initial condition is an O(N) python loop over a 3-D grid get a smaller grid on the command line
ARGS = {"demo_3d_LBM_two_droplets.py": ["--res", "12", "12", "24"],
        "demo_3d_LBM_droplet_spread.py": ["--res", "16", "16", "16"],
        "demo_3d_LBM_multiphase.py": ["--res", "16", "16", "16"],
        "demo_3d_LBM_Rosensweig_instability.py": ["--res", "16", "12", "16"]}


@pytest.mark.parametrize("name", DEMOS)
def test_unmodified_demo_reaches_the_first_solver_op(name, tmp_path):
    demo_dir = tmp_path / "demo"
    demo_dir.mkdir()
    shutil.copy(os.path.join(REF_DEMOS, name), demo_dir / name)
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([ROOT, os.path.join(ROOT, "compat")]), CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, name, "--total_steps", "2", *ARGS.get(name, [])], cwd=demo_dir, env=env,
                       capture_output=True, text=True, timeout=600)
    tail = (r.stderr or "").strip().splitlines()[-1] if r.stderr.strip() else ""
    assert r.returncode != 0 and "no CPU fallback" in tail, f"{name} stopped elsewhere:\n{r.stderr[-2000:]}"
    assert "lbm_ferrofluid_b200" in r.stderr      # the op that raised is ours, reached through src.LBM.*
