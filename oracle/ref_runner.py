#!/usr/bin/env python
"""Test / measurement infrastructure: run the UNMODIFIED reference solver in its own process.

The repo ships a ``src/LBM`` package of its own (the drop-in shim), so the reference's ``src.LBM`` can never be
imported into a process that has the repo root on sys.path.  This script is therefore always run as a *script*
(sys.path[0] = oracle/, repo root absent); it imports the reference from /root/reference when that exists (build
container) or from the byte-compiled copy oracle/_ref/pyref.zip (GPU box; oracle/build_ref_solver.py) and drives the
reference's own classes through the demo loop (make_golden.run_hcz).

    python oracle/ref_runner.py traj --dim 3 --res 130 66 130 --mag 100 --steps 10 50 --dtype float32 \
                                     --device cuda --out /tmp/ref32        # writes /tmp/ref32_<step>.npz
    python oracle/ref_runner.py time --dim 3 --res 96 96 96 --steps 3 --warmup 1 --device cpu   # one JSON line

Used by tests/test_gpu_reference.py (parity at demo size against the reference on CUDA in fp32 and fp64) and by
bench.py's reference arm / cpu_baseline (kind "reference").  Never imported by the product path.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
# the repo root must NOT be importable here (its src/ shim would shadow the reference)
sys.path[:] = [p for p in sys.path if os.path.abspath(p or os.getcwd()) != ROOT]
if HERE not in sys.path:
    sys.path.insert(0, HERE)

import numpy as np  # noqa: E402
import torch  # noqa: E402

import make_golden as M  # noqa: E402

KEYS = ("f", "g", "rho", "vel", "density", "pressure", "psi", "rho0", "flags", "mflags")


def case_kwargs(a):
    kw = dict(magnetic=not a.no_magnetic, mag=a.mag, gravity=a.gravity, kappa=a.kappa, tau_f=a.tau_f, tau_g=a.tau_g,
              contact_angle=a.contact_angle * math.pi, frac=a.frac, open_axes=tuple(a.open_axes), kbc=a.kbc,
              droplets=a.droplets, n_sweeps=a.n_sweeps, seed=a.seed, device=a.device, noise=a.noise)
    return kw


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd", choices=["traj", "time", "where"])
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--res", type=int, nargs="+", default=[130, 66, 130])
    ap.add_argument("--steps", type=int, nargs="+", default=[10])
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--dtype", default="float32", choices=["float32", "float64"])
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--out", default=None)
    ap.add_argument("--keys", nargs="+", default=list(KEYS))
    ap.add_argument("--mag", type=float, default=100.0)
    ap.add_argument("--gravity", type=float, default=1e-4)
    ap.add_argument("--kappa", type=float, default=0.01)
    ap.add_argument("--tau-f", type=float, default=0.68)
    ap.add_argument("--tau-g", type=float, default=0.68)
    ap.add_argument("--contact-angle", type=float, default=0.5, help="in units of pi")
    ap.add_argument("--frac", type=float, default=0.5)
    ap.add_argument("--open-axes", type=int, nargs="*", default=[1])
    ap.add_argument("--kbc", action="store_true")
    ap.add_argument("--droplets", action="store_true")
    ap.add_argument("--no-magnetic", action="store_true")
    ap.add_argument("--n-sweeps", type=int, default=30)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--noise", type=float, default=1e-3)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--flush-denormal", action="store_true",
                    help="TIMING only: psi decays exponentially away from the interface and the host cores process "
                         "the resulting denormals ~50x slower; flushing favours the CPU arm")
    a = ap.parse_args()
    M.import_reference()          # raises if `src` does not resolve inside the reference
    if a.cmd == "where":
        print(json.dumps({"reference": M.REF}))
        return
    dtype = getattr(torch, a.dtype)
    res = tuple(a.res)
    cores = a.threads or (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    if a.flush_denormal:
        torch.set_flush_denormal(True)
    cuda = a.device.startswith("cuda")
    if a.cmd == "traj":
        steps = sorted(set(a.steps))
        out = M.run_hcz(a.dim, res, steps[-1], dtype, snapshots=steps, **case_kwargs(a))
        for st, d in out.items():
            np.savez(f"{a.out}_{st}.npz", **{k: v for k, v in d.items() if k in a.keys or np.ndim(v) == 0 or k == "res"})
        print(json.dumps({"written": [f"{a.out}_{st}.npz" for st in out], "reference": M.REF}))
        return
    # ---- time: warm-up + timed outer steps, wall clock (device synchronised on CUDA)
    n, w = a.steps[0], a.warmup
    marks = {}

    def timer(step):
        if step == w + 1 or step == w + n + 1:
            if cuda:
                torch.cuda.synchronize()
            marks[step] = time.perf_counter()

    M.run_hcz(a.dim, res, w + n, dtype, timer=timer, **case_kwargs(a))
    dt = marks[w + n + 1] - marks[w + 1]
    cells = int(np.prod(res))
    print(json.dumps({"mlups": cells * n / dt / 1e6, "ms_per_step": dt / n * 1e3, "steps": n, "warmup": w,
                      "res": list(res), "device": a.device, "dtype": a.dtype, "cores": cores,
                      "flush_denormal": bool(a.flush_denormal), "reference": M.REF,
                      "torch": torch.__version__}))


if __name__ == "__main__":
    main()
