#!/usr/bin/env python
"""The loop of demo/demo_3d_LBM_Rosensweig_instability.py on the fused API of this repo: FerroSolver.step for the
solver, save_rendered_image (device marching cubes + tiled soft rasteriser) for the frames, export_asset for the
final mesh.  Same parameters and frame cadence as the demo (:26-60, :102-113, :126-132, :208-232).

    python examples/rosensweig_3d.py --res 130 130 130 --total_steps 350 --mag_strength 100 --out /tmp/ros3d
"""
import argparse
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbm_ferrofluid_b200 import initializers as I  # noqa: E402
from lbm_ferrofluid_b200.io import export_asset, mkdir, save_rendered_image  # noqa: E402
from lbm_ferrofluid_b200.render import SoftRenderer  # noqa: E402
from lbm_ferrofluid_b200.solver import FerroSolver  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, nargs=3, default=[130, 130, 130])
    ap.add_argument("--total_steps", type=int, default=350)
    ap.add_argument("--mag_strength", type=float, default=100.0)
    ap.add_argument("--gravity_strength", type=float, default=1e-4)
    ap.add_argument("--frame_every", type=int, default=10)
    ap.add_argument("--magnetic_path", default="auto", choices=["auto", "faithful", "scalar", "converged"])
    ap.add_argument("--perturb", type=float, default=0.0, help="cells of sinusoidal interface perturbation (demo: 0)")
    ap.add_argument("--graph", action="store_true", help="replay a CUDA graph of two steps instead of ~45 launches per step")
    ap.add_argument("--out", default="ros3d_out")
    a = ap.parse_args()
    dev = torch.device("cuda")
    res = tuple(a.res)
    flags = I.shell_flags(res, device=dev)                       # wall = xXyYzZ            (demo :114-117)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)      # magnetic wall = xX zZ    (demo :118-120)
    rho = I.rosensweig_rho(res, flags, perturb=a.perturb)        # liquid below y = H/2     (demo :128-132)
    solver = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, a.mag_strength, 0.0], kappa=0.01, tau_f=0.68,
                         tau_g=0.68, contact_angle=torch.tensor([0.5 * math.pi]), gravity=a.gravity_strength,
                         n_sweeps=30, magnetic_path=a.magnetic_path)
    solver.initialize(rho)
    renderer = SoftRenderer(camera_mode="look_at", eye=torch.tensor([1.6, 0.8, 1.6], device=dev),
                            camera_direction=torch.tensor([-1 / math.sqrt(2), 0.0, -1 / math.sqrt(2)], device=dev),
                            near=1.0, bg_color=torch.tensor([1.0, 1.0, 1.0], device=dev), device=dev)
    mkdir(a.out)
    t0 = time.time()
    frames = 0
    done = 0
    while done < a.total_steps:
        # the demo writes a frame after step index 0, frame_every, 2 * frame_every, ...: run up to the next one
        due = ((done + a.frame_every - 1) // a.frame_every) * a.frame_every
        n = min(due - done + 1, a.total_steps - done)
        solver.step(n, graph=a.graph)
        done += n
        if (done - 1) % a.frame_every == 0:
            save_rendered_image(renderer, solver.phi, os.path.join(a.out, f"{done:03}.png"), res, 1.0)
            frames += 1
    verts, faces = save_rendered_image(renderer, solver.phi, os.path.join(a.out, "final_frame.png"), res, 1.0)
    torch.cuda.synchronize()
    dt = time.time() - t0
    export_asset(os.path.join(a.out, "Rosensweig_instability.obj"), verts, faces)
    cells = res[0] * res[1] * res[2]
    h = (solver.phi[0, 0, 1:-1, 1:-1, 1:-1] < 0).float().sum(dim=1)             # liquid column heights
    print(f"{a.total_steps} steps + {frames + 1} frames at {res}: {dt:.2f} s wall "
          f"({cells * a.total_steps / dt / 1e6:.0f} MLUPS including frames), path={solver.magnetic_path}, "
          f"mass={solver.total_mass():.6f}, finite={bool(torch.isfinite(solver.rho).all())}, "
          f"faces={faces.shape[0]}, column height min/mean/max = {h.min().item():.0f}/{h.mean().item():.1f}/"
          f"{h.max().item():.0f}")


if __name__ == "__main__":
    main()
