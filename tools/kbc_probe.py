"""Development probe: first non-finite step of the fused D2Q9 KBC solver on the demo_2d_LBM_KBC initial condition."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbm_ferrofluid_b200 import initializers as I  # noqa: E402
from lbm_ferrofluid_b200.solver import SinglePhaseSolver  # noqa: E402
from lbm_ferrofluid_b200.types import KBCType  # noqa: E402

for n, steps in ((256, 60), (512, 30), (1024, 30), (4096, 30)):
    dev = torch.device("cuda")
    flags = torch.ones((1, 1, n, n), dtype=torch.uint8, device=dev)
    tau = 0.5 + 3 * 0.2 * n / 3000
    vel = I.double_shear_layer((n, n), device=dev)
    rho = torch.full((1, 1, n, n), 0.265, device=dev)
    s = SinglePhaseSolver(flags, tau=tau, KBC_type=int(KBCType.KBC_A)).initialize(rho, vel)
    bad = None
    for t in range(steps):
        s.step(1)
        if not torch.isfinite(s.f).all():
            bad = t + 1
            break
    nn = int((~torch.isfinite(s.f)).any(dim=1).sum())
    print(f"n={n} tau={tau:.4f} first non-finite step: {bad} (cells {nn}) max|u|={float(torch.nan_to_num(s.vel).abs().max()):.6f}",
          flush=True)
    if bad:
        idx = (~torch.isfinite(s.f)).any(dim=1).nonzero()[:5]
        print("  first bad cells (b,y,x):", idx.tolist())
