import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from lbm_ferrofluid_b200 import initializers as I
from lbm_ferrofluid_b200.solver import FerroSolver
dev = torch.device("cuda")
res = (12, 20, 36)
f32 = lambda v: torch.tensor(v, dtype=torch.float32, device=dev)
for kw in ({}, dict(density_liquid=0.31, density_gas=0.017, rho_liquid=0.29, rho_gas=0.02)):
    s = FerroSolver(I.shell_flags(res, device=dev), **kw)
    torch.manual_seed(1)
    s.rho.copy_(0.3 * torch.rand((1, 1, *res), device=dev) - 0.02)
    rg, rl, dg, dl = s._rho_gas, s._rho_liquid, s._density_gas, s._density_liquid
    x = s.rho
    print(kw or "defaults")
    t1 = x - rg
    print("  sub", int(((x - f32(rg)) != t1).sum()))
    t2 = t1 / (rl - rg)
    print("  div: 1f/f32(b)", int((t1 * (f32(1.0) / f32(rl - rg)) != t2).sum()), " f32(1/b)", int((t1 * f32(1.0 / (rl - rg)) != t2).sum()),
          " true", int((t1 / f32(rl - rg) != t2).sum()))
    t3 = (dl - dg) * t2
    print("  mul", int((t2 * f32(dl - dg) != t3).sum()))
    t4 = dg + t3
    print("  add", int((t3 + f32(dg) != t4).sum()))
    t5 = t4 - dg
    t6 = 2.0 * t5
    t7 = t6 / (dl - dg)
    print("  div2: 1f/f32(b)", int((t6 * (f32(1.0) / f32(dl - dg)) != t7).sum()), " f32(1/b)", int((t6 * f32(1.0 / (dl - dg)) != t7).sum()))
    print("  kernel vs torch", int((s.phi != s.phi_torch).sum()))
