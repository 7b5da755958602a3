#!/usr/bin/env python
"""Kernel micro-benchmarks through the C ABI (development tool, not part of the product path).

    python tools/kbench.py sweep D H W [n_sweeps]     -> ms per population-faithful sweep, GB/s at 152 B/cell
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbm_ferrofluid_b200 import _host as H  # noqa: E402
from lbm_ferrofluid_b200 import initializers as I  # noqa: E402


def sweep(res, n=30, reps=3):
    D, Hh, W = res
    dev = torch.device("cuda")
    N = D * Hh * W
    hA = torch.rand((1, 19, D, Hh, W), device=dev) * 1e-3
    hB = torch.zeros_like(hA)
    mfl = I.shell_flags(res, open_axes=(1,), device=dev)
    rhs = torch.rand((1, 1, D, Hh, W), device=dev) * 1e-3
    psi = torch.zeros_like(rhs)
    run = lambda: H.launch("lbmff_magnetic_sweeps_d3q19", dev, H.ptr(hA), H.ptr(hB), H.ptr(mfl), H.ptr(rhs),  # noqa
                           H.ptr(psi), n, 1.0, 1.0, 1.0, 1, D, Hh, W)
    run()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / n)
    print(f"sweep variant={os.environ.get('LBMFF_SWEEP_VARIANT', 'default')} res={res} ms={best:.4f} "
          f"GB/s={152 * N / best / 1e6:.1f} pages/pop={N * 4 / 2**21:.2f}", flush=True)


def ssweep(res, n=30, reps=3):
    """tau = 1 scalar sweeps through lbmff_magnetic_sweeps_scalar_3d: psi-only intermediate states (default) against
    the two-float sweeps (LBMFF_MAG_PSI=0), ms per sweep (the expand kernel included in the psi-only figure)."""
    D, Hh, W = res
    dev = torch.device("cuda")
    N = D * Hh * W
    aA = torch.rand((1, 2, D, Hh, W), device=dev) * 1e-3
    aB = torch.rand((1, 2, D, Hh, W), device=dev) * 1e-3
    mfl = I.shell_flags(res, open_axes=(1,), device=dev)
    rhs = torch.rand((1, 1, D, Hh, W), device=dev) * 1e-3
    psi = torch.zeros_like(rhs)
    run = lambda: H.launch("lbmff_magnetic_sweeps_scalar_3d", dev, H.ptr(aA), H.ptr(aB), H.ptr(mfl), H.ptr(rhs),  # noqa
                           H.ptr(psi), n, 1.0, 1.0, 1, D, Hh, W)
    for mode, rows in (("1", "4"), ("1", "2"), ("1", "1"), ("0", "-")):
        os.environ["LBMFF_MAG_PSI"] = mode
        os.environ["LBMFF_MAG_PSI_ROWS"] = rows
        ms = timed(run, reps) / n
        model = 12 if mode == "1" else 20
        print(f"scalar sweeps psi_only={mode} rows={rows} res={res} n={n}: {ms:.4f} ms per sweep, "
              f"{model * N / ms / 1e6:.0f} GB/s of the {model} B/cell model", flush=True)
    os.environ.pop("LBMFF_MAG_PSI", None)
    os.environ.pop("LBMFF_MAG_PSI_ROWS", None)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def streams(res):
    """Ceilings: one-stream copy (torch), 19-in/19-out shifted copy (propagation kernel, no arithmetic)."""
    D, Hh, W = res
    dev = torch.device("cuda")
    N = D * Hh * W
    a = torch.rand((1, 19, D, Hh, W), device=dev)
    b = torch.zeros_like(a)
    ms = timed(lambda: b.copy_(a))
    print(f"torch copy 19N floats: ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)
    ms = timed(lambda: H.launch("lbmff_propagation_d3q19", dev, H.ptr(a), H.ptr(b), 1, D, Hh, W))
    print(f"propagation d3q19 (38 streams, no math): ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)
    # same bytes as 19 separate one-population copies issued back to back (2 streams at a time)
    ms = timed(lambda: [b[0, q].copy_(a[0, q]) for q in range(19)])
    print(f"19 one-population copies: ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)


def single2d(n, steps=50):
    """demo_2d_LBM(_KBC) at n x n: SinglePhaseSolver BGK and KBC, MLUPS and GB/s at 72 B/cell."""
    from lbm_ferrofluid_b200.solver import SinglePhaseSolver
    from lbm_ferrofluid_b200.types import KBCType
    dev = torch.device("cuda")
    flags = torch.ones((1, 1, n, n), dtype=torch.uint8, device=dev)
    tau = 0.5 + 3 * 0.2 * n / 3000
    vel = I.double_shear_layer((n, n), device=dev)
    rho = torch.full((1, 1, n, n), 0.265, device=dev)
    for name, kbc in (("BGK", None), ("KBC", int(KBCType.KBC_A))):
        s = SinglePhaseSolver(flags, tau=tau, KBC_type=kbc).initialize(rho, vel)
        ms = timed(lambda: s.step(steps), reps=3) / steps
        print(f"D2Q9 {name} {n}^2: {ms:.4f} ms/step {n * n / ms / 1e3:.0f} MLUPS {72 * n * n / ms / 1e6:.0f} GB/s "
              f"finite={bool(torch.isfinite(s.f).all())}", flush=True)


def ros2d(n, steps=10):
    """demo_2d_LBM_Rosensweig_instability at n x n (HCZ + magnetic + KBC on g), both magnetic paths."""
    import math
    from lbm_ferrofluid_b200.solver import FerroSolver
    from lbm_ferrofluid_b200.types import KBCType
    dev = torch.device("cuda")
    res = (n, n)
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res, flags, frac=0.45)
    for path in ("faithful", "scalar"):
        s = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, 100.0], tau_f=0.7, tau_g=0.8,
                        contact_angle=0.75 * math.pi, KBC_type=int(KBCType.KBC_A), magnetic_path=path)
        s.initialize(rho)
        ms = timed(lambda: s.step(steps), reps=3) / steps
        print(f"2-D Rosensweig {path} {n}^2: {ms:.4f} ms/step {n * n / ms / 1e3:.0f} MLUPS "
              f"({32 * 72 * n * n / ms / 1e6:.0f} GB/s of the 2304 B/cell model) finite={bool(torch.isfinite(s.rho).all())}",
              flush=True)


def cg(res, steps=6, tol=1e-4):
    """3-D Rosensweig with magnetic_path="converged": iterations of the cold and the warm-started solves, ms per CG
    iteration (5 launches) against one population-faithful sweep, and the whole step against the 30-sweep step."""
    import math
    from lbm_ferrofluid_b200.poisson import PoissonCG
    from lbm_ferrofluid_b200.solver import FerroSolver
    dev = torch.device("cuda")
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res, flags)
    N = res[0] * res[1] * res[2]
    mk = lambda path, **kw: FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, 100.0, 0.0], magnetic_path=path,  # noqa
                                        contact_angle=0.5 * math.pi, gravity=1e-4, keep_force=False, **kw).initialize(rho)
    s = mk("converged", cg_tol=tol, cg_max_iters=3000)
    its = []
    for _ in range(steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); s.step(); b.record()
        torch.cuda.synchronize()
        i = s.cg.info()[0]
        its.append((i["iters"], i["converged"], round(a.elapsed_time(b), 2)))
    print(f"cg res={res} tol={tol}: (iters, converged, ms) per step = {its}", flush=True)
    # cost of one iteration: a solve that cannot converge (tol = 0) with a fixed iteration count
    rhs = s._rhs.clone()
    fixed = PoissonCG(mflags, tol=0.0, max_iters=50)
    psi = torch.zeros_like(rhs)
    ms = timed(lambda: fixed.solve(rhs, psi), reps=3) / 50
    print(f"cg iteration: {ms:.3f} ms = {N / ms / 1e6:.1f} GLUPS-iter, {56 * N / ms / 1e6:.0f} GB/s of a 56 B/cell model",
          flush=True)
    del s, fixed
    t = mk("scalar")
    ms30 = timed(lambda: t.step(), reps=3)
    print(f"30-sweep scalar step: {ms30:.2f} ms", flush=True)


def pcg(res, tol=1e-4, steps=True):
    """Cold solve of the first Rosensweig step's magnetic potential: plain CG against multigrid-preconditioned CG
    (iterations, ms, residual history), and the warm-started steps of a converged-path solver with each."""
    import math
    from lbm_ferrofluid_b200.poisson import PoissonCG
    from lbm_ferrofluid_b200.solver import FerroSolver
    dev = torch.device("cuda")
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res, flags)
    mk = lambda **kw: FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, 100.0, 0.0], magnetic_path="converged",  # noqa
                                  contact_angle=0.5 * math.pi, gravity=1e-4, keep_force=False, cg_tol=tol,
                                  **kw).initialize(rho)
    s = mk(cg_max_iters=3000)
    s.step()
    rhs = s._rhs.clone()
    del s
    for name, kw in (("cg", dict(max_iters=3000)), ("pcg-mg", dict(max_iters=200, precondition="mg"))):
        if name == "cg" and not steps:
            continue
        sol = PoissonCG(mflags, tol=tol, keep_history=True, **kw)
        psi = torch.zeros_like(rhs)
        sol.solve(rhs, psi)                      # builds the hierarchy / graph
        torch.cuda.synchronize()
        psi.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sol.solve(rhs, psi); b.record()
        torch.cuda.synchronize()
        i = sol.info()[0]
        h = sol.history()[0]
        hist = [f"{float(v) / i['rhs_norm']:.1e}" for v in h[: i["iters"] + 1]]
        print(f"{name} cold res={res} tol={tol}: iters={i['iters']} converged={i['converged']} "
              f"{a.elapsed_time(b):.2f} ms ({a.elapsed_time(b) / max(1, i['iters']):.3f} ms/iter)", flush=True)
        print(f"  |r|/|b| history: {hist if len(hist) <= 40 else hist[:20] + ['...'] + hist[-5:]}", flush=True)
        del sol
    for name, kw in (("cg", dict(cg_max_iters=3000)), ("pcg-mg", dict(cg_max_iters=200, cg_precondition="mg"))):
        if not steps:
            break
        s = mk(**kw)
        its = []
        for _ in range(6):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); s.step(); b.record()
            torch.cuda.synchronize()
            i = s.cg.info()[0]
            its.append((i["iters"], i["converged"], round(a.elapsed_time(b), 2)))
        print(f"{name} steps res={res}: (iters, converged, ms) = {its}", flush=True)
        del s


def pcgprof(res, iters=4):
    """A fixed number of multigrid-PCG iterations on the first Rosensweig step's right-hand side (for ncu launch lists;
    run with LBMFF_CG_NO_GRAPH=1 so that every kernel is a plain launch)."""
    import math
    from lbm_ferrofluid_b200.poisson import PoissonCG
    from lbm_ferrofluid_b200.solver import FerroSolver
    dev = torch.device("cuda")
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res, flags)
    s = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, 100.0, 0.0], magnetic_path="scalar",
                    contact_angle=0.5 * math.pi, gravity=1e-4, keep_force=False).initialize(rho)
    s.step()
    rhs = s._rhs.clone()
    del s
    sol = PoissonCG(mflags, tol=0.0, max_iters=iters, precondition="mg")
    psi = torch.zeros_like(rhs)
    sol.solve(rhs, psi)
    torch.cuda.synchronize()
    print("pcgprof", sol.info()[0], flush=True)


def ext(res, steps=20, reps=3):
    """single-phase BGK on D3Q19 / D3Q27, forcing none / Guo, A/B pull vs esoteric pull in place."""
    from lbm_ferrofluid_b200.extensions import ExtSinglePhaseSolver
    from lbm_ferrofluid_b200.solver import SinglePhaseSolver
    dev = torch.device("cuda")
    D, Hh, W = res
    N = D * Hh * W
    flags = torch.ones((1, 1, D, Hh, W), dtype=torch.uint8, device=dev)
    flags[:, :, :, 0] = 2; flags[:, :, :, -1] = 2
    rho = torch.full((1, 1, D, Hh, W), 0.265, device=dev)
    vel = torch.zeros((1, 3, D, Hh, W), device=dev)
    vel[:, 0] = 0.05
    s = SinglePhaseSolver(flags, tau=0.8).initialize(rho, vel)
    ms = timed(lambda: s.step(steps), reps=reps) / steps
    print(f"pinned D3Q19 BGK A/B {res}: {ms:.4f} ms/step {N / ms / 1e3:.0f} MLUPS {152 * N / ms / 1e6:.0f} GB/s", flush=True)
    del s
    for lattice, Q in (("D3Q19", 19), ("D3Q27", 27)):
        for forcing in (None, "guo"):
            for streaming in ("ab", "esoteric"):
                s = ExtSinglePhaseSolver(flags, tau=0.8, lattice=lattice, forcing=forcing, streaming=streaming,
                                         force=(1e-6, 0.0, 0.0) if forcing else None).initialize(rho, vel)
                ms = timed(lambda: s.step(steps), reps=reps) / steps
                print(f"ext {lattice} forcing={forcing} {streaming} {res}: {ms:.4f} ms/step {N / ms / 1e3:.0f} MLUPS "
                      f"{8 * Q * N / ms / 1e6:.0f} GB/s pop_bytes={s.population_bytes / 1e9:.2f} GB "
                      f"finite={bool(torch.isfinite(s.rho).all())}", flush=True)
                del s
                torch.cuda.empty_cache()
    for lattice, Q in (("D3Q19", 19), ("D3Q27", 27)):
        for streaming in ("ab", "esoteric"):
            s = ExtSinglePhaseSolver(flags, tau=0.8, lattice=lattice, streaming=streaming, storage="fp16").initialize(rho, vel)
            ms = timed(lambda: s.step(steps), reps=reps) / steps
            print(f"ext {lattice} fp16 storage {streaming} {res}: {ms:.4f} ms/step {N / ms / 1e3:.0f} MLUPS "
                  f"{4 * Q * N / ms / 1e6:.0f} GB/s pop_bytes={s.population_bytes / 1e9:.2f} GB "
                  f"finite={bool(torch.isfinite(s.rho).all())}", flush=True)
            del s
            torch.cuda.empty_cache()
    vel = vel + 0.01 * torch.rand_like(vel)      # KBC needs a non-equilibrium state to have work to do
    for streaming in ("ab", "esoteric"):
        s = ExtSinglePhaseSolver(flags, tau=0.51, lattice="D3Q27", collision="kbc", streaming=streaming).initialize(rho, vel)
        ms = timed(lambda: s.step(steps), reps=reps) / steps
        print(f"ext D3Q27 KBC {streaming} {res}: {ms:.4f} ms/step {N / ms / 1e3:.0f} MLUPS {8 * 27 * N / ms / 1e6:.0f} GB/s "
              f"finite={bool(torch.isfinite(s.rho).all())}", flush=True)
        del s
        torch.cuda.empty_cache()
    for lattice, Q in (("D3Q19", 19), ("D3Q27", 27)):
        for streaming in ("ab", "esoteric"):
            s = ExtSinglePhaseSolver(flags, tau=0.55, lattice=lattice, collision="mrt", tau_bulk=0.9, tau_high=1.0,
                                     streaming=streaming).initialize(rho, vel)
            ms = timed(lambda: s.step(steps), reps=reps) / steps
            print(f"ext {lattice} MRT {streaming} {res}: {ms:.4f} ms/step {N / ms / 1e3:.0f} MLUPS {8 * Q * N / ms / 1e6:.0f} GB/s "
                  f"finite={bool(torch.isfinite(s.rho).all())}", flush=True)
            del s
            torch.cuda.empty_cache()


def cs(res, steps=5):
    """The HCZ collide-stream stage inside real steps (scalar magnetic path): cp.async kernel vs TMA kernel, per kz."""
    import math
    import statistics
    from lbm_ferrofluid_b200.solver import FerroSolver
    dev = torch.device("cuda")
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res, flags)
    N = res[0] * res[1] * res[2]
    s = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, 100.0, 0.0], magnetic_path="scalar",
                    contact_angle=0.5 * math.pi, gravity=1e-4, keep_force=False).initialize(rho)
    s.step(3)
    for tma in os.environ.get("KB_TMA", "0 1").split():
        for kz in os.environ.get("KB_KZ", "8 16 32").split():
            os.environ["LBMFF_HCZ_TMA"], os.environ["LBMFF_HCZ_KZ"] = tma, kz
            s.step(1)
            s.enable_kernel_timing(True)
            s.step(steps)
            torch.cuda.synchronize()
            ms = statistics.median(s.kernel_times_ms()["collide_stream"])
            s.enable_kernel_timing(False)
            print(f"collide-stream tma={tma} kz={kz} {res}: {ms:.3f} ms  {304 * N / ms / 1e6:.0f} GB/s of the 304 B/cell "
                  f"model  finite={bool(torch.isfinite(s.rho).all())}", flush=True)


def render(n, image=256):
    """The demo frame (save_rendered_image): iso-surface of a perturbed Rosensweig phi at n^3, then the SoftRenderer
    chain at image^2 (x2 supersampling), each stage timed with CUDA events."""
    import math
    from lbm_ferrofluid_b200.isosurface import marching_cubes
    from lbm_ferrofluid_b200.render import SoftRenderer, soft_rasterize
    dev = torch.device("cuda")
    res = (n, n, n)
    flags = I.shell_flags(res, device=dev)
    rho = I.rosensweig_rho(res, flags, perturb=0.03 * n)
    u = (2.0 * (rho - 0.02381) / (0.2508 - 0.02381) - 1.0)[0, 0].contiguous() * (2.0 / n)
    ms_mc = timed(lambda: marching_cubes(u, 0.0), reps=3)
    verts, faces = marching_cubes(u, 0.0)
    verts = (verts - n / 2.0) * (2.0 / n)
    ren = SoftRenderer(image_size=image, camera_mode="look_at", eye=torch.tensor([1.6, 0.8, 1.6], device=dev), near=1.0,
                       bg_color=torch.tensor([1.0, 1.0, 1.0], device=dev), device=dev)
    tex = 0.68 * torch.ones(1, faces.shape[0], 2, 3, device=dev)
    ms_all = timed(lambda: ren.forward(verts[None], faces[None], tex), reps=3)
    lit = ren.lighting(verts[None], faces[None], tex)
    fv = ren.vertices_to_faces(ren.project_to_image(ren.transform_to_camera_frame(verts[None])), faces[None]).contiguous()
    ms_ras = timed(lambda: soft_rasterize(fv, lit, 2 * image, (1.0, 1.0, 1.0), 1.0, 100), reps=3)
    F = faces.shape[0]
    print(f"render n={n}: faces={F} marching_cubes {ms_mc:.3f} ms ({4 * n ** 3 / ms_mc / 1e6:.0f} GB/s of the field read "
          f"twice = {8 * n ** 3 / ms_mc / 1e6:.0f} GB/s), renderer.forward {ms_all:.3f} ms, rasterise {(2 * image)}^2: "
          f"{ms_ras:.3f} ms", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "render":
        render(int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 256)
    if sys.argv[1] == "cs":
        cs(tuple(int(v) for v in sys.argv[2:5]))
    if sys.argv[1] == "ext":
        ext(tuple(int(v) for v in sys.argv[2:5]), *( [int(sys.argv[5]), 1] if len(sys.argv) > 5 else []))
    if sys.argv[1] == "pcgcold":
        pcg(tuple(int(v) for v in sys.argv[2:5]), steps=False)
    if sys.argv[1] == "pcgprof":
        pcgprof(tuple(int(v) for v in sys.argv[2:5]))
    if sys.argv[1] == "pcg":
        pcg(tuple(int(v) for v in sys.argv[2:5]), tol=float(sys.argv[5]) if len(sys.argv) > 5 else 1e-4)
    if sys.argv[1] == "cg":
        cg(tuple(int(v) for v in sys.argv[2:5]), tol=float(sys.argv[5]) if len(sys.argv) > 5 else 1e-4)
    if sys.argv[1] == "single2d":
        single2d(int(sys.argv[2]))
    if sys.argv[1] == "ros2d":
        ros2d(int(sys.argv[2]))
    if sys.argv[1] == "streams":
        streams(tuple(int(v) for v in sys.argv[2:5]))
    if sys.argv[1] == "ssweep":
        ssweep(tuple(int(v) for v in sys.argv[2:5]), int(sys.argv[5]) if len(sys.argv) > 5 else 30)
    if sys.argv[1] == "sweep":
        sweep(tuple(int(v) for v in sys.argv[2:5]), int(sys.argv[5]) if len(sys.argv) > 5 else 30)
