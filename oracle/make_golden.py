"""Generate golden vectors from the UNMODIFIED reference (run in the build container only).

    python oracle/make_golden.py            # writes tests/golden/*.npz

The reference (``/root/reference``, pure PyTorch) ships no tests or golden
vectors, so parity is pinned by executing its own code
here on CPU and committing the (small) outputs.  The reference cannot travel to
the GPU box; these fixtures can.  Viz-only imports that are not installed
(matplotlib, mcubes, imageio; ``src/LBM/utils/__init__.py:1-3``) are stubbed.
Nothing in this file is used by the product path.
"""
from __future__ import annotations

import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
PYREF = os.path.join(HERE, "_ref", "pyref.zip")      # sourceless bytecode of the reference (build_ref_solver.py)
REF = os.environ.get("LBMFF_REFERENCE") or ("/root/reference" if os.path.isdir("/root/reference/src") else PYREF)
OUT = os.path.join(HERE, "..", "tests", "golden")


def import_reference():
    """Import the UNMODIFIED reference solver from REF.  The repo ships its own ``src/LBM`` (the drop-in shim, a
    regular package), which would shadow the reference's namespace package whenever the repo root is on sys.path --
    so the result is checked: ``src`` must resolve inside REF, else this raises instead of silently testing the
    product against itself."""
    for name in ("matplotlib", "matplotlib.colors", "matplotlib.ticker", "mcubes", "imageio"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib.colors"].ListedColormap = object
    sys.modules["matplotlib"].colors = sys.modules["matplotlib.colors"]
    sys.modules["matplotlib"].ticker = sys.modules["matplotlib.ticker"]
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import src.LBM.simulation as _sim
    where = os.path.realpath(os.path.dirname(_sim.__file__))
    if not where.startswith(os.path.realpath(REF) + os.sep):
        raise ImportError(f"`src.LBM` resolved to {where}, not inside the reference at {REF}: the repo's own src/ shim "
                          "is shadowing it (run this file as a script, or through oracle/ref_runner.py, with the "
                          "repo root OFF sys.path)")
    from src.LBM.simulation import SimulationParameters, SimulationRunner  # noqa
    from src.LBM.utils import CellType, KBCType, get_staggered  # noqa
    return SimulationParameters, SimulationRunner, CellType, KBCType, get_staggered


def rng_tensor(rs, shape, lo, hi, dtype):
    return torch.from_numpy(rs.uniform(lo, hi, size=shape)).to(dtype)


def shell_flags(res, open_axes=(), interior_obstacles=0, rs=None):
    """OBSTACLE shell (except along ``open_axes``: 0=x,1=y,2=z), FLUID inside."""
    dim = len(res)
    fl = torch.full((1, 1, *res), 2, dtype=torch.uint8)
    sl = [Ellipsis] + [slice(1, -1)] * dim
    for a in open_axes:
        sl[len(sl) - 1 - a] = slice(None)
    fl[tuple(sl)] = 1
    if interior_obstacles:
        for _ in range(interior_obstacles):
            idx = tuple(int(rs.randint(1, n - 1)) for n in res)
            fl[(0, 0) + idx] = 2
    return fl


def make_runner(SimulationParameters, SimulationRunner, dim, dtype, res, **kw):
    device = torch.device(kw.get("device", "cpu"))
    params = SimulationParameters(
        dim=dim, dtype=dtype, device=device, simulation_size=(1, 1, *res),
        dt=kw.get("dt", 1.0),
        density_gas=kw.get("density_gas", 0.02381), density_fluid=kw.get("density_fluid", 0.2508),
        contact_angle=torch.tensor([kw.get("contact_angle", 0.5 * math.pi)]).to(dtype).to(device),
        Q=9 if dim == 2 else 19, tau=kw.get("tau", 1.0),
        rho_gas=kw.get("rho_gas", 0.02381), rho_fluid=kw.get("rho_fluid", 0.2508),
        kappa=kw.get("kappa", 0.01), tau_g=kw.get("tau_g", 0.68), tau_f=kw.get("tau_f", 0.68), k=0.33)
    return SimulationRunner(parameters=params)


def t2n(d):
    return {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in d.items()}


# ---------------------------------------------------------------------------
# per-op vectors
# ---------------------------------------------------------------------------
def gen_ops(dim, dtype, seed, dx, dt, tag):
    SP, SR, CellType, KBCType, get_staggered = import_reference()
    rs = np.random.RandomState(seed)
    res = (6, 7, 9) if dim == 3 else (9, 11)
    Q = 19 if dim == 3 else 9
    kw = dict(dt=dt, tau=0.8, tau_f=0.7, tau_g=0.85, kappa=0.02, contact_angle=0.6 * math.pi)
    runner = make_runner(SP, SR, dim, dtype, res, **kw)
    prop, macro = runner.create_propagation(), runner.create_macro_compute()
    col, hcz, mag = runner.create_collision(), runner.create_collision_HCZ(), runner.create_LBM_magnetic()
    hcz.set_gravity(gravity=2e-4)

    out = {"dx": dx, "dt": dt, "res": np.array(res)}
    f = rng_tensor(rs, (1, Q, *res), 0.005, 0.05, dtype)
    g = rng_tensor(rs, (1, Q, *res), 0.005, 0.05, dtype)
    h = rng_tensor(rs, (1, Q, *res), -1.0, 1.0, dtype)
    rho = rng_tensor(rs, (1, 1, *res), 0.03, 0.26, dtype)
    density = rng_tensor(rs, (1, 1, *res), 0.03, 0.26, dtype)
    pressure = rng_tensor(rs, (1, 1, *res), 0.001, 0.01, dtype)
    vel = rng_tensor(rs, (1, dim, *res), -0.08, 0.08, dtype) * (dx / dt)
    force = rng_tensor(rs, (1, dim, *res), -1e-3, 1e-3, dtype)
    phi = rng_tensor(rs, (1, 1, *res), -1.2, 1.2, dtype)
    # put a band of phi inside the smoothing width so the sin branch is exercised
    phi[..., 2, :] = rng_tensor(rs, phi[..., 2, :].shape, -0.1 * dx, 0.1 * dx, dtype)
    H2 = rng_tensor(rs, (1, 1, *res), 50.0, 150.0, dtype)
    Hext = rng_tensor(rs, (1, dim, *res), -1.0, 1.0, dtype)
    flags_io = shell_flags(res, (), 5, rs)            # shell + interior obstacles
    flags = shell_flags(res)                          # closed box
    mflags = shell_flags(res, open_axes=(1,))         # y open (Rosensweig magnetic flags)
    out.update(t2n(dict(f=f, g=g, h=h, rho=rho, density=density, pressure=pressure, vel=vel, force=force,
                        phi=phi, H2=H2, Hext=Hext, flags_io=flags_io, flags=flags, mflags=mflags)))

    out["propagation"] = prop.propagation(f=f.clone())
    out["rebounce"] = prop.rebounce_obstacle(f=f.clone(), flags=flags_io)
    r, v, d = macro.macro_compute(dx=dx, dt=dt, f=f, rho=rho.clone(), vel=vel.clone(), flags=flags_io,
                                  density=density.clone())
    out["macro_rho"], out["macro_vel"], out["macro_density"] = r, v, d
    out["get_pressure"] = macro.get_pressure(dx=dx, dt=dt, density=density)
    out["eos"] = col.equation_of_states(dx=dx, dt=dt, rho=rho)
    out["feq"] = col.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel, force=None)
    out["feq_force"] = col.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel, force=force)
    out["geq"] = col.get_geq_(dx=dx, dt=dt, rho=rho, vel=vel, density=density, pressure=pressure, force=None)
    out["grad_io"] = type(col).get_grad(input_=rho, dx=dx, flags=flags_io)
    out["grad_open"] = type(col).get_grad(input_=rho, dx=dx, flags=mflags)
    out["laplacian"] = col.get_laplacian(input_=rho, dx=dx, flags=flags)
    out["bgk"] = col.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags_io, force=force)
    out["smooth_phi"] = hcz.smooth_phi(phi=phi, eps=0.1 * dx)
    out["gamma"] = hcz.compute_Gamma(dx=dx, dt=dt, vel=vel)
    stag = get_staggered(Hext)
    for a, s in enumerate(stag):
        out[f"stag{a}"] = s
    mag._tau = 1.0   # the value every magnetic demo uses (SimulationParameters default)
    Hi, hn = mag.get_H_int(dt=dt, dx=dx, phi=phi, flags=mflags, H_ext_mac=stag, h=h.clone())
    out["mag_H_int"], out["mag_h"] = Hi, hn
    mag._tau = 0.9   # exercise the tau != 1 branch too
    Hi, hn = mag.get_H_int(dt=dt, dx=dx, phi=phi, flags=flags_io, H_ext_mac=stag, h=h.clone())
    out["mag09_H_int"], out["mag09_h"] = Hi, hn

    for name, use_h2 in (("cap", True), ("capnoh", False)):
        rr = rho.clone()
        res_ = hcz.capillary_process(rho=rr, vel=vel.clone(), flags=flags, force=force.clone(), dt=dt, dx=dx,
                                     g=g, density=density.clone(), pressure=pressure.clone(),
                                     H2=H2 if use_h2 else None, phi=phi if use_h2 else None)
        for k, t in zip(("rho", "vel", "density", "pressure", "force", "dfai", "dprho"), res_):
            out[f"{name}_{k}"] = t
    if dim == 2:
        hcz.preset_KBC(dx=dx, dt=dt)
    for name, kbc in (("hcz", None),) + ((("hczkbc", int(KBCType.KBC_A)),) if dim == 2 else ()):
        fn, gn = hcz.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, density=density, flags=flags_io, force=force,
                               g=g, pressure=pressure, dfai=out["cap_dfai"], dprho=out["cap_dprho"], KBC_type=kbc)
        out[f"{name}_f"], out[f"{name}_g"] = fn, gn
    if dim == 2:
        mrt = runner.create_collision_MRT()
        mrt.preset_KBC(dx=dx, dt=dt)
        # near-equilibrium f so that the entropic stabiliser is well conditioned
        feq = mrt.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel, force=None)
        fk = feq * (1.0 + rng_tensor(rs, feq.shape, -0.05, 0.05, dtype))
        out["kbc_in_f"] = fk
        for nm, kt in (("kbcA", KBCType.KBC_A), ("kbcD", KBCType.KBC_D)):
            out[f"{nm}_s"] = mrt.get_s_by_KBC(dx=dx, dt=dt, f=fk, rho=rho, vel=vel, KBC_type=int(kt))
            out[nm] = mrt.collision(dx=dx, dt=dt, f=fk, rho=rho, vel=vel, flags=flags_io, force=force,
                                    KBC_type=int(kt))
        out["kbcNone_s"] = mrt.get_s_by_KBC(dx=dx, dt=dt, f=fk, rho=rho, vel=vel, KBC_type=None)
        sc = runner.create_collision_SC()
        out["sc_force"] = sc.calculate_force(dx=dx, dt=dt, density=density, flags=flags_io, force=force)
        out["sc"] = sc.collision(dx=dx, dt=dt, f=f, rho=rho, density=density, vel=vel, flags=flags_io, force=force)
        out["vort"] = macro.get_vort(vel=vel, dx=dx)
    np.savez_compressed(os.path.join(OUT, f"ops_{dim}d_{tag}.npz"), **t2n(out))


# ---------------------------------------------------------------------------
# trajectories (the demo loops, vectorised synthetic initial conditions)
# ---------------------------------------------------------------------------
def perturbed_interface(res, frac, rho_l, rho_g, dtype, droplets=False):
    """flat interface at frac*H perturbed by 2 sin(4 pi x/W) cos(4 pi z/D) cells,
    or (droplets) two spheres of radius min(res)//4 at 3/8 and 5/8 of W."""
    dim = len(res)
    grids = torch.meshgrid(*[torch.arange(n, dtype=torch.float64) for n in res], indexing="ij")
    if dim == 3:
        z, y, x = grids
        D, H, W = res
    else:
        y, x = grids
        z = torch.zeros_like(x)
        H, W = res
        D = 1
    if droplets:   # the rule of demo_3d_LBM_two_droplets.py:130-149: integer centres, <= radius^2
        r = min(res) // 4
        liquid = torch.zeros_like(x, dtype=torch.bool)
        for cx in (3 * W // 8, 5 * W // 8):
            d2 = (x - cx) ** 2 + (y - H // 2) ** 2 + ((z - D // 2) ** 2 if dim == 3 else 0)
            liquid |= d2 <= r * r
    else:
        height = frac * H + 2.0 * torch.sin(4 * math.pi * x / W) * (torch.cos(4 * math.pi * z / D) if dim == 3 else 1.0)
        liquid = y < height
    rho = torch.where(liquid, torch.tensor(rho_l, dtype=torch.float64), torch.tensor(rho_g, dtype=torch.float64))
    return rho.to(dtype).reshape(1, 1, *res)


def run_hcz(dim, res, steps, dtype=torch.float32, *, magnetic=True, mag=100.0, gravity=1e-4, kappa=0.01,
            tau_f=0.68, tau_g=0.68, contact_angle=0.5 * math.pi, frac=0.5, open_axes=(1,), kbc=False,
            droplets=False, n_sweeps=30, seed=0, device="cpu", snapshots=(), noise=1e-3, timer=None):
    """The demo loop (demo_3d_LBM_Rosensweig_instability.py:151-209 and its 2-D / two-droplet / multiphase siblings)
    driven through the reference's own classes on ``device``.  Returns {step: state dict} for every step listed in
    ``snapshots`` plus the final one; ``timer(step)`` (optional) is called before every step (ref_runner's clock)."""
    SP, SR, CellType, KBCType, get_staggered = import_reference()
    rho_g, rho_l = 0.02381, 0.2508
    dev = torch.device(device)
    runner = make_runner(SP, SR, dim, dtype, res, kappa=kappa, tau_f=tau_f, tau_g=tau_g, contact_angle=contact_angle,
                         device=dev)
    prop, macro = runner.create_propagation(), runner.create_macro_compute()
    col, mgf = runner.create_collision_HCZ(), runner.create_LBM_magnetic()
    if kbc:
        col.preset_KBC(dx=1.0, dt=1.0)
    col.set_gravity(gravity=gravity)
    Q = 9 if dim == 2 else 19
    dx = dt = 1.0
    flags = shell_flags(res)
    mflags = shell_flags(res, open_axes=open_axes)
    rho = perturbed_interface(res, frac, rho_l, rho_g, dtype, droplets)
    torch.manual_seed(seed)
    rho = rho * (1.0 + noise * (torch.rand(rho.shape, dtype=torch.float64).to(dtype) - 0.5))
    rho[flags == 2] = rho_l
    init = dict(rho0=rho.clone(), flags=flags, mflags=mflags)
    flags, mflags, rho = flags.to(dev), mflags.to(dev), rho.to(dev)
    density = rho.clone()
    vel = torch.zeros((1, dim, *res), dtype=dtype, device=dev)
    force = torch.zeros((1, dim, *res), dtype=dtype, device=dev)
    pressure = macro.get_pressure(dx=dx, dt=dt, density=density)
    f = col.get_feq_(dx=dx, dt=dt, rho=density, vel=vel, force=force)
    g = col.get_geq_(dx=dx, dt=dt, rho=rho, vel=vel, density=density, pressure=pressure, force=force, feq=f)
    h = torch.zeros((1, Q, *res), dtype=dtype, device=dev)
    Hext = torch.zeros((1, dim, *res), dtype=dtype, device=dev)
    Hext[:, 1] = mag
    Hmac = get_staggered(Hext, mode="replicate")
    psi = torch.zeros_like(rho)
    mass0 = float(f.double().sum().item())       # total mass over ALL cells, walls included
    meta = dict(res=np.array(res), mag=mag, gravity=gravity, kappa=kappa, tau_f=tau_f, tau_g=tau_g,
                contact_angle=contact_angle, magnetic=int(magnetic), kbc=int(kbc), n_sweeps=n_sweeps)
    out = {}
    for step in range(1, steps + 1):
        if timer is not None:
            timer(step)
        f = prop.propagation(f=f)
        g = prop.propagation(f=g)
        rho, vel, density = macro.macro_compute(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, density=density)
        f = prop.rebounce_obstacle(f=f, flags=flags)
        g = prop.rebounce_obstacle(f=g, flags=flags)
        H2 = phi = None
        if magnetic:
            phi = -(2.0 * (density - rho_g) / (rho_l - rho_g) - 1.0)
            for i in range(n_sweeps):
                h = prop.propagation(f=h)
                h = prop.rebounce_obstacle(f=h, flags=mflags)
                if i == n_sweeps - 1:
                    psi = h[:, 1:].sum(dim=1).unsqueeze(1) / (1.0 - mgf._weight[:, 0:1])
                H_int, h = mgf.get_H_int(dt=dt, dx=dx, phi=phi, flags=mflags, H_ext_mac=Hmac, h=h)
            H2 = ((Hext + H_int) * (Hext + H_int)).sum(dim=1).unsqueeze(1)
        rho, vel, density, pressure, force, dfai, dprho = col.capillary_process(
            rho=rho, vel=vel, flags=flags, force=force, dt=dt, dx=dx, g=g, density=density, pressure=pressure,
            H2=H2, phi=phi)
        f, g = col.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, density=density, flags=flags, force=force, g=g,
                             pressure=pressure, dfai=dfai, dprho=dprho,
                             KBC_type=int(KBCType.KBC_A) if kbc else None)
        if step in snapshots or step == steps:
            st = dict(init, f=f, g=g, rho=rho, vel=vel, density=density, pressure=pressure, force=force, steps=step,
                      **meta)
            if device != "cpu" or snapshots:     # (the committed traj_* goldens keep their original key set)
                st.update(mass0=mass0, mass=float(f.double().sum().item()))
            if magnetic:
                st.update(h=h, psi=psi, H2=H2, phi=phi)
            out[step] = t2n(st)
    if timer is not None:
        timer(steps + 1)
    return out


def traj_hcz(name, dim, res, steps, dtype=torch.float32, **kw):
    out = run_hcz(dim, res, steps, dtype, **kw)[steps]
    np.savez_compressed(os.path.join(OUT, f"traj_{name}.npz"), **out)


def shear_layer(res, dtype):
    """Double shear layer of demo_2d_LBM.py:85-99, vectorised."""
    H, W = res
    j = torch.arange(H, dtype=torch.float64).reshape(H, 1).expand(H, W)
    i = torch.arange(W, dtype=torch.float64).reshape(1, W).expand(H, W)
    u0 = 0.2
    ux = torch.where(j <= H / 2, u0 * torch.tanh(80.0 * (j / H - 0.25)), u0 * torch.tanh(80.0 * (0.75 - j / H)))
    uy = 0.05 * u0 * torch.sin(2 * math.pi * (i / W + 0.25))
    return torch.stack([ux, uy]).to(dtype).reshape(1, 2, H, W)


def traj_single(name, res, steps, kbc, sc=False, dtype=torch.float32):
    SP, SR, CellType, KBCType, get_staggered = import_reference()
    tau = 0.5 + 3.0 * 0.2 * max(res) / 3000.0 if not sc else 1.0
    runner = make_runner(SP, SR, 2, dtype, res, tau=tau, density_gas=0.038, density_fluid=0.265,
                         rho_gas=0.038, rho_fluid=0.265)
    prop, macro = runner.create_propagation(), runner.create_macro_compute()
    dx = dt = 1.0
    if sc:
        col = runner.create_collision_SC()
        flags = shell_flags(res)
        rho = perturbed_interface(res, 0.5, 0.265, 0.038, dtype, droplets=True)
        vel = torch.zeros((1, 2, *res), dtype=dtype)
    else:
        col = runner.create_collision_MRT() if kbc else runner.create_collision()
        if kbc:
            col.preset_KBC(dx=dx, dt=dt)
        flags = torch.ones((1, 1, *res), dtype=torch.uint8)
        rho = torch.full((1, 1, *res), 0.265, dtype=dtype)
        vel = shear_layer(res, dtype)
    force = torch.zeros((1, 2, *res), dtype=dtype)
    f = col.get_feq_(dx=dx, dt=dt, rho=rho, vel=vel, force=force)
    init = dict(rho0=rho.clone(), vel0=vel.clone(), flags=flags, tau=tau)
    density = rho.clone()
    for _ in range(steps):
        f = prop.propagation(f=f)
        if sc:
            rho, vel, density = macro.macro_compute(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, density=density)
        else:
            rho, vel = macro.macro_compute(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags)
        f = prop.rebounce_obstacle(f=f, flags=flags)
        if sc:
            f = col.collision(dx=dx, dt=dt, f=f, rho=rho, density=density, vel=vel, flags=flags, force=force)
        elif kbc:
            f = col.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=force,
                              KBC_type=int(KBCType.KBC_A))
        else:
            f = col.collision(dx=dx, dt=dt, f=f, rho=rho, vel=vel, flags=flags, force=force)
    np.savez_compressed(os.path.join(OUT, f"traj_{name}.npz"),
                        **t2n(dict(init, f=f, rho=rho, vel=vel, steps=steps, res=np.array(res))))


# ---------------------------------------------------------------------------
# initial conditions: the demos' OWN python loops, executed unmodified
# ---------------------------------------------------------------------------
class _Captured(Exception):
    pass


def demo_initial_fields(demo, res):
    """Run ``<reference>/demo/<demo>.py::main(res=res, total_steps=0)`` unmodified on the CPU and capture the
    arguments of its first ``get_feq_`` call -- the fields its O(N) python initialisation loops produced
    (demo_3d_LBM_two_droplets.py:130-157, demo_2d_LBM.py:85-99).  The run is cut short right there."""
    import importlib.util

    import_reference()
    rs = types.ModuleType("renderutils")
    rs.SoftRenderer = lambda *a, **k: None
    sys.modules.setdefault("renderutils", rs)
    from src.LBM.LBM_collision import LBMCollision2d, LBMCollision3d

    got = {}

    def recorder(self, dx, dt, rho, vel, force=None):
        got.update(rho=rho.clone(), vel=vel.clone())
        raise _Captured

    base = os.path.join(REF, "demo", demo)
    path = base + ".py" if os.path.exists(base + ".py") else base + ".pyc"
    spec = importlib.util.spec_from_file_location("_demo_" + demo, path)
    mod = importlib.util.module_from_spec(spec)
    saved = LBMCollision2d.get_feq_, LBMCollision3d.get_feq_, torch.cuda.is_available
    cwd = os.getcwd()
    try:
        import tempfile
        os.chdir(tempfile.mkdtemp())             # the demos mkdir next to __file__ only; cwd is irrelevant but keep clean
        torch.cuda.is_available = lambda: False  # the demos pick their device from this
        spec.loader.exec_module(mod)
        LBMCollision2d.get_feq_ = LBMCollision3d.get_feq_ = recorder
        try:
            mod.main(res=list(res), total_steps=0)
        except _Captured:
            pass
    finally:
        LBMCollision2d.get_feq_, LBMCollision3d.get_feq_, torch.cuda.is_available = saved
        os.chdir(cwd)
    return got


def gen_inits():
    out = {}
    for tag, demo, res in (("drop3d_a", "demo_3d_LBM_two_droplets", (12, 10, 26)),
                           ("drop3d_b", "demo_3d_LBM_two_droplets", (9, 13, 21)),
                           ("drop2d_a", "demo_2d_LBM_two_droplets", (22, 50)),
                           ("drop2d_b", "demo_2d_LBM_two_droplets", (17, 37)),
                           ("shear_a", "demo_2d_LBM", (20, 26)),
                           ("shear_b", "demo_2d_LBM_KBC", (33, 17))):
        got = demo_initial_fields(demo, res)
        out[tag + "_res"] = np.array(res)
        out[tag + "_rho"] = got["rho"].numpy()
        out[tag + "_vel"] = got["vel"].numpy()
    np.savez_compressed(os.path.join(OUT, "init_demos.npz"), **out)


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for dim in (2, 3):
        gen_ops(dim, torch.float32, 1234 + dim, 1.0, 1.0, "unit")
        gen_ops(dim, torch.float32, 4321 + dim, 0.5, 0.25, "scaled")
        gen_ops(dim, torch.float64, 1234 + dim, 1.0, 1.0, "unit64")
    traj_hcz("ros3d", 3, (12, 14, 16), 6)
    traj_hcz("ros3d_f64", 3, (12, 14, 16), 6, dtype=torch.float64)
    traj_hcz("drop3d", 3, (12, 12, 20), 4, kappa=0.5, gravity=0.0, open_axes=(), droplets=True)
    traj_hcz("multiphase3d", 3, (10, 12, 12), 5, magnetic=False, contact_angle=0.55 * math.pi)
    traj_hcz("ros2d", 2, (24, 28), 8, tau_f=0.7, tau_g=0.8, contact_angle=0.75 * math.pi, frac=0.45,
             open_axes=(1,), kbc=True)
    traj_hcz("hcz2d", 2, (20, 22), 6, magnetic=False, tau_f=0.7, tau_g=0.8, contact_angle=0.75 * math.pi)
    traj_single("bgk2d", (32, 32), 20, kbc=False)
    traj_single("kbc2d", (32, 32), 20, kbc=True)
    traj_single("sc2d", (24, 24), 10, kbc=False, sc=True)
    gen_inits()
    print("golden vectors written to", os.path.normpath(OUT))
    for fn in sorted(os.listdir(OUT)):
        print(f"  {fn:28s} {os.path.getsize(os.path.join(OUT, fn)) / 1024:8.1f} KiB")


if __name__ == "__main__":
    main()
