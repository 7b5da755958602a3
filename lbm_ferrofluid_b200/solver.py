"""Fused time-stepping drivers on top of the C ABI's fused tier (include/lbmff.h, second half).

The reference has no solver object: each demo script owns the state tensors and spells the step out
op by op (demo_3d_LBM_Rosensweig_instability.py:151-209).  ``FerroSolver.step`` executes exactly that
sequence -- stream f,g / macro / bounce / phi / n_sweeps x (stream h, bounce h, get_H_int) / H2 /
capillary_process / HCZ collision -- with 9 + n_sweeps kernel launches instead of ~4400, and
``SinglePhaseSolver.step`` the demo_2d_LBM(_KBC) loop with one launch.  The op-by-op classes in this
package remain the drop-in API; these drivers are what bench.py times.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional, Sequence

import torch

from . import _host as H
from .collision import LBMCollision2d, LBMCollision3d
from .grid import get_staggered
from .macro_compute import LBMMacroCompute2d, LBMMacroCompute3d
from .types import CellType, KBCType

FLUID, OBSTACLE = int(CellType.FLUID), int(CellType.OBSTACLE)


def _res(t):
    return tuple(t.shape[2:])


class FerroSolver:
    """HCZ two-phase flow (+ optional ferrofluid magnetic subsystem), D2Q9 or D3Q19.

    Parameters mirror the demos: ``flags`` / ``magnetic_flags`` uint8 [B,1,(D,)H,W]; ``H_ext`` either a
    cell-centred field [B,dim,...] (then its staggered faces are built with get_staggered, as the demos
    do) or a uniform vector of ``dim`` floats; ``gravity`` the scalar given to set_gravity.
    ``magnetic_path``: "faithful" (population lattice h, any tau), "scalar" (tau == 1 collapse, walls on
    the shell only; bit-identical psi) or "auto" (scalar when valid).  "converged" is NOT the reference's
    algorithm: instead of n_sweeps relaxation sweeps it solves for their fixed point with conjugate gradients
    (poisson.PoissonCG; tau == 1 only) to relative residual ``cg_tol`` in at most ``cg_max_iters`` iterations,
    warm-started from the previous step; ``cg_precondition="mg"`` adds one geometric-multigrid V(2,2) cycle per
    iteration (csrc/poisson_mg.cuh).
    """

    def __init__(self, flags: torch.Tensor, *, density_liquid=0.2508, density_gas=0.02381, rho_liquid=0.2508,
                 rho_gas=0.02381, kappa=0.01, tau_f=0.68, tau_g=0.68, tau=1.0, contact_angle=0.5 * math.pi,
                 gravity=1e-4, dx=1.0, dt=1.0, magnetic_flags: Optional[torch.Tensor] = None,
                 H_ext=None, n_sweeps: int = 30, KBC_type: Optional[int] = None, magnetic_path: str = "auto",
                 keep_force: bool = True, cg_tol: float = 1e-4, cg_max_iters: int = 60,
                 cg_precondition: Optional[str] = None):
        self.flags = H.need_cuda(flags, "flags", torch.uint8)
        self.device = self.flags.device
        self.dim, self.B, _, self.D, self.Hh, self.W = H.dims(self.flags)
        self.res = _res(self.flags)
        self.Q = 9 if self.dim == 2 else 19
        self.dx, self.dt = float(dx), float(dt)
        self._density_liquid, self._density_gas = density_liquid, density_gas
        self._rho_liquid, self._rho_gas = rho_liquid, rho_gas
        self._kappa, self._tau_f, self._tau_g, self._tau = kappa, tau_f, tau_g, tau
        self.gravity = float(gravity)
        self.cot = H.cot_of_contact_angle(contact_angle)
        self.n_sweeps = int(n_sweeps)
        self.KBC_type = KBC_type
        if self.dim == 3 or KBC_type is None:
            self._kbc_mode = 0
        else:
            self._kbc_mode = 1 if KBCType.is_KBC(KBC_type) else 2
        bad = ((self.flags != FLUID) & (self.flags != OBSTACLE)).any().item()
        if bad:
            raise RuntimeError("FerroSolver supports FLUID / OBSTACLE flags only; use the op-by-op API otherwise")
        self.params = H.fluid_params(self, self.dx, self.dt, cot=self.cot, gravity=self.gravity)
        self.magnetic = magnetic_flags is not None
        opts = dict(dtype=torch.float32, device=self.device)
        sc = lambda c: torch.zeros((self.B, c, *self.res), **opts)  # noqa: E731
        self.rho, self.pressure, self.vel = sc(1), sc(1), sc(self.dim)
        self.force = sc(self.dim) if keep_force else None
        self._ws = sc(3)
        self._f = [sc(self.Q), sc(self.Q)]
        self._g = [sc(self.Q), sc(self.Q)]
        self._cur = 0
        self.steps_done = 0
        self._timing = None
        if self.magnetic:
            self.mflags = H.need_cuda(magnetic_flags, "magnetic_flags", torch.uint8)
            if H_ext is None:
                raise RuntimeError("magnetic_flags given without H_ext")
            if torch.is_tensor(H_ext):
                self.H_ext = H.need_cuda(H_ext, "H_ext")
                self._H_uniform = None
            else:
                vec = [float(v) for v in H_ext]
                if len(vec) != self.dim:
                    raise RuntimeError(f"uniform H_ext needs {self.dim} components")
                self._H_uniform = (ctypes.c_double * 3)(*(vec + [0.0] * (3 - self.dim)))
                self.H_ext = None
                full = torch.empty((self.B, self.dim, *self.res), **opts)
                for a, v in enumerate(vec):
                    full[:, a] = v
                H_ext = full
            self.H_ext_mac = get_staggered(H_ext if torch.is_tensor(H_ext) else full, mode="replicate")
            inner = self.mflags[(slice(None), slice(None)) + (slice(1, -1),) * self.dim]
            walls_on_shell = not (inner == OBSTACLE).any().item()
            scalar_ok = float(tau) == 1.0 and walls_on_shell
            if magnetic_path == "auto":
                magnetic_path = "scalar" if scalar_ok else "faithful"
            if magnetic_path == "scalar" and not scalar_ok:
                raise RuntimeError("magnetic_path='scalar' needs tau == 1 and every OBSTACLE cell of "
                                   "magnetic_flags on the outer shell")
            if magnetic_path == "converged" and float(tau) != 1.0:
                raise RuntimeError("magnetic_path='converged' solves the fixed point of the tau == 1 sweeps only")
            if magnetic_path not in ("scalar", "faithful", "converged"):
                raise ValueError("magnetic_path must be 'auto', 'faithful', 'scalar' or 'converged'")
            self.magnetic_path = magnetic_path
            self.psi, self._rhs, self._chi = sc(1), sc(1), sc(1)
            if magnetic_path == "faithful":
                self._h = [sc(self.Q), sc(self.Q)]
            elif magnetic_path == "scalar":
                self._a = [sc(2), sc(2)]
            else:
                from .poisson import PoissonCG
                self._h = self._a = ()
                self.cg = PoissonCG(self.mflags, dx=self.dx, dt=self.dt, tol=cg_tol, max_iters=cg_max_iters,
                                    precondition=cg_precondition)
            self._mcur = 0
        else:
            self.magnetic_path = None
            self.mflags = None

    # ---- state ------------------------------------------------------------------------------
    @property
    def f(self):
        return self._f[self._cur]

    @property
    def g(self):
        return self._g[self._cur]

    @property
    def density(self):
        """density = affine(rho), the map macro_compute / capillary_process apply (HCZ_3d.py:213-215)."""
        dg, dl, rg, rl = self._density_gas, self._density_liquid, self._rho_gas, self._rho_liquid
        return dg + (dl - dg) * ((self.rho - rg) / (rl - rg))

    @property
    def phi(self):
        """Order parameter the demos render (demo_3d...Rosensweig...py:162): one fused pass over rho
        (lbmff_order_parameter), bit-identical to the nine torch kernels of ``phi_torch``."""
        out = torch.empty_like(self.rho)
        H.launch("lbmff_order_parameter", self.device, H.ptr(self.rho), float(self._rho_gas), float(self._rho_liquid),
                 float(self._density_gas), float(self._density_liquid), H.ptr(out), self.rho.numel())
        return out

    @property
    def phi_torch(self):
        """The same field as the demos' own torch expression (the checker of ``phi``)."""
        dg, dl = self._density_gas, self._density_liquid
        return -(2.0 * (self.density - dg) / (dl - dg) - 1.0)

    @property
    def h(self):
        """Magnetic lattice populations; rebuilt from the scalar state when the scalar path is active."""
        if not self.magnetic:
            return None
        if self.magnetic_path == "faithful":
            return self._h[self._mcur]
        if self.magnetic_path == "converged":
            raise RuntimeError("magnetic_path='converged' keeps no lattice populations (psi is solved for directly)")
        out = torch.empty((self.B, self.Q, *self.res), dtype=torch.float32, device=self.device)
        H.launch(f"lbmff_magnetic_materialize_h_{self.dim}d", self.device, H.ptr(self._a[self._mcur]),
                 H.ptr(self._a[1 - self._mcur]), H.ptr(self.psi), H.ptr(self._rhs), H.ptr(self.mflags), H.ptr(out),
                 self.dx, self.dt, *H.sizes(self.dim, self.B, self.D, self.Hh, self.W))
        return out

    def initialize(self, rho: torch.Tensor, density: Optional[torch.Tensor] = None,
                   vel: Optional[torch.Tensor] = None):
        """Initial pressure, f, g from rho / density / vel exactly as the demos do
        (demo_3d_LBM_Rosensweig_instability.py:128-145): p = EOS(density), f = feq(density, vel),
        g = geq(rho, vel, density, p, feq = f); h = 0."""
        rho = H.need_cuda(rho, "rho")
        density = rho if density is None else H.need_cuda(density, "density")
        self.rho.copy_(rho)
        if vel is not None:
            self.vel.copy_(H.need_cuda(vel, "vel"))
        else:
            self.vel.zero_()
        kw = dict(density_liquid=self._density_liquid, density_gas=self._density_gas, rho_liquid=self._rho_liquid,
                  rho_gas=self._rho_gas, device=self.device)
        macro = (LBMMacroCompute2d if self.dim == 2 else LBMMacroCompute3d)(**kw)
        col = (LBMCollision2d if self.dim == 2 else LBMCollision3d)(tau=1.0, **kw)
        self.pressure.copy_(macro.get_pressure(dx=self.dx, dt=self.dt, density=density))
        zero = torch.zeros_like(self.vel)
        f0 = col.get_feq_(dx=self.dx, dt=self.dt, rho=density, vel=self.vel, force=zero)
        g0 = col.get_geq_(dx=self.dx, dt=self.dt, rho=self.rho, vel=self.vel, density=density,
                          pressure=self.pressure, force=zero, feq=f0)
        self._cur = 0
        self._f[0].copy_(f0)
        self._g[0].copy_(g0)
        if self.magnetic:
            for t in (self._h if self.magnetic_path == "faithful" else self._a):
                t.zero_()
            self.psi.zero_()
            self._mcur = 0
        self.steps_done = 0
        return self

    # ---- stepping ---------------------------------------------------------------------------
    def _sizes(self):
        return H.sizes(self.dim, self.B, self.D, self.Hh, self.W)

    def enable_kernel_timing(self, on: bool = True):
        """Bracket the magnetic sweeps and the collide-stream kernel with CUDA events on the launching
        stream (bench.py's roofline numbers); read them back with kernel_times_ms()."""
        self._timing = {"sweeps": [], "collide_stream": []} if on else None

    def _timed(self, key):
        solver = self

        class _T:
            def __enter__(self_):
                if solver._timing is not None:
                    self_.a = torch.cuda.Event(enable_timing=True)
                    self_.b = torch.cuda.Event(enable_timing=True)
                    self_.a.record()

            def __exit__(self_, *exc):
                if solver._timing is not None:
                    self_.b.record()
                    solver._timing[key].append((self_.a, self_.b))
        return _T()

    def kernel_times_ms(self):
        """{name: [ms per bracket]} -- call after a synchronize."""
        return {k: [a.elapsed_time(b) for a, b in v] for k, v in (self._timing or {}).items()}

    def _step_once(self, before_source=None, after_rho_final=None):
        """One outer step.  Hooks (used by step_host to overlap host copies): ``before_source`` runs just before
        the first kernel that reads the applied-field faces, ``after_rho_final`` right after the wetting boundary,
        when rho (hence phi) has its end-of-step value."""
        d, sz, suf = self.dim, self._sizes(), H.suffix(self.dim)
        cur, nxt = self._cur, 1 - self._cur
        dev = self.device
        with H.nvtx("lbmff:moments"):
            H.launch("lbmff_hcz_moments_" + suf, dev, H.ptr(self._f[cur]), H.ptr(self.flags), H.ptr(self.rho),
                     H.ptr(self._chi if self.magnetic else None), self.params, *sz)
        # capillary_process part 1 does not depend on the magnetic subsystem: issue it first
        with H.nvtx("lbmff:capillary_fields"):
            H.launch(f"lbmff_hcz_capillary_fields_{d}d", dev, H.ptr(self.rho), H.ptr(self.flags),
                     H.ptr(self.pressure), self.params, H.ptr(self._ws), *sz)
        if after_rho_final is not None:
            after_rho_final()
        chi = psi = mfl = hext = None
        hu = None
        if self.magnetic:
            if before_source is not None:
                before_source()
            with H.nvtx("lbmff:magnetic_source"):
                H.launch(f"lbmff_magnetic_source_{d}d", dev, H.ptr(self._chi), H.ptr(self.mflags),
                         *[H.ptr(t) for t in self.H_ext_mac], self.dx, H.ptr(self._rhs), *sz)
            if before_source is not None:   # the faces are free again: the next upload may start under the sweeps
                self._faces_read = torch.cuda.Event()
                self._faces_read.record()
            m = self._mcur
            with H.nvtx("lbmff:magnetic_sweeps"), self._timed("sweeps"):
                if self.magnetic_path == "converged":
                    self.cg.solve(self._rhs, self.psi)
                elif self.magnetic_path == "faithful":
                    H.launch("lbmff_magnetic_sweeps_" + suf, dev, H.ptr(self._h[m]), H.ptr(self._h[1 - m]),
                             H.ptr(self.mflags), H.ptr(self._rhs), H.ptr(self.psi), self.n_sweeps,
                             float(self._tau), self.dx, self.dt, *sz)
                else:
                    H.launch(f"lbmff_magnetic_sweeps_scalar_{d}d", dev, H.ptr(self._a[m]), H.ptr(self._a[1 - m]),
                             H.ptr(self.mflags), H.ptr(self._rhs), H.ptr(self.psi), self.n_sweeps, self.dx,
                             self.dt, *sz)
            if self.n_sweeps & 1 and self.magnetic_path != "converged":
                self._mcur = 1 - m
            chi, psi, mfl, hext, hu = self._chi, self.psi, self.mflags, self.H_ext, self._H_uniform
        args = [H.ptr(self._f[cur]), H.ptr(self._g[cur]), H.ptr(self._f[nxt]), H.ptr(self._g[nxt]),
                H.ptr(self.rho), H.ptr(self.flags), H.ptr(mfl), H.ptr(self._ws), H.ptr(chi), H.ptr(psi),
                H.ptr(hext), ctypes.cast(hu, ctypes.c_void_p) if hu is not None else ctypes.c_void_p(0),
                self.params]
        if d == 2:
            args.append(self._kbc_mode)
        with H.nvtx("lbmff:collide_stream"), self._timed("collide_stream"):
            H.launch("lbmff_hcz_collide_stream_" + suf, dev, *args, H.ptr(self.vel), H.ptr(self.pressure),
                     H.ptr(self.force), *sz)
        self._cur = nxt
        self.steps_done += 1

    def step(self, n: int = 1, graph: bool = False):
        """Advance n outer steps.  Asynchronous on torch's current stream.

        ``graph=True`` replays a CUDA graph of TWO steps (an A/B buffer round trip) captured on first use: the
        ~45 launches of a step then cost one graph launch per pair.  Measured at the reference's default 130^3
        demo size the step is already GPU-bound (2.2 ms of under-filled kernels), so the gain is ~5 %; it grows as
        the grid shrinks.  Bit-identical to the eager path; any odd remainder runs eagerly."""
        n = int(n)
        if graph and n >= 2:
            g = self._step_graph()
            if g is not None:                 # None: the capture failed (see _step_graph); eager launches below
                for _ in range(n // 2):
                    g.replay()
                self.steps_done += 2 * (n // 2)
                n -= 2 * (n // 2)
        for _ in range(n):
            self._step_once()
        return self

    def _step_graph(self):
        """CUDA graph of two consecutive steps starting from the CURRENT buffer parity (one graph per parity)."""
        if self.magnetic_path == "converged":
            raise RuntimeError("graph=True is not available with magnetic_path='converged' "
                               "(the CG solve launches its own conditional graph)")
        if self._timing is not None:
            raise RuntimeError("graph=True cannot be combined with enable_kernel_timing()")
        graphs = self.__dict__.setdefault("_graphs", {})
        key = (self._cur, getattr(self, "_mcur", 0))
        if key not in graphs:
            import gc
            saved = (self._cur, getattr(self, "_mcur", 0), self.steps_done)
            g = None
            # A capture is invalidated by any capture-unsafe CUDA call made while it is open -- in "global" mode by ANY
            # thread, and in this thread by whatever a garbage-collection pass happens to finalise (a pinned buffer, an
            # event of an earlier solver): seen once in four full test-suite runs.  So: thread-local capture mode, the
            # collector off for the duration, one retry, and if the capture still fails the caller steps eagerly -- a
            # graph is an optimisation, never a reason for step() to raise.
            for _attempt in range(2):
                torch.cuda.synchronize(self.device)
                gc_on = gc.isenabled()
                gc.collect()
                gc.disable()
                try:
                    cand = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(cand, capture_error_mode="thread_local"):   # records only: nothing executes
                        self._step_once()
                        self._step_once()
                    g = cand
                except Exception as e:                     # noqa: BLE001 -- any capture failure means "no graph"
                    import warnings
                    warnings.warn(f"CUDA graph capture of the step failed ({type(e).__name__}); "
                                  f"{'retrying' if _attempt == 0 else 'stepping eagerly'}")
                    g = None
                finally:
                    if gc_on:
                        gc.enable()
                    self._cur, self.steps_done = saved[0], saved[2]
                    if self.magnetic:
                        self._mcur = saved[1]
                if g is not None:
                    break
                torch.cuda.synchronize(self.device)
            graphs[key] = g
        return graphs[key]

    def set_H_ext_faces_from_host(self, faces_host: Sequence[torch.Tensor]):
        """Upload the staggered applied field (the H_ext_mac argument of get_H_int) from pinned host memory,
        asynchronously on the current stream.  Only the FACES are replaced: they feed the Poisson source
        (get_H_int).  The cell-centred field of the Kelvin force, H2 = |H_ext + H_int|^2
        (demo_3d_LBM_Rosensweig_instability.py:174-178), stays the ``H_ext`` given at construction -- exactly as in
        the demos, where ``H_ext_const_real`` and ``H_ext_mac`` are two separate tensors the caller keeps
        consistent.  For a time-varying applied field also update ``solver.H_ext`` (field mode) yourself."""
        if not self.magnetic:
            raise RuntimeError("solver was built without a magnetic subsystem")
        for dst, src in zip(self.H_ext_mac, faces_host):
            if tuple(dst.shape) != tuple(src.shape):
                raise RuntimeError(f"face array of shape {tuple(src.shape)} given, {tuple(dst.shape)} expected")
            dst.copy_(src, non_blocking=True)

    def step_host(self, faces_host: Optional[Sequence[torch.Tensor]], phi_host: Optional[torch.Tensor]):
        """One outer step through host buffers: H2D of this step's applied-field faces (pinned), the step, D2H of
        the order parameter phi the demos consume (pinned).  Asynchronous.  The copies ride a second stream: the
        upload waits only for the previous step's last reader of the faces and so runs under that step's
        magnetic sweeps (the host enqueues steps faster than the device executes them), the download of phi starts
        as soon as rho is final and overlaps the sweeps and the collision of its own step."""
        if faces_host is not None and not self.magnetic:
            raise RuntimeError("solver was built without a magnetic subsystem")
        cur = torch.cuda.current_stream(self.device)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)      # downloads
            self._up_stream = torch.cuda.Stream(device=self.device)        # uploads (other copy engine)
            self._faces_read = None
        cs, us = self._copy_stream, self._up_stream
        up = None
        if faces_host is not None:
            # WAR on the face arrays: wait only for the previous step's source kernel (their last reader), not for
            # the whole previous step -- the upload then runs under the previous step's magnetic sweeps
            if self._faces_read is not None:
                us.wait_event(self._faces_read)
            else:
                us.wait_stream(cur)
            with torch.cuda.stream(us):
                for dst, src in zip(self.H_ext_mac, faces_host):
                    dst.copy_(src, non_blocking=True)
                up = torch.cuda.Event()
                up.record(us)
        phi_done = [None]

        def download():
            if phi_host is None:
                return
            ev = torch.cuda.Event()
            ev.record(cur)
            with torch.cuda.stream(cs):
                cs.wait_event(ev)
                phi = self.phi
                phi_done[0] = torch.cuda.Event()
                phi_done[0].record(cs)       # rho may be overwritten once phi has been formed
                phi_host.copy_(phi, non_blocking=True)
                phi.record_stream(cs)

        self._step_once(before_source=(lambda: cur.wait_event(up)) if up is not None else None,
                        after_rho_final=download)
        if phi_done[0] is not None:
            cur.wait_event(phi_done[0])
        self._host_copy_pending = True
        return self

    def wait_host_copies(self):
        """Make the current stream wait for the host copies queued by step_host."""
        if getattr(self, "_copy_stream", None) is not None:
            torch.cuda.current_stream(self.device).wait_stream(self._copy_stream)
            torch.cuda.current_stream(self.device).wait_stream(self._up_stream)

    @property
    def launches_per_step(self) -> int:
        wet = 5 if self.dim == 3 else 3
        shell = (6 if self.dim == 3 else 2) if self.magnetic else 0   # rim (+ two z-shell planes) for psi and chi
        if self.magnetic_path == "converged":
            mag = 1 + self.cg.launches_per_solve
        else:
            mag = 1 + self.n_sweeps if self.magnetic else 0
        return 1 + mag + wet + 1 + shell + 1

    def total_mass(self) -> float:
        """sum over ALL cells (walls included) of sum_q f: conserved to round-off."""
        return float(self.f.double().sum().item())

    # ---- checkpoint / resume (absent in the reference, which only writes frames) ----------
    def state_dict(self) -> dict:
        """Everything a later ``load_state_dict`` needs to continue BIT-IDENTICALLY: the current populations, the
        fields one step reads from the previous one (rho, vel, pressure, psi: the warm start) and the magnetic
        lattice state of the active path.  Tensors are clones on the solver's device; ``torch.save`` it as is."""
        sd = dict(version=1, dim=self.dim, res=tuple(self.res), batch=self.B, steps_done=self.steps_done,
                  magnetic_path=self.magnetic_path, f=self.f.clone(), g=self.g.clone(), rho=self.rho.clone(),
                  vel=self.vel.clone(), pressure=self.pressure.clone())
        if self.magnetic:
            sd["psi"] = self.psi.clone()
            if self.magnetic_path == "faithful":
                sd["h"] = self._h[self._mcur].clone()
            elif self.magnetic_path == "scalar":
                sd["a_cur"], sd["a_prev"] = self._a[self._mcur].clone(), self._a[1 - self._mcur].clone()
        return sd

    def load_state_dict(self, sd: dict):
        if sd.get("version") != 1 or sd["dim"] != self.dim or tuple(sd["res"]) != tuple(self.res) or \
                sd["batch"] != self.B:
            raise RuntimeError("checkpoint does not match this solver (version / dim / res / batch)")
        if sd["magnetic_path"] != self.magnetic_path:
            raise RuntimeError(f"checkpoint was written by magnetic_path={sd['magnetic_path']!r}, "
                               f"this solver runs {self.magnetic_path!r}")
        self._cur = 0
        self._f[0].copy_(sd["f"]); self._g[0].copy_(sd["g"])
        self.rho.copy_(sd["rho"]); self.vel.copy_(sd["vel"]); self.pressure.copy_(sd["pressure"])
        if self.magnetic:
            self.psi.copy_(sd["psi"])
            self._mcur = 0
            if self.magnetic_path == "faithful":
                self._h[0].copy_(sd["h"])
            elif self.magnetic_path == "scalar":
                self._a[0].copy_(sd["a_cur"]); self._a[1].copy_(sd["a_prev"])
        self.steps_done = int(sd["steps_done"])
        return self


class SinglePhaseSolver:
    """Single-phase BGK / KBC flow, D2Q9 or D3Q19: demo_2d_LBM.py / demo_2d_LBM_KBC.py loop in one kernel."""

    def __init__(self, flags: torch.Tensor, *, tau: float, KBC_type: Optional[int] = None, dx=1.0, dt=1.0,
                 force: Optional[torch.Tensor] = None):
        self.flags = H.need_cuda(flags, "flags", torch.uint8)
        self.device = self.flags.device
        self.dim, self.B, _, self.D, self.Hh, self.W = H.dims(self.flags)
        self.res = _res(self.flags)
        self.Q = 9 if self.dim == 2 else 19
        self.tau, self.dx, self.dt = float(tau), float(dx), float(dt)
        if KBC_type is not None and self.dim == 3:
            raise RuntimeError("the reference has no 3-D KBC/MRT collision (simulation_runner.py:107-108)")
        self._kbc = -1 if KBC_type is None else (1 if KBCType.is_KBC(KBC_type) else 0)
        self.force = H.need_cuda(force, "force") if force is not None else None
        opts = dict(dtype=torch.float32, device=self.device)
        self.rho = torch.zeros((self.B, 1, *self.res), **opts)
        self.vel = torch.zeros((self.B, self.dim, *self.res), **opts)
        self._f = [torch.zeros((self.B, self.Q, *self.res), **opts) for _ in range(2)]
        self._cur = 0
        self.steps_done = 0

    @property
    def f(self):
        return self._f[self._cur]

    def initialize(self, rho: torch.Tensor, vel: torch.Tensor):
        """f = feq(rho, vel, force = 0) with self.tau (demo_2d_LBM.py:99)."""
        self.rho.copy_(H.need_cuda(rho, "rho"))
        self.vel.copy_(H.need_cuda(vel, "vel"))
        col = (LBMCollision2d if self.dim == 2 else LBMCollision3d)(tau=self.tau, device=self.device)
        self._cur = 0
        self._f[0].copy_(col.get_feq_(dx=self.dx, dt=self.dt, rho=self.rho, vel=self.vel,
                                      force=torch.zeros_like(self.vel)))
        self.steps_done = 0
        return self

    def step(self, n: int = 1):
        sz = H.sizes(self.dim, self.B, self.D, self.Hh, self.W)
        for _ in range(int(n)):
            cur, nxt = self._cur, 1 - self._cur
            args = [H.ptr(self._f[cur]), H.ptr(self._f[nxt]), H.ptr(self.flags), H.ptr(self.force), H.ptr(self.rho),
                    H.ptr(self.vel), self.tau]
            if self.dim == 2:
                args.append(self._kbc)
            H.launch("lbmff_stream_collide_" + H.suffix(self.dim), self.device, *args, self.dx, self.dt, *sz)
            self._cur = nxt
            self.steps_done += 1
        return self


class ShanChenSolver:
    """Shan-Chen pseudopotential two-phase flow on D2Q9: the loop of demo_2d_LBM_multiphase_SC.py:113-131
    (propagation, macro_compute, rebounce_obstacle, LBMCollisionSC2d.collision) in two kernels per step --
    lbmff_sc_moments_d2q9 (rho, density, psi) and lbmff_sc_collide_stream_d2q9 (bounce, interaction force, BGK).
    The op-by-op class LBMCollisionSC2d stays the drop-in API."""

    def __init__(self, flags: torch.Tensor, *, tau: float = 1.0, density_liquid=0.265, density_gas=0.038,
                 rho_liquid=0.265, rho_gas=0.038, dx=1.0, dt=1.0, keep_force: bool = False):
        self.flags = H.need_cuda(flags, "flags", torch.uint8)
        self.device = self.flags.device
        self.dim, self.B, _, self.D, self.Hh, self.W = H.dims(self.flags)
        if self.dim != 2:
            raise RuntimeError("the reference's Shan-Chen model is 2-D (simulation_runner.py:107-108); "
                               "extensions.LBMCollisionSC3d is the 3-D extension")
        self.res = _res(self.flags)
        self.dx, self.dt = float(dx), float(dt)
        self._tau, self._density_liquid, self._density_gas = float(tau), density_liquid, density_gas
        self._rho_liquid, self._rho_gas = rho_liquid, rho_gas
        self.params = H.fluid_params(self, self.dx, self.dt)
        opts = dict(dtype=torch.float32, device=self.device)
        sc = lambda c: torch.zeros((self.B, c, *self.res), **opts)  # noqa: E731
        self.rho, self.density, self.psi, self.vel = sc(1), sc(1), sc(1), sc(2)
        self.force = sc(2) if keep_force else None
        self._f = [sc(9), sc(9)]
        self._cur = 0
        self.steps_done = 0

    @property
    def f(self):
        return self._f[self._cur]

    def initialize(self, rho: torch.Tensor, density: Optional[torch.Tensor] = None,
                   vel: Optional[torch.Tensor] = None):
        """f = feq(density, vel, force = 0) as the demo does (demo_2d_LBM_multiphase_SC.py:111)."""
        self.rho.copy_(H.need_cuda(rho, "rho"))
        self.density.copy_(self.rho if density is None else H.need_cuda(density, "density"))
        if vel is None:
            self.vel.zero_()
        else:
            self.vel.copy_(H.need_cuda(vel, "vel"))
        col = LBMCollision2d(tau=self._tau, device=self.device)
        self._cur = 0
        self._f[0].copy_(col.get_feq_(dx=self.dx, dt=self.dt, rho=self.density, vel=self.vel,
                                      force=torch.zeros_like(self.vel)))
        self.steps_done = 0
        return self

    def step(self, n: int = 1):
        sz = (self.B, self.Hh, self.W)
        for _ in range(int(n)):
            cur, nxt = self._cur, 1 - self._cur
            H.launch("lbmff_sc_moments_d2q9", self.device, H.ptr(self._f[cur]), H.ptr(self.flags), H.ptr(self.rho),
                     H.ptr(self.density), H.ptr(self.psi), self.params, *sz)
            H.launch("lbmff_sc_collide_stream_d2q9", self.device, H.ptr(self._f[cur]), H.ptr(self._f[nxt]),
                     H.ptr(self.flags), H.ptr(self.psi), H.ptr(self.rho), H.ptr(self.vel), H.ptr(self.force),
                     self.params, *sz)
            self._cur = nxt
            self.steps_done += 1
        return self
