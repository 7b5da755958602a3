"""Converged magnetic potential: host side of ``lbmff_poisson_cg_{2d,3d}`` (include/lbmff.h, last section).

Opt-in alternative to the reference's fixed 30 relaxation sweeps per outer step
(demo_3d_LBM_Rosensweig_instability.py:163-173): conjugate gradients on the fixed-point equation of those sweeps
(tau == 1), run entirely on the device.  Not part of the drop-in surface -- the reference has no such solver.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _host as H
from . import _lib


class PoissonCG:
    """Workspace + launcher.  ``solve`` is asynchronous on torch's current stream; ``info()`` synchronises."""

    def __init__(self, mflags: torch.Tensor, *, dx=1.0, dt=1.0, tol: float = 1e-4, max_iters: int = 200,
                 keep_history: bool = False, precondition: Optional[str] = None):
        self.mflags = H.need_cuda(mflags, "magnetic_flags", torch.uint8)
        self.device = self.mflags.device
        self.dim, self.B, _, self.D, self.Hh, self.W = H.dims(self.mflags)
        self.dx, self.dt = float(dx), float(dt)
        self.tol, self.max_iters = float(tol), int(max_iters)
        n = _lib.load().lbmff_poisson_cg_scratch_doubles(self.D, self.Hh, self.W)
        if n <= 0:
            raise RuntimeError("lbmff_poisson_cg_scratch_doubles: bad extents")
        N = self.D * self.Hh * self.W
        if precondition not in (None, "mg"):
            raise ValueError("precondition must be None or 'mg'")
        self.precondition = precondition
        self._mg = None
        self._hierarchy_built = False
        if precondition == "mg":   # geometric-multigrid V(2,2) cycle per iteration (csrc/poisson_mg.cuh)
            m = _lib.load().lbmff_poisson_mg_workspace_floats(self.D, self.Hh, self.W)
            self._mg = torch.empty((self.B, max(1, m)), dtype=torch.float32, device=self.device)
        self._work = torch.empty((self.B, 7 if precondition else 5, N), dtype=torch.float32, device=self.device)
        self._scratch = torch.zeros((self.B, n), dtype=torch.float64, device=self.device)
        self._info = torch.zeros((self.B, 5), dtype=torch.float64, device=self.device)
        self._hist = (torch.zeros((self.B, self.max_iters + 1), dtype=torch.float32, device=self.device)
                      if keep_history else None)

    @property
    def launches_per_solve(self) -> int:
        return self.B * (4 + 1)   # rhs, start, two reductions + ONE graph launch (the iterations are a device-side WHILE loop)

    def solve(self, rhs: torch.Tensor, psi: torch.Tensor, max_iters: Optional[int] = None):
        """psi (warm start) is updated in place to the solution of the sweeps' fixed-point equation."""
        rhs, psi_c = H.need_cuda(rhs, "rhs"), H.need_cuda(psi, "psi")
        if psi_c.data_ptr() != psi.data_ptr():
            raise RuntimeError("psi must be contiguous (it is updated in place)")
        if tuple(rhs.shape) != tuple(self.mflags.shape) or tuple(psi.shape) != tuple(self.mflags.shape):
            raise RuntimeError("rhs / psi must have the shape of magnetic_flags")
        it = self.max_iters if max_iters is None else int(max_iters)
        if it > self.max_iters:
            raise RuntimeError("max_iters above the one the workspace was sized for")
        self._last_iters = it
        if self.precondition == "mg":
            H.launch(f"lbmff_poisson_pcg_{self.dim}d", self.device, H.ptr(rhs), H.ptr(self.mflags), H.ptr(psi),
                     H.ptr(self._work), H.ptr(self._mg), H.ptr(self._scratch), it, self.tol, 1.0, self.dx, self.dt,
                     H.ptr(self._hist), H.ptr(self._info), int(self._hierarchy_built),
                     *H.sizes(self.dim, self.B, self.D, self.Hh, self.W))
            self._hierarchy_built = True     # mflags and the workspace belong to this object: build once
            return psi
        H.launch(f"lbmff_poisson_cg_{self.dim}d", self.device, H.ptr(rhs), H.ptr(self.mflags), H.ptr(psi),
                 H.ptr(self._work), H.ptr(self._scratch), it, self.tol, 1.0, self.dx, self.dt,
                 H.ptr(self._hist), H.ptr(self._info), *H.sizes(self.dim, self.B, self.D, self.Hh, self.W))
        return psi

    def invalidate_hierarchy(self):
        """Call after changing ``mflags`` in place: the multigrid hierarchy (built by the first solve from the flags)
        is rebuilt by the next one."""
        self._hierarchy_built = False

    def info(self):
        """Per batch item: dict(iters, converged, residual, rhs_norm, drift) of the last solve (synchronises).
        ``drift`` is the constant the reference's sweeps would add to psi per sweep in their stationary state."""
        rows = self._info.cpu().tolist()
        return [dict(iters=int(r[0]), converged=bool(r[1]), residual=r[2] ** 0.5, rhs_norm=r[3] ** 0.5, drift=r[4])
                for r in rows]

    def history(self):
        """|r| after each iteration of the last solve, [B][iters + 1] (needs keep_history=True)."""
        if self._hist is None:
            raise RuntimeError("constructed with keep_history=False")
        it = getattr(self, "_last_iters", self.max_iters)   # the C entry strides the history by its own max_iters + 1
        h = self._hist.flatten()[: self.B * (it + 1)].view(self.B, it + 1).cpu()
        return [h[b, : i["iters"] + 1].tolist() for b, i in enumerate(self.info())]
