"""Streaming and bounce-back (reference: src/LBM/LBM_propagation/LBM_propagation_{2,3}d.py)."""
from __future__ import annotations

import torch

from . import _host as H


class _Propagation:
    type_flags = ["x", "X", "y", "Y", "z", "Z"]
    _dim = 0

    def __init__(self, *args, **kwargs):
        pass

    def _check(self, f):
        f = H.need_cuda(f, "f")
        dim, B, Q, D, Hh, W = H.dims(f)
        if dim != self._dim or Q != (9 if dim == 2 else 19):
            raise RuntimeError(f"expected a [B, {9 if self._dim == 2 else 19}, ...] {self._dim}-D lattice tensor")
        return f, B, D, Hh, W

    def propagation(self, f: torch.Tensor) -> torch.Tensor:
        """Periodic pull streaming; returns a new tensor (LBM_propagation_3d.py:18-111)."""
        f, B, D, Hh, W = self._check(f)
        out = torch.empty_like(f)
        H.launch("lbmff_propagation_" + H.suffix(self._dim), f.device, H.ptr(f), H.ptr(out),
                 *H.sizes(self._dim, B, D, Hh, W))
        return out

    def rebounce_obstacle(self, f: torch.Tensor, flags: torch.Tensor) -> torch.Tensor:
        """Swap opposite populations on OBSTACLE cells (LBM_propagation_3d.py:113-142)."""
        f, B, D, Hh, W = self._check(f)
        flags = H.need_cuda(flags, "flags", torch.uint8)
        out = torch.empty_like(f)
        H.launch("lbmff_rebounce_obstacle_" + H.suffix(self._dim), f.device, H.ptr(f), H.ptr(flags), H.ptr(out),
                 *H.sizes(self._dim, B, D, Hh, W))
        return out


class LBMPropagation2d(_Propagation):
    rank = 2
    _dim = 2


class LBMPropagation3d(_Propagation):
    rank = 2   # sic: the reference's 3-D class also says rank = 2 (LBM_propagation_3d.py:9)
    _dim = 3
