"""Cell-centred -> staggered (face) averages (reference: src/LBM/utils/grid.py:7-64)."""
from __future__ import annotations

from typing import List

import torch

from . import _host as H


def _staggered(input: torch.Tensor, axis: int, mode: str) -> torch.Tensor:
    if input.dim() not in (4, 5) or (axis == 2 and input.dim() != 5):
        raise RuntimeError("A grid has to be 2D(3D) [B, C, (D), H, W] to be staggered")
    if mode != "replicate":
        raise NotImplementedError("only mode='replicate' (the reference's default and only use) is implemented")
    x = H.need_cuda(input, "input")
    dim, B, C, D, Hh, W = H.dims(x)
    shape = list(x.shape)
    shape[x.dim() - 1 - axis] += 1
    out = torch.empty(shape, dtype=x.dtype, device=x.device)
    H.launch("lbmff_get_staggered", x.device, H.ptr(x), H.ptr(out), axis, B * C, D, Hh, W)
    return out


def get_staggered_x(input: torch.Tensor, mode: str = "replicate") -> torch.Tensor:
    return _staggered(input, 0, mode)


def get_staggered_y(input: torch.Tensor, mode: str = "replicate") -> torch.Tensor:
    return _staggered(input, 1, mode)


def get_staggered_z(input: torch.Tensor, mode: str = "replicate") -> torch.Tensor:
    if input.dim() != 5:
        raise RuntimeError("A grid has to be 3D [B, C, D, H, W] to be staggered")
    return _staggered(input, 2, mode)


def get_staggered(input: torch.Tensor, mode: str = "replicate") -> List[torch.Tensor]:
    ncomp = input.shape[1]
    if ncomp < 2 or ncomp > 3:
        raise RuntimeError("Only 2D or 3D scene supported")
    return [_staggered(input[:, a:a + 1], a, mode) for a in range(ncomp)]
