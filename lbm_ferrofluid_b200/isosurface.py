"""Iso-surface extraction on the device: host side of ``lbmff_iso_surface_*`` (include/lbmff.h).

``marching_cubes(volume, isovalue)`` has the call shape of ``mcubes.marching_cubes`` as the reference uses it
(src/LBM/utils/io.py:60-63) but takes a CUDA tensor and returns CUDA tensors: no host round trip of the field.
"""
from __future__ import annotations

import ctypes

import numpy as np
import torch

from . import _host as H
from . import _lib


def case_table():
    """(ntri [256], edges [256, 8, 3]) of the derived marching-cubes table (host only, no GPU needed)."""
    ntri = (ctypes.c_int * 256)()
    edges = (ctypes.c_int * (256 * 24))()
    _lib.call("lbmff_mc_case_table", ctypes.cast(ntri, ctypes.c_void_p), ctypes.cast(edges, ctypes.c_void_p))
    return np.array(ntri, dtype=np.int32), np.array(edges, dtype=np.int32).reshape(256, 8, 3)


def triangle_soup(volume: torch.Tensor, isovalue: float = 0.0, flip: bool = False) -> torch.Tensor:
    """[F, 3, 3] fp32 CUDA tensor: the triangles of {volume == isovalue} in array-index coordinates (axis 0, 1, 2).
    Counter-clockwise normals point towards DECREASING values -- out of the liquid for the -phi volume the frame
    writer passes, so that its upper surface is lit (``flip`` reverses).  One 4-byte device->host read
    (the triangle count) separates the two passes."""
    vol = H.need_cuda(volume, "volume")
    if vol.dim() != 3:
        raise RuntimeError("volume must be a [D, H, W] tensor")
    D, Hh, W = vol.shape
    nb = _lib.load().lbmff_iso_surface_blocks(D, Hh, W)
    if nb <= 0:
        raise RuntimeError("marching_cubes: every extent must be >= 2")
    dev = vol.device
    offsets = torch.empty(nb, dtype=torch.int32, device=dev)
    total = torch.zeros(1, dtype=torch.int32, device=dev)
    H.launch("lbmff_iso_surface_count_3d", dev, H.ptr(vol), float(isovalue), H.ptr(offsets), H.ptr(total), D, Hh, W)
    F = int(total.item())
    tri = torch.empty((F, 3, 3), dtype=torch.float32, device=dev)
    H.launch("lbmff_iso_surface_emit_3d", dev, H.ptr(vol), float(isovalue), int(bool(flip)), H.ptr(offsets),
             H.ptr(tri if F else None), F, D, Hh, W)
    return tri


def marching_cubes(volume: torch.Tensor, isovalue: float = 0.0, flip: bool = False):
    """(vertices [3F, 3] fp32, faces [F, 3] int32), both on the device: the soup with trivial indexing
    (vertices are not welded; the soft rasteriser and the OBJ writer only ever gather vertices[faces])."""
    tri = triangle_soup(volume, isovalue, flip)
    F = tri.shape[0]
    return tri.reshape(3 * F, 3), torch.arange(3 * F, dtype=torch.int32, device=tri.device).reshape(F, 3)
