"""Small output helpers with the names the reference's demos import from ``src.LBM.utils``
(src/LBM/utils/io.py).  Viz/export is outside the solver hot path: these are
thin, lazily-importing conveniences so the demo scripts reach their time loops in this image, which has
no matplotlib / PyMCubes / imageio."""
from __future__ import annotations

import os

import numpy as np
import torch

__all__ = ["mkdir", "save_img", "export_asset", "save_rendered_image", "dump_2d_plt_file_single"]


def mkdir(path: str):
    os.makedirs(path, exist_ok=True)


def save_img(tensor_input: torch.Tensor, filename: str):
    """Normalise a [B,C,H,W] (or z-averaged [B,C,D,H,W]) field to 8 bit and write it, y up."""
    import cv2

    if tensor_input.dim() == 5:
        tensor_input = tensor_input.mean(dim=2)
    elif tensor_input.dim() != 4:
        raise RuntimeError("To save an image, the tensor shape should be 4 or 5")
    img = tensor_input[0].detach().float().cpu()
    lo, hi = img.min(), img.max()
    img = (img - lo) / (hi - lo + 1e-12)
    arr = (img.permute(1, 2, 0).numpy() * 255.0)
    if arr.shape[2] == 2:
        arr = np.concatenate([arr, np.zeros_like(arr[..., :1])], axis=2)
    cv2.imwrite(filename, np.flipud(arr))


def export_asset(save_path: str, vertices: torch.Tensor, faces: torch.Tensor):
    """Write a triangle mesh as Wavefront OBJ (1-based face indices)."""
    v = vertices.reshape(-1, 3).detach().cpu().numpy()
    f = faces.reshape(-1, 3).to(torch.int64).cpu().numpy()
    if f.size and f.min() == 0:
        f = f + 1
    with open(save_path, "w") as fo:
        fo.write("# OBJ file\n")
        for row in v:
            fo.write("v {} {} {}\n".format(*row))
        for row in f:
            fo.write("f {} {} {}\n".format(*row))


def save_rendered_image(renderer, phi, filename, res, dx):
    """Headless stand-in for the reference's marching-cubes + soft-rasteriser frame (io.py:50-84): writes a
    y-up projection of the liquid column height.  Returns empty (verts, faces)."""
    import cv2

    liquid = (phi[0, 0, 1:-1, 1:-1, 1:-1] < 0).float().sum(dim=1)   # [D-2, W-2] column heights
    img = (liquid / max(1.0, float(phi.shape[-2] - 2))).clamp(0, 1).cpu().numpy()
    cv2.imwrite(filename, (img * 255.0).astype(np.uint8))
    dev = phi.device
    return [torch.zeros((0, 3), device=dev), torch.zeros((0, 3), dtype=torch.int32, device=dev)]


def dump_2d_plt_file_single(filename, np_density, np_uv, B):
    """Tecplot ASCII dump of a 2-D scalar + vector field."""
    H, W = np_density.shape[-2:]
    with open(filename, "w") as fo:
        fo.write('TITLE ="Magnetic internal force"\n')
        fo.write('VARIABLES = "X", "Y", "density", "U", "V"\n')
        fo.write(f"ZONE I={W}, J={H}, F=POINT\n")
        d = np.asarray(np_density).reshape(-1, H, W)[0]
        uv = np.asarray(np_uv).reshape(-1, 2, H, W)[0]
        for j in range(H):
            for i in range(W):
                fo.write(f"{i} {j} {d[j, i]} {uv[0, j, i]} {uv[1, j, i]}\n")
