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
    """The reference's frame writer (src/LBM/utils/io.py:50-84) with every stage on the device: the interior of phi
    padded by phi.max(), the zero level of -phi * 2 / (max(res) dx) extracted by the CUDA marching cubes
    (isosurface.py; the reference goes through the host and PyMCubes), vertices centred and scaled into [-1, 1],
    a uniform 0.68 grey surface texture, ``renderer.forward`` (render.SoftRenderer: tiled CUDA soft rasteriser), and
    the RGB image written as 8 bit.  Returns [verts, faces] like the reference."""
    import cv2
    import torch.nn.functional as F

    from .isosurface import marching_cubes

    phi = F.pad(phi[..., 1:-1, 1:-1, 1:-1], pad=(1, 1, 1, 1, 1, 1), mode="constant", value=phi.max().item())
    device = phi.device
    verts, faces = marching_cubes((-phi * (2.0 / max(res) / dx))[0, 0].contiguous(), 0.0)
    verts = (verts - torch.tensor([float(r) for r in res], device=device) / 2.0) * (2.0 / max(res))
    textures = 0.68 * torch.ones(1, faces.shape[-2], 2, 3, dtype=torch.float32, device=device)
    image = renderer.forward(verts.unsqueeze(0), faces.unsqueeze(0), textures)
    rgb = (image[0, :3].permute(1, 2, 0).detach().clamp(0, 1).cpu().numpy() * 255).astype(np.uint8)
    cv2.imwrite(filename, rgb[..., ::-1])            # imageio.imwrite takes RGB, cv2 BGR
    return [verts, faces]


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
