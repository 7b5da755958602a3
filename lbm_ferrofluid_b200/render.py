"""Host mirror of the reference's ``renderutils`` for the frames the 3-D demos write.

``SoftRenderer`` takes the reference's constructor arguments and exposes its methods with the same names and
argument meaning (renderutils/softrenderer.py:24-265 constructor, :266-330 render, :332-365 lighting, :367-428 camera,
:430-467 rasterize, :469-610 look_at / look / perspective_distortion / vertices_to_faces); ``soft_rasterize`` is
renderutils/soft_rasterize.py:185-233.  The per-vertex camera / lighting algebra is a handful of torch device
expressions (a few flops per vertex -- plumbing); the per-pixel work is the hand-written tiled CUDA rasteriser behind
``lbmff_soft_rasterize_*`` (csrc/soft_raster.cu).  Forward only: the demos render frames, nothing differentiates
through them.  CUDA tensors only -- the reference says the same of itself ("SoftRenderer works only in GPU mode").
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as Fn

from . import _host as H
from . import _lib

_DIST = {"hard": 0, "barycentric": 1, "euclidean": 2}
_RGB = {"hard": 0, "softmax": 1}
_ALPHA = {"hard": 0, "sum": 1, "prod": 2}
_SAMPLE = {"surface": 0, "vertex": 1}


def soft_rasterize(face_vertices, textures, image_size=256, background_color=None, near=1, far=100, fill_back=True,
                   eps=1e-3, sigma_val=1e-5, dist_func="euclidean", dist_eps=1e-4, gamma_val=1e-4,
                   aggr_func_rgb="softmax", aggr_func_alpha="prod", texture_type="surface", return_aggrs=False):
    """[B, 4, image_size, image_size] RGBA, as SoftRasterizeFunction.forward (soft_rasterize.py:26-110)."""
    fv = H.need_cuda(face_vertices, "face_vertices")
    tex = H.need_cuda(textures, "textures")
    if fv.dim() != 4 or fv.shape[2:] != (3, 3):
        raise RuntimeError("face_vertices must be [B, F, 3, 3]")
    B, F = fv.shape[:2]
    if tex.dim() != 4 or tex.shape[0] != B or tex.shape[1] != F or tex.shape[3] != 3:
        raise RuntimeError("textures must be [B, F, texture_size, 3]")
    for name, table, key in (("dist_func", _DIST, dist_func), ("aggr_func_rgb", _RGB, aggr_func_rgb),
                             ("aggr_func_alpha", _ALPHA, aggr_func_alpha), ("texture_type", _SAMPLE, texture_type)):
        if key not in table:
            raise KeyError(f"{name}={key!r}: expected one of {sorted(table)}")
    if background_color is None:
        background_color = [0, 0, 0]
    dev = fv.device
    size = int(image_size)
    colors = torch.ones((B, 4, size, size), dtype=torch.float32, device=dev)
    for k in range(3):
        colors[:, k] *= float(background_color[k])
    aggrs = torch.zeros((B, 2, size, size), dtype=torch.float32, device=dev) if return_aggrs else None
    if F == 0:
        colors[:, 3] = 0.0                                         # no face: every aggregation leaves alpha 0
        return (colors, aggrs) if return_aggrs else colors
    log_eps = math.log(1.0 / dist_eps - 1.0)                       # soft_rasterize.py:54
    n_off = _lib.load().lbmff_soft_rasterize_offsets(B, size)
    if n_off <= 0:
        raise RuntimeError("soft_rasterize: image too large")
    info = torch.empty((B, F, 24), dtype=torch.float32, device=dev)
    box = torch.empty((B * F + B * ((F + 31) // 32), 4), dtype=torch.float32, device=dev)   # face + group boxes
    offsets = torch.empty(n_off, dtype=torch.int32, device=dev)
    H.launch("lbmff_soft_rasterize_bin", dev, H.ptr(fv), H.ptr(info), H.ptr(box), H.ptr(offsets), B, F, size,
             float(sigma_val), log_eps)
    total = int(offsets[-1].item())                                # one 4-byte read: the list length
    faces_of_tiles = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    H.launch("lbmff_soft_rasterize_forward", dev, H.ptr(fv), H.ptr(tex), H.ptr(info), H.ptr(box), H.ptr(offsets),
             H.ptr(faces_of_tiles), H.ptr(colors), H.ptr(aggrs), B, F, size, int(tex.shape[2]), float(near),
             float(far), float(eps), float(sigma_val), _DIST[dist_func], log_eps, float(gamma_val),
             _RGB[aggr_func_rgb], _ALPHA[aggr_func_alpha], _SAMPLE[texture_type], int(bool(fill_back)))
    return (colors, aggrs) if return_aggrs else colors


def _normalize(v, eps):
    return Fn.normalize(v, eps=eps)


class SoftRenderer:
    """renderutils.SoftRenderer with the reference's constructor arguments and defaults (softrenderer.py:48-79)."""

    def __init__(self, image_size: int = 256, anti_aliasing: bool = True, bg_color: torch.Tensor = None,
                 fill_back: bool = True, camera_mode: str = "look_at", K=None, rmat=None, tvec=None,
                 perspective_distort: bool = True, sigma_val: float = 1e-5, dist_func: str = "euclidean",
                 dist_eps: float = 1e-4, gamma_val: float = 1e-4, aggr_func_rgb: str = "softmax",
                 aggr_func_alpha: str = "prod", texture_type: str = "surface", viewing_angle: float = 30.0,
                 viewing_scale: float = 1.0, eye: torch.Tensor = None, camera_direction: torch.Tensor = None,
                 near: float = 1, far: float = 100, light_mode: str = "surface",
                 light_intensity_ambient: float = 0.5, light_intensity_directional: float = 0.5,
                 light_color_ambient: torch.Tensor = None, light_color_directional: torch.Tensor = None,
                 light_direction: torch.Tensor = None, device="cpu"):
        self.image_size, self.anti_aliasing, self.device = image_size, anti_aliasing, device
        self.bg_color = torch.zeros(3, device=device) if bg_color is None else bg_color
        self.fill_back, self.camera_mode = fill_back, camera_mode
        if camera_mode == "projection":
            raise NotImplementedError("camera_mode='projection' is not used by any demo of the reference "
                                      "(its render() never calls perspective_projection with K either)")
        if camera_mode not in ("look", "look_at"):
            raise ValueError("camera_mode must be 'look' or 'look_at'")
        self.perspective_distort, self.viewing_angle, self.viewing_scale = perspective_distort, viewing_angle, viewing_scale
        if eye is None:
            self.eye = torch.tensor([0, 0, -(1.0 / math.tan(math.radians(viewing_angle)) + 1)], device=device)
        else:
            self.eye = eye.to(device)
        # the reference overwrites the constructor's camera_direction in both look modes (softrenderer.py:228)
        self.camera_direction = torch.tensor([0, 0, 1.0], device=device)
        self.near, self.far = near, far
        self.sigma_val, self.dist_func, self.dist_eps, self.gamma_val = sigma_val, dist_func, dist_eps, gamma_val
        self.aggr_func_rgb, self.aggr_func_alpha, self.texture_type = aggr_func_rgb, aggr_func_alpha, texture_type
        self.light_intensity_ambient = light_intensity_ambient
        self.light_intensity_directional = light_intensity_directional
        ones = torch.ones(3, device=device)
        self.light_color_ambient = ones if light_color_ambient is None else light_color_ambient.to(device)
        self.light_color_directional = ones.clone() if light_color_directional is None else light_color_directional.to(device)
        self.light_direction = (torch.tensor([0, 1.0, 0], device=device) if light_direction is None
                                else light_direction.to(device))
        self.rasterizer_eps = 1e-3

    # ---- softrenderer.py:266-330 ------------------------------------------------------------------------------
    def forward(self, vertices, faces, textures=None, mode=None, K=None, rmat=None, tvec=None):
        return self.render(vertices, faces, textures, mode, K, rmat, tvec)

    __call__ = forward

    def render(self, vertices, faces, textures=None, mode=None, K=None, rmat=None, tvec=None):
        if mode not in ["depth", "silhouette"]:
            textures = self.lighting(vertices, faces, textures)
        vertices = self.transform_to_camera_frame(vertices)
        vertices = self.project_to_image(vertices)
        return self.rasterize(vertices, faces, textures)

    # ---- softrenderer.py:332-365, lighting.py:23-196 ------------------------------------------------------------
    def lighting(self, vertices, faces, textures):
        fv = self.vertices_to_faces(vertices, faces)
        B, F = fv.shape[:2]
        dev = fv.device
        ambient = torch.zeros(B, F, 3, device=dev)
        ia = min(max(float(self.light_intensity_ambient), 0.0), 1.0)
        if ia != 0:
            ambient = ambient + ia * self.light_color_ambient.to(dev)[None, None, :]
        directional = torch.zeros(B, F, 3, device=dev)
        idr = min(max(float(self.light_intensity_directional), 0.0), 1.0)
        if idr != 0:
            # the reference calls compute_directional_light WITHOUT a direction (softrenderer.py:355-360): the light
            # always comes from lighting.py's default (0, 1, 0), whatever light_direction the renderer was given
            direction = torch.tensor([0.0, 1.0, 0.0], device=dev)[None, None, :]
            v10 = fv[:, :, 0] - fv[:, :, 1]
            v12 = fv[:, :, 2] - fv[:, :, 1]
            normals = Fn.normalize(torch.cross(v12, v10, dim=2), p=2, dim=2, eps=1e-6)
            cos = Fn.relu(torch.sum(normals * direction, dim=2))
            directional = directional + idr * (self.light_color_directional.to(dev)[None, None, :] * cos[:, :, None])
        return (ambient[:, :, None, :] + directional[:, :, None, :]) * textures

    # ---- softrenderer.py:367-428 --------------------------------------------------------------------------------
    def transform_to_camera_frame(self, vertices):
        if self.camera_mode == "look_at":
            return self.look_at(vertices, self.eye)
        return self.look(vertices, self.eye, self.camera_direction)

    def project_to_image(self, vertices):
        return self.perspective_distortion(vertices, angle=self.viewing_angle)

    # ---- softrenderer.py:430-467 --------------------------------------------------------------------------------
    def rasterize(self, vertices, faces, textures):
        face_vertices = self.vertices_to_faces(vertices, faces)
        face_textures = textures
        if self.texture_type == "vertex":
            face_textures = self.textures_to_faces(textures, faces)
        image_size = self.image_size * (2 if self.anti_aliasing else 1)
        out = soft_rasterize(face_vertices, face_textures, image_size, self.bg_color, self.near, self.far,
                             self.fill_back, self.rasterizer_eps, self.sigma_val, self.dist_func, self.dist_eps,
                             self.gamma_val, self.aggr_func_rgb, self.aggr_func_alpha, self.texture_type)
        if self.anti_aliasing:
            out = Fn.avg_pool2d(out, kernel_size=2, stride=2)
        return out

    # ---- softrenderer.py:469-610 --------------------------------------------------------------------------------
    @staticmethod
    def _frame(z_dir, up):
        z_axis = _normalize(z_dir, 1e-5)
        x_axis = _normalize(torch.cross(up, z_axis, dim=1), 1e-5)
        y_axis = _normalize(torch.cross(z_axis, x_axis, dim=1), 1e-5)
        return torch.cat((x_axis[:, None, :], y_axis[:, None, :], z_axis[:, None, :]), dim=1)

    def look_at(self, vertices, eye, at: Optional[torch.Tensor] = None, up: Optional[torch.Tensor] = None):
        dev, B = vertices.device, vertices.shape[0]
        eye = eye.to(dev)
        at = torch.tensor([0.0, 0.0, 0.0], device=dev) if at is None else at.to(dev)
        up = torch.tensor([0.0, 1.0, 0.0], device=dev) if up is None else up.to(dev)
        eye, at, up = (t[None, :].repeat(B, 1) if t.dim() == 1 else t for t in (eye, at, up))
        R = self._frame(at - eye, up)
        if vertices.shape != eye.shape:
            eye = eye[:, None, :]
        return torch.matmul(vertices - eye, R.transpose(1, 2))

    def look(self, vertices, eye, direction: Optional[torch.Tensor] = None, up: Optional[torch.Tensor] = None):
        dev = vertices.device
        direction = torch.tensor([0, 1.0, 0], device=dev) if direction is None else direction.to(dev)
        up = torch.tensor([0, 1.0, 0], device=dev) if up is None else up.to(dev)
        eye = eye.to(dev)
        eye, direction, up = (t[None, :] if t.dim() == 1 else t for t in (eye, direction, up))
        R = self._frame(direction, up)
        if vertices.shape != eye.shape:
            eye = eye[:, None, :]
        return torch.matmul(vertices - eye, R.transpose(1, 2))

    def perspective_distortion(self, vertices, angle: Optional[float] = 30.0):
        width = torch.tan(torch.tensor([angle / 180 * math.pi], device=vertices.device))[:, None]
        z = vertices[..., 2]
        xy = vertices[..., :2] / (z * width).unsqueeze(-1)
        return torch.cat((xy, z.unsqueeze(-1)), dim=2)

    def vertices_to_faces(self, vertices, faces):
        B, V = vertices.shape[:2]
        faces = faces.long() + (torch.arange(B, dtype=torch.long, device=vertices.device) * V)[:, None, None]
        return vertices.reshape(B * V, 3)[faces]

    def textures_to_faces(self, textures, faces):
        B, F = textures.shape[:2]
        faces = faces.long() + (torch.arange(B, device=textures.device) * F)[:, None, None]
        return textures.reshape(B * F, 2, 3)[faces]

    def set_eye_from_angles(self, distance, elevation, azimuth, degrees: bool = True):
        if degrees:
            elevation, azimuth = math.radians(float(elevation)), math.radians(float(azimuth))
        d = float(distance)
        self.eye = torch.tensor([d * math.cos(elevation) * math.sin(azimuth), d * math.sin(elevation),
                                 -d * math.cos(elevation) * math.cos(azimuth)], device=self.device)
