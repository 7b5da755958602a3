"""Host-side plumbing shared by the operator classes: argument checks, raw pointers, launches."""
from __future__ import annotations

import ctypes
import math

import torch

from . import _lib

FluidParams = _lib.FluidParams


def need_cuda(t: torch.Tensor, name: str, dtype=torch.float32) -> torch.Tensor:
    if not torch.is_tensor(t):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(
            f"{name} is on {t.device}: lbm_ferrofluid_b200 runs hand-written CUDA kernels only (no CPU fallback)")
    if t.dtype != dtype:
        if dtype == torch.uint8:
            t = t.to(torch.uint8)
        else:
            raise TypeError(f"{name} must be {dtype} (got {t.dtype}); the sm_100a kernels compute in fp32")
    return t if t.is_contiguous() else t.contiguous()


def dims(t: torch.Tensor):
    """(dim, B, C, D, H, W) of a [B, C, (D,) H, W] tensor (D = 1 in 2-D)."""
    if t.dim() == 4:
        B, C, H, W = t.shape
        return 2, B, C, 1, H, W
    if t.dim() == 5:
        B, C, D, H, W = t.shape
        return 3, B, C, D, H, W
    raise RuntimeError("expected a [B, C, (D,) H, W] tensor")


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch(name: str, device, *args):
    """Call a C-ABI entry point with `device` current, on torch's current stream (appended last)."""
    with torch.cuda.device(device):
        _lib.call(name, *args, stream())


def sizes(dim, B, D, H, W):
    return (B, H, W) if dim == 2 else (B, D, H, W)


def suffix(dim):
    return "d2q9" if dim == 2 else "d3q19"


def cot_of_contact_angle(contact_angle) -> float:
    """tan(pi/2 - theta) evaluated the way the reference does it (fp32 tensor arithmetic, HCZ_3d.py:97)."""
    if torch.is_tensor(contact_angle):
        ca = contact_angle.detach().to(torch.float32).reshape(-1)[:1].cpu()
    else:
        ca = torch.tensor([float(contact_angle)], dtype=torch.float32)
    return float(torch.tan(math.pi / 2.0 - ca).item())


def fluid_params(op, dx=1.0, dt=1.0, cot=0.0, gravity=0.0) -> FluidParams:
    return FluidParams(
        dx=float(dx), dt=float(dt),
        density_liquid=float(op._density_liquid), density_gas=float(op._density_gas),
        rho_liquid=float(op._rho_liquid), rho_gas=float(op._rho_gas),
        kappa=float(getattr(op, "_kappa", 0.0)), tau=float(op._tau),
        tau_f=float(getattr(op, "_tau_f", 1.0)), tau_g=float(getattr(op, "_tau_g", 1.0)),
        gravity=float(gravity), cot_theta=float(cot))


class nvtx:
    """NVTX range around a stage of the step (visible in nsys / ncu --nvtx; a no-op without a profiler)."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        torch.cuda.nvtx.range_pop()
