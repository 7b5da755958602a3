"""lbmff_get_H_int_d2q9_static: the allocation-free variant of lbmff_get_H_int_d2q9 (include/lbmff.h).

No reference counterpart: the psi scratch is a static __device__ array
inside liblbmff.so instead of a caller-provided buffer.  The test calls both entry points through the C ABI on the
same inputs, twice in a row (the second call sees the first call's scratch contents and a different grid size), and
requires bit-identical outputs: the two entry points launch the same kernels, only the scratch address differs.
Collected last (file name) so that a failure here cannot hide another test under `pytest -x`.
"""
import ctypes

import pytest
import torch

from lbm_ferrofluid_b200 import _host as H
from lbm_ferrofluid_b200 import _lib

pytestmark = pytest.mark.gpu


def _inputs(B, Hh, W, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    dev = torch.device("cuda:0")
    phi = torch.rand((B, 1, Hh, W), generator=g).to(dev)
    flags = torch.ones((B, 1, Hh, W), dtype=torch.uint8)      # CellType.FLUID
    flags[..., 0, :] = flags[..., -1, :] = 2                  # CellType.OBSTACLE rim
    flags[..., :, 0] = flags[..., :, -1] = 2
    flags[0, 0, Hh // 2, W // 3] = 2                          # one interior wall cell
    Hx = (0.5 * torch.randn((B, 1, Hh, W + 1), generator=g)).to(dev)
    Hy = (100.0 + 0.5 * torch.randn((B, 1, Hh + 1, W), generator=g)).to(dev)
    h = (0.1 * torch.randn((B, 9, Hh, W), generator=g)).to(dev)
    return phi, flags.to(dev), Hx, Hy, h


def _run_both(B, Hh, W, seed, tau):
    phi, flags, Hx, Hy, h = _inputs(B, Hh, W, seed)
    out = {}
    for name in ("lbmff_get_H_int_d2q9", "lbmff_get_H_int_d2q9_static"):
        H_int = torch.full((B, 2, Hh, W), float("nan"), device=h.device)
        h_new = torch.full_like(h, float("nan"))
        args = [H.ptr(phi), H.ptr(flags), H.ptr(Hx), H.ptr(Hy), H.ptr(h), tau, 1.0, 1.0, H.ptr(H_int), H.ptr(h_new)]
        if name.endswith("_static"):
            H.launch(name, h.device, *args, B, Hh, W)
        else:
            psi = torch.empty((B, 1, Hh, W), device=h.device)
            H.launch(name, h.device, *args, H.ptr(psi), B, Hh, W)
        torch.cuda.synchronize()
        out[name] = (H_int, h_new)
    return out


def test_static_scratch_matches_caller_scratch_two_calls_in_a_row():
    # second call: smaller grid, other inputs -- the array still holds the first call's psi beyond the new extent
    for B, Hh, W, seed, tau in ((2, 130, 258, 1, 1.0), (1, 37, 53, 2, 0.9)):
        out = _run_both(B, Hh, W, seed, tau)
        (Ha, ha), (Hb, hb) = out["lbmff_get_H_int_d2q9"], out["lbmff_get_H_int_d2q9_static"]
        assert torch.isfinite(Hb).all() and torch.isfinite(hb).all()
        # same kernels, same arithmetic: the tolerance is zero
        assert torch.equal(Ha, Hb), f"H_int differs: max |d| = {(Ha - Hb).abs().max().item():.3e}"
        assert torch.equal(ha, hb), f"h differs: max |d| = {(ha - hb).abs().max().item():.3e}"


def test_static_scratch_rejects_sizes_beyond_the_limit():
    lib = _lib.load()
    cap = lib.lbmff_static_psi_max_floats()
    assert cap == 1 << 20
    B, Hh, W = 1, 1025, 1024                         # one row over the limit; rejected before any pointer is used
    assert B * Hh * W > cap
    null = ctypes.c_void_p(0)
    with pytest.raises(RuntimeError, match="exceeds LBMFF_STATIC_PSI_MAX_FLOATS"):
        H.launch("lbmff_get_H_int_d2q9_static", torch.device("cuda:0"), null, null, null, null, null, 1.0, 1.0, 1.0,
                 null, null, B, Hh, W)
    # exactly at the limit is accepted
    phi, flags, Hx, Hy, h = _inputs(1, 1024, 1024, 3)
    H_int = torch.empty((1, 2, 1024, 1024), device=h.device)
    h_new = torch.empty_like(h)
    H.launch("lbmff_get_H_int_d2q9_static", h.device, H.ptr(phi), H.ptr(flags), H.ptr(Hx), H.ptr(Hy), H.ptr(h), 1.0,
             1.0, 1.0, H.ptr(H_int), H.ptr(h_new), 1, 1024, 1024)
    torch.cuda.synchronize()
    assert torch.isfinite(H_int).all()
