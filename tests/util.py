"""Shared helpers for the parity tests (test infrastructure)."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    out = {}
    for k in z.files:
        a = z[k]
        out[k] = torch.from_numpy(a) if a.ndim > 0 else a.item()
    return out


def rel_l2(a, b):
    a = a.double().cpu()
    b = b.double().cpu()
    den = b.norm().item()
    return (a - b).norm().item() / (den if den > 0 else 1.0)


def rel_linf(a, b):
    """max |a-b| scaled by max |b| (scale-relative, robust to sign changes/zeros)."""
    a = a.double().cpu()
    b = b.double().cpu()
    den = b.abs().max().item()
    return (a - b).abs().max().item() / (den if den > 0 else 1.0)


def assert_close(a, b, tol, what=""):
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    assert torch.isfinite(a.double()).all(), f"{what}: non-finite values"
    e = rel_linf(a, b)
    assert e <= tol, f"{what}: scaled max error {e:.3e} > {tol:.1e}"
