#!/usr/bin/env python
"""Kernel micro-benchmarks through the C ABI (development tool, not part of the product path).

    python tools/kbench.py sweep D H W [n_sweeps]     -> ms per population-faithful sweep, GB/s at 152 B/cell
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbm_ferrofluid_b200 import _host as H  # noqa: E402
from lbm_ferrofluid_b200 import initializers as I  # noqa: E402


def sweep(res, n=30, reps=3):
    D, Hh, W = res
    dev = torch.device("cuda")
    N = D * Hh * W
    hA = torch.rand((1, 19, D, Hh, W), device=dev) * 1e-3
    hB = torch.zeros_like(hA)
    mfl = I.shell_flags(res, open_axes=(1,), device=dev)
    rhs = torch.rand((1, 1, D, Hh, W), device=dev) * 1e-3
    psi = torch.zeros_like(rhs)
    run = lambda: H.launch("lbmff_magnetic_sweeps_d3q19", dev, H.ptr(hA), H.ptr(hB), H.ptr(mfl), H.ptr(rhs),  # noqa
                           H.ptr(psi), n, 1.0, 1.0, 1.0, 1, D, Hh, W)
    run()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / n)
    print(f"sweep variant={os.environ.get('LBMFF_SWEEP_VARIANT', 'default')} res={res} ms={best:.4f} "
          f"GB/s={152 * N / best / 1e6:.1f} pages/pop={N * 4 / 2**21:.2f}", flush=True)


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


def streams(res):
    """Ceilings: one-stream copy (torch), 19-in/19-out shifted copy (propagation kernel, no arithmetic)."""
    D, Hh, W = res
    dev = torch.device("cuda")
    N = D * Hh * W
    a = torch.rand((1, 19, D, Hh, W), device=dev)
    b = torch.zeros_like(a)
    ms = timed(lambda: b.copy_(a))
    print(f"torch copy 19N floats: ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)
    ms = timed(lambda: H.launch("lbmff_propagation_d3q19", dev, H.ptr(a), H.ptr(b), 1, D, Hh, W))
    print(f"propagation d3q19 (38 streams, no math): ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)
    # same bytes as 19 separate one-population copies issued back to back (2 streams at a time)
    ms = timed(lambda: [b[0, q].copy_(a[0, q]) for q in range(19)])
    print(f"19 one-population copies: ms={ms:.4f} GB/s={152 * N / ms / 1e6:.1f}", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "streams":
        streams(tuple(int(v) for v in sys.argv[2:5]))
    if sys.argv[1] == "sweep":
        sweep(tuple(int(v) for v in sys.argv[2:5]), int(sys.argv[5]) if len(sys.argv) > 5 else 30)
