"""z-slab decomposition on ONE GPU: P virtual ranks (halos by device copies) must reproduce the one-GPU
FerroSolver run BIT FOR BIT on every owned cell.  The multi-process NCCL path uses the
same generator and the exchanger tested in tests/test_slab_host.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def build(res, P, path, walls_open, steps, kappa=0.01, gravity=1e-4, droplets=False, push=False, **slab_kw):
    from lbm_ferrofluid_b200 import initializers as I
    from lbm_ferrofluid_b200.distributed import SlabFerroSolver, SlabLayout, slab_shell_flags, step_virtual_ranks
    from lbm_ferrofluid_b200.solver import FerroSolver

    dev = torch.device("cuda")
    flags = I.shell_flags(res, device=dev)
    mflags = I.shell_flags(res, open_axes=walls_open, device=dev)
    if droplets:
        rho = I.two_droplets_rho(res, flags)
    else:
        rho = I.rosensweig_rho(res, flags, noise=1e-3, seed=3)
    kw = dict(H_ext=[0.0, 100.0, 0.0], n_sweeps=6, magnetic_path=path, kappa=kappa, gravity=gravity,
              contact_angle=0.55 * 3.141592653589793)
    one = FerroSolver(flags, magnetic_flags=mflags, keep_force=False, **kw).initialize(rho).step(steps)
    slabs = []
    for r in range(P):
        lay = SlabLayout(res[0], P, r)
        lf = slab_shell_flags(lay, res[1], res[2], device=dev)
        lm = slab_shell_flags(lay, res[1], res[2], open_axes=walls_open, device=dev)
        assert torch.equal(lf, lay.take(flags)) and torch.equal(lm, lay.take(mflags))
        slabs.append(SlabFerroSolver(lay, lf, magnetic_flags=lm, **kw, **slab_kw).initialize(lay.take(rho)))
    step_virtual_ranks(slabs, steps, push=push)
    torch.cuda.synchronize()
    return one, slabs


@pytest.mark.parametrize("P", [2, 3])
@pytest.mark.parametrize("path", ["faithful", "scalar"])
@pytest.mark.parametrize("walls_open", [(1,), ()])
def test_virtual_ranks_bit_identical_to_one_gpu(P, path, walls_open):
    res = (24, 14, 18)
    one, slabs = build(res, P, path, walls_open, steps=3)
    fluid = (one.mflags != 2)
    for name in ("f", "g", "rho", "pressure", "vel", "psi"):
        got = torch.cat([s.owned(name) for s in slabs], dim=2)
        want = getattr(one, name)
        if name == "psi":   # wall-cell psi is never observed (shell rule overwrites it)
            got, want = got * fluid, want * fluid
        assert torch.equal(got, want), f"{name}: P={P} {path} differs from the one-GPU run " \
                                       f"(max {float((got - want).abs().max()):.3e})"
    if path == "faithful":
        got = torch.cat([s.owned("h") for s in slabs], dim=2)
        assert torch.equal(got, one.h)


@pytest.mark.parametrize("P", [2, 3])
@pytest.mark.parametrize("path", ["faithful", "scalar"])
@pytest.mark.parametrize("walls_open", [(1,), ()])
def test_fused_boundary_sweep_and_halo_push_is_bit_identical(P, path, walls_open):
    """lbmff_slab_magnetic_sweep_push_d3q19: the boundary-plane sweep stores the z-moving populations straight into the
    neighbour's halo plane (here: the neighbour virtual rank's tensor; across GPUs the same pointer arithmetic runs on
    symmetric-memory mappings, SymmPushExchanger) -- no exchange of h at all, and still the one-GPU result bit for bit,
    h included.  "scalar": the psi-only sweeps push their one psi plane the same way (lbmff_slab_magnetic_sweep_psi_3d)."""
    res = (24, 14, 18)
    one, slabs = build(res, P, path, walls_open, steps=3, push=True)
    fluid = (one.mflags != 2)
    for name in ("f", "g", "rho", "pressure", "vel", "psi") + (("h",) if path == "faithful" else ()):
        got = torch.cat([s.owned(name) for s in slabs], dim=2)
        want = getattr(one, name)
        if name == "psi":
            got, want = got * fluid, want * fluid
        assert torch.equal(got, want), f"{name}: P={P} differs (max {float((got - want).abs().max()):.3e})"


def test_virtual_ranks_two_droplets_no_gravity():
    one, slabs = build((32, 16, 40), 4, "faithful", (), steps=2, kappa=0.5, gravity=0.0, droplets=True)
    for name in ("f", "g", "rho", "pressure"):
        assert torch.equal(torch.cat([s.owned(name) for s in slabs], dim=2), getattr(one, name)), name


def test_slab_mass_is_conserved():
    one, slabs = build((24, 14, 18), 2, "faithful", (1,), steps=4)
    assert abs(sum(s.total_mass() for s in slabs) - one.total_mass()) <= 1e-9 * abs(one.total_mass())


def test_slab_entry_argument_errors():
    import ctypes

    from lbm_ferrofluid_b200 import _host as H
    from lbm_ferrofluid_b200._lib import Slab

    t = torch.zeros((1, 1, 8, 8, 8), device="cuda")
    fl = torch.ones((1, 1, 8, 8, 8), dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match="plane range"):
        H.launch("lbmff_slab_shell_rule_3d", t.device, H.ptr(t), H.ptr(fl), ctypes.byref(Slab(16, 0, 4, 8)),
                 1, 8, 8, 8)
    with pytest.raises(RuntimeError, match="D <= Dg"):
        H.launch("lbmff_slab_shell_rule_3d", t.device, H.ptr(t), H.ptr(fl), ctypes.byref(Slab(4, 0, 0, 8)),
                 1, 8, 8, 8)


@pytest.mark.parametrize("world", [2, 4])
def test_nccl_ranks_bit_identical_to_one_gpu(world):
    """The REAL data plane (one process per GPU, NCCL halos on the side stream overlapping the interior kernels)
    against the one-GPU run, bit for bit (tools/verify_slab.py).  Needs `world` GPUs on the box; bench.py runs the
    same check before every N > 1 timing."""
    import json
    import os
    import socket
    import subprocess
    import sys

    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(root, "tools", "verify_slab.py")], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    rep = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert rep["slab_bit_identical"] and all(p["bit_identical"] for p in rep["paths"].values())


def test_slab_scalar_path_refused_with_interior_magnetic_wall():
    """the scalar sweep kernels assume magnetic walls on the global shell only; the slab solver must
    check it like FerroSolver does."""
    from lbm_ferrofluid_b200.distributed import SlabFerroSolver, SlabLayout, slab_shell_flags

    dev = torch.device("cuda")
    lay = SlabLayout(24, 2, 0)
    lf = slab_shell_flags(lay, 14, 18, device=dev)
    lm = slab_shell_flags(lay, 14, 18, open_axes=(1,), device=dev)
    lm[0, 0, 6, 5, 5] = 2
    kw = dict(H_ext=[0.0, 100.0, 0.0], n_sweeps=6)
    with pytest.raises(RuntimeError, match="outer shell"):
        SlabFerroSolver(lay, lf, magnetic_flags=lm, magnetic_path="scalar", **kw)
    assert SlabFerroSolver(lay, lf, magnetic_flags=lm, magnetic_path="auto", **kw).magnetic_path == "faithful"
    lm[0, 0, 6, 5, 5] = 7
    with pytest.raises(RuntimeError, match="FLUID / OBSTACLE"):
        SlabFerroSolver(lay, lf, magnetic_flags=lm, magnetic_path="faithful", **kw)


@pytest.mark.parametrize("path", ["faithful", "scalar"])
def test_virtual_ranks_bit_identical_with_the_tma_kernel(path, monkeypatch):
    """Same contract on a grid wide enough (W = 72) for k_hcz_collide_stream_tma to run on the interior planes."""
    one, slabs = build((24, 12, 72), 2, path, (1,), steps=3)
    fluid = (one.mflags != 2)
    for name in ("f", "g", "rho", "pressure", "vel", "psi"):
        got, want = torch.cat([s.owned(name) for s in slabs], dim=2), getattr(one, name)
        if name == "psi":
            got, want = got * fluid, want * fluid
        assert torch.equal(got, want), name


@pytest.mark.parametrize("path", ["faithful", "scalar"])
@pytest.mark.parametrize("P,res,chunk,steps", [(2, (24, 14, 18), 3, 3), (2, (32, 12, 72), 5, 4), (3, (48, 10, 20), 7, 3)])
def test_compressed_streaming_is_bit_identical(P, res, chunk, steps, path):
    """streaming="compressed": ONE f and ONE g buffer with chunk + 1 spare planes, passes issued in chunks whose output
    is displaced by -/+ (chunk + 1) planes (ascending / descending steps alternate).  Same kernels, so every owned
    cell must equal the one-GPU A/B run bit for bit -- after odd and even numbers of steps (both displacements)."""
    one, slabs = build(res, P, path, (1,), steps, streaming="compressed", chunk=chunk)
    fluid = (one.mflags != 2)
    for name in ("f", "g", "rho", "pressure", "vel", "psi"):
        got, want = torch.cat([s.owned(name) for s in slabs], dim=2), getattr(one, name)
        if name == "psi":
            got, want = got * fluid, want * fluid
        assert torch.equal(got, want), f"{name} differs (P={P}, chunk={chunk}, {steps} steps)"
    ab = build(res, P, path, (1,), 1)[1][0]
    assert slabs[0].state_bytes < ab.state_bytes      # one f and one g buffer (+ spare planes) instead of two each


def test_compressed_streaming_argument_errors():
    from lbm_ferrofluid_b200.distributed import SlabFerroSolver, SlabLayout, slab_shell_flags

    lay = SlabLayout(24, 2, 0)
    lf = slab_shell_flags(lay, 14, 18, device="cuda")
    with pytest.raises(RuntimeError, match="spare planes"):
        SlabFerroSolver(lay, lf, streaming="compressed", chunk=16)
    with pytest.raises(ValueError):
        SlabFerroSolver(lay, lf, streaming="esoteric")
