#!/usr/bin/env python
"""bench.py -- MLUPS of the 3-D Rosensweig HCZ + magnetic D3Q19 step (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--res D H W] [--magnetic-path faithful|scalar|auto]

One "step" = one outer time step of demo_3d_LBM_Rosensweig_instability.py (stream f,g; moments; 30 magnetic
sweeps; capillary process; HCZ collision) on a synthetic perturbed-interface field.  MLUPS counts every cell
of the array (wall shell included) x outer steps.  Prints ONE JSON line (rank 0).

 * value     : device-resident throughput (state already in HBM), CUDA events, max over ranks.
 * e2e       : same metric through the public API with HOST buffers: every step uploads that step's applied
               field (the three H_ext_mac face arrays get_H_int takes) from pinned memory and reads the order
               parameter phi (what the demos render) back to pinned memory.
 * roofline  : dominant kernel, algorithmic bytes / CUDA-event duration vs MEASURED_PEAKS.json.
 * cpu_baseline / --impl reference : the REFERENCE's own PyTorch implementation on the host cores (kind "reference":
               the byte-compiled copy oracle/_ref/pyref.zip run by oracle/ref_runner.py in its own process), bounded
               sample; the oracle port (oracle/lbm_oracle.py, kind "port") only if that copy is absent.
 * configs   : (N = 1) the other BASELINE.json configurations, each a short timed run with its own byte model.
 * slab_verify (N > 1) : before timing, a small case is run through the real NCCL exchanger and compared bit for
               bit with a one-GPU run (lbm_ferrofluid_b200.distributed.verify_against_single_gpu).
The timed region of `value` is repeated --reps times (default 3); value / ms_per_step are the MEDIAN repetition.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MLUPS (3D Rosensweig HCZ+magnetic)"
N_SWEEPS = 30
MAG = 100.0
BYTES_D3Q19 = 152          # one read + one write of 19 fp32 populations
BYTES_SCALAR_SWEEP = 20    # scalar sweep: 8 B state read + 8 B written + 4 B source
BYTES_HCZ = 304            # f + g, one read + one write each


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--res", type=int, nargs=3, default=None, help="D H W per GPU (default 512 512 512)")
    ap.add_argument("--magnetic-path", default=os.environ.get("LBMFF_MAGNETIC_PATH", "faithful"),
                    choices=["faithful", "scalar", "auto"])
    ap.add_argument("--cpu-res", type=int, nargs=3, default=[96, 96, 96], help="grid of the bounded CPU sample")
    ap.add_argument("--workload", default="rosensweig", choices=["rosensweig", "two_droplets"],
                    help="rosensweig: the metric's configuration (default).  two_droplets: BASELINE.json configs[4], "
                         "demo_3d_LBM_two_droplets scaled up, e.g. --gpus 8 --res 128 1024 1024 for 1024^3")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fast-path", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the block of the other BASELINE configs")
    ap.add_argument("--no-ref-eager", action="store_true", help="skip the reference-eager-on-this-GPU number")
    ap.add_argument("--no-verify", action="store_true", help="skip the N-rank bit-identity check")
    ap.add_argument("--streaming", default="ab", choices=["ab", "compressed"],
                    help="N > 1: 'compressed' keeps ONE f and ONE g buffer with chunk+1 spare planes (1024^3 on 2 GPUs)")
    ap.add_argument("--chunk", type=int, default=32, help="planes per launch of the compressed passes")
    ap.add_argument("--reps", type=int, default=3, help="repetitions of the K-step timed region (median reported)")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ---------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# ---------------------------------------------------------------------------------------------------
def _ref_runner(argv, timeout):
    """Run oracle/ref_runner.py (the unmodified reference in its own process) and parse its JSON line."""
    env = {k: v for k, v in os.environ.items() if k not in ("PYTHONPATH", "RANK", "WORLD_SIZE", "LOCAL_RANK")}
    r = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "ref_runner.py"), *argv], capture_output=True,
                       text=True, env=env, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(r.stderr.strip().splitlines()[-1] if r.stderr.strip() else "ref_runner failed")
    return json.loads(r.stdout.strip().splitlines()[-1])


def reference_present():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "pyref.zip")) or \
        os.path.isdir("/root/reference/src/LBM")


def cpu_rosensweig_mlups(res, steps, warmup):
    """The reference's own CPU path on a bounded grid: (MLUPS, ms/step, cores, kind).  kind "reference" = the
    unmodified reference classes driven through the demo loop (oracle/ref_runner.py); "port" = oracle.hcz_step,
    only when the byte-compiled reference did not travel."""
    cores = os.cpu_count() or 1
    if reference_present():
        try:
            j = _ref_runner(["time", "--dim", "3", "--res", *map(str, res), "--mag", str(MAG), "--steps", str(steps),
                             "--warmup", str(warmup), "--device", "cpu", "--flush-denormal"], timeout=1500)
            return j["mlups"], j["ms_per_step"], j["cores"], "reference"
        except Exception as e:                               # fall back to the port, loudly
            sys.stderr.write(f"bench.py: reference CPU run failed ({e}); timing the oracle port instead\n")
    from oracle import lbm_oracle as O
    from oracle.make_golden import perturbed_interface, shell_flags

    torch.set_num_threads(cores)
    # psi decays exponentially away from the interface, so the first steps are full of denormals, which the
    # host cores process ~50x slower (measured: 22 s vs 0.39 s per step at 48^3).  Flush-to-zero favours the
    # CPU arm and is used for TIMING only (parity runs never set it).
    torch.set_flush_denormal(True)
    p = O.HCZParams(contact_angle=torch.tensor([0.5 * math.pi]), n_sweeps=N_SWEEPS)
    flags, mflags = shell_flags(res), shell_flags(res, open_axes=(1,))
    rho = perturbed_interface(res, 0.5, p.rho_liquid, p.rho_gas, torch.float32)
    rho[flags == 2] = p.rho_liquid
    vel = torch.zeros((1, 3, *res))
    pressure, f, g = O.hcz_init(rho, rho.clone(), vel, p)
    Hext = torch.zeros((1, 3, *res))
    Hext[:, 1] = MAG
    s = dict(f=f, g=g, h=torch.zeros_like(f), rho=rho, vel=vel, density=rho.clone(), pressure=pressure,
             flags=flags, mflags=mflags, H_ext=Hext, H_ext_mac=O.get_staggered(Hext))
    for _ in range(warmup):
        s = O.hcz_step(s, p)
    t0 = time.perf_counter()
    for _ in range(steps):
        s = O.hcz_step(s, p)
    dt = time.perf_counter() - t0
    cells = res[0] * res[1] * res[2]
    return cells * steps / dt / 1e6, dt / steps * 1e3, cores, "port"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    res = tuple(args.cpu_res)
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    # bound the whole run to a few minutes: ~3.5 s per step at 96^3 on 8 cores
    steps, warmup = min(steps, 10), min(warmup, 3)
    mlups, ms, cores, kind = cpu_rosensweig_mlups(res, steps, warmup)
    sample = (f"same workload at {res[0]}x{res[1]}x{res[2]} (bounded sample), {warmup} warm-up + {steps} timed steps, "
              f"torch CPU threads = {cores}, denormals flushed to zero (timing only)")
    note = ("the reference's own src/LBM classes driven through the demo loop on the host cores (byte-compiled copy "
            "oracle/_ref/pyref.zip, unmodified; 512^3 is out of its reach: three [1,19,3,D,H,W] temporaries of 30 GB)"
            if kind == "reference" else
            "oracle port (torch-CPU restatement pinned against the reference's golden vectors): the byte-compiled "
            "reference oracle/_ref/pyref.zip was not found")
    line = {
        "impl": "reference", "metric": METRIC, "value": mlups, "unit": "MLUPS", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "demo_3d_LBM_Rosensweig_instability D3Q19 HCZ+magnetic mg100, 30 sweeps/step",
                   "res": list(res), "note": note},
        "cpu_baseline": {"value": mlups, "unit": "MLUPS", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": mlups, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------
# clocks sampler
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append([c.strip() for c in ln.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------
def build_solver(res, path, device, world=1, rank=0, workload="rosensweig", mag=MAG, **slab_kw):
    from lbm_ferrofluid_b200 import initializers as I
    from lbm_ferrofluid_b200.solver import FerroSolver

    if world > 1:
        from lbm_ferrofluid_b200.distributed import SlabFerroSolver
        if workload == "two_droplets":
            return SlabFerroSolver.two_droplets(res, path, device, world, rank, mag=1.0, n_sweeps=N_SWEEPS, **slab_kw)
        return SlabFerroSolver.rosensweig(res, path, device, world, rank, mag=mag, n_sweeps=N_SWEEPS, **slab_kw)
    if workload == "two_droplets":
        flags = I.shell_flags(res, device=device)
        s = FerroSolver(flags, magnetic_flags=flags.clone(), H_ext=[0.0, 1.0, 0.0], n_sweeps=N_SWEEPS,
                        magnetic_path=path, contact_angle=0.5 * math.pi, gravity=1e-4, kappa=0.5, keep_force=False)
        return s.initialize(I.two_droplets_rho(res, flags))
    flags = I.shell_flags(res, device=device)
    mflags = I.shell_flags(res, open_axes=(1,), device=device)
    rho = I.rosensweig_rho(res, flags)
    s = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, mag, 0.0], n_sweeps=N_SWEEPS, magnetic_path=path,
                    contact_angle=0.5 * math.pi, gravity=1e-4, keep_force=False)
    s.initialize(rho)
    return s


def _event_ms(fn, steps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    fn(steps)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b)


def other_configs(device, peak):
    """The other BASELINE.json configurations on one GPU: a short timed run each (3 warm-up
    steps, median of 3 repetitions, CUDA events; every state is far larger than the 126 MB L2), with the byte model of
    DESIGN.md section 3 and the fraction of the measured copy bandwidth that model reaches."""
    from lbm_ferrofluid_b200 import initializers as I
    from lbm_ferrofluid_b200.solver import FerroSolver, SinglePhaseSolver
    from lbm_ferrofluid_b200.types import KBCType

    out = {}

    def measure(name, solver, cells, bytes_cell, steps, note):
        solver.step(3)
        ms = statistics.median(_event_ms(solver.step, steps) for _ in range(3)) / steps
        v = cells / ms / 1e3
        fld = solver.rho if hasattr(solver, "rho") else solver.f
        out[name] = {"value": v, "unit": "MLUPS", "ms_per_step": ms, "cells": cells, "steps_timed": steps,
                     "model_bytes_per_cell": bytes_cell, "model_gbs": v * 1e6 * bytes_cell / 1e9,
                     "frac_of_hbm_peak": v * 1e6 * bytes_cell / 1e9 / peak, "note": note,
                     "finite": bool(torch.isfinite(fld).all().item())}

    n = 4096
    flags = torch.ones((1, 1, n, n), dtype=torch.uint8, device=device)
    tau = 0.5 + 3 * 0.2 * n / 3000
    vel = I.double_shear_layer((n, n), device=device)
    rho = torch.full((1, 1, n, n), 0.265, device=device)
    measure("demo_2d_LBM 4096^2 D2Q9 BGK", SinglePhaseSolver(flags, tau=tau).initialize(rho, vel), n * n, 72, 100,
            "double shear layer, Re 3000, one fused stream+macro+collide kernel per step")
    measure("demo_2d_LBM_KBC 4096^2 D2Q9 KBC_A", SinglePhaseSolver(flags, tau=tau, KBC_type=int(KBCType.KBC_A))
            .initialize(rho, vel), n * n, 72, 100,
            "entropic KBC; the reference's 0/0 = NaN stabiliser cells are reproduced, so finite may be false")
    del flags, vel, rho
    res = (n, n)
    flags = I.shell_flags(res, device=device)
    mflags = I.shell_flags(res, open_axes=(1,), device=device)
    s = FerroSolver(flags, magnetic_flags=mflags, H_ext=[0.0, MAG], tau_f=0.7, tau_g=0.8,
                    contact_angle=0.75 * math.pi, KBC_type=int(KBCType.KBC_A), magnetic_path="faithful",
                    n_sweeps=N_SWEEPS, keep_force=False).initialize(I.rosensweig_rho(res, flags, frac=0.45))
    measure("demo_2d_LBM_Rosensweig_instability 4096^2 mg100 (HCZ + magnetic + KBC on g)", s, n * n,
            (2 + N_SWEEPS) * 72, 10, "population-faithful 30 sweeps; model (2+30) x 72 B")
    del s, flags, mflags
    torch.cuda.empty_cache()
    res = (256, 256, 256)
    s = build_solver(res, "faithful", device, mag=85.0)
    measure("demo_3d_LBM_Rosensweig_instability 256^3 mg85", s, 256 ** 3, (2 + N_SWEEPS) * BYTES_D3Q19, 10,
            "population-faithful 30 sweeps; model (2+30) x 152 B")
    del s
    torch.cuda.empty_cache()
    return out


def reference_eager_on_gpu(res=(256, 256, 256)):
    """BASELINE.md section 3's optional second baseline: the unmodified reference with device='cuda' on this GPU
    (eager PyTorch, ~4 400 launches per step), in its own process."""
    if not reference_present():
        return None
    try:
        j = _ref_runner(["time", "--dim", "3", "--res", *map(str, res), "--mag", str(MAG), "--steps", "3", "--warmup",
                         "1", "--device", "cuda"], timeout=900)
        return {"value": j["mlups"], "unit": "MLUPS", "ms_per_step": j["ms_per_step"], "res": list(res),
                "what": "unmodified reference (oracle/_ref/pyref.zip), eager PyTorch, device='cuda', fp32, same "
                        "workload; 512^3 is out of its reach (three [1,19,3,D,H,W] temporaries of 30 GB each)"}
    except Exception as e:
        return {"value": None, "error": str(e)[:200]}


def bind_to_gpu_numa_node(index):
    """Pin this rank's host threads to the CPU cores next to its GPU (NVML's ideal affinity) BEFORE any pinned buffer
    is allocated, so the e2e path's staging memory is first-touched on the GPU's own NUMA node instead of wherever the
    launcher started the process.  Returns the number of cores bound to, or None if NVML does not cooperate."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        n = len(os.sched_getaffinity(0))
        pynvml.nvmlShutdown()
        return n
    except Exception:
        return None


def run_ours(args):
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local) if world > 1 else None
    if world > 1:
        # halo messages ride NCCL's own stream: make it high priority so its few CTAs are scheduled ahead of
        # the interior-plane kernel they overlap with
        os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
        # NCCL prints its version banner to STDOUT at NCCL_DEBUG=VERSION; stdout carries exactly one JSON line here
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=device)
    res = tuple(args.res) if args.res else (512, 512, 512)
    cells_rank = res[0] * res[1] * res[2]
    cells = cells_rank * world

    # ---- N > 1: the data plane is checked against a one-GPU run, bit for bit, before anything is timed
    slab_verify = None
    if world > 1 and not args.no_verify:
        from lbm_ferrofluid_b200.distributed import verify_against_single_gpu
        slab_verify = verify_against_single_gpu(device, world, rank, (8 * world if world > 8 else 64, 48, 56), 3)
        torch.cuda.empty_cache()

    slab_kw = dict(streaming=args.streaming, chunk=args.chunk) if world > 1 else {}
    solver = build_solver(res, args.magnetic_path, device, world, rank, args.workload, **slab_kw)
    path = solver.magnetic_path
    state_gb = solver.state_bytes / 1e9 if hasattr(solver, "state_bytes") else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn(steps)
        b.record()
        barrier()
        ms = torch.tensor([a.elapsed_time(b)], device=device)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- device-resident throughput: --reps repetitions of the K-step timed region, median reported
    solver.step(args.warmup)
    solver.enable_kernel_timing(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    reps_ms = [timed(solver.step, args.steps) for _ in range(max(1, args.reps))]
    clocks = sampler.stop() if rank == 0 else None
    ktimes = solver.kernel_times_ms()
    solver.enable_kernel_timing(False)
    ms_total = statistics.median(reps_ms)
    ms_step = ms_total / args.steps
    mlups = cells * args.steps / (ms_total * 1e-3) / 1e6
    finite = bool(torch.isfinite(solver.rho).all().item())

    # ---- roofline of the dominant kernel (CUDA events bracketing it inside the timed region, all repetitions)
    peak, peak_src = peaks()
    sweeps_ms = statistics.mean(ktimes["sweeps"]) if ktimes.get("sweeps") else 0.0
    cs_ms = statistics.mean(ktimes["collide_stream"])
    if sweeps_ms >= cs_ms:
        per_launch_ms = sweeps_ms / N_SWEEPS
        bytes_cell = BYTES_D3Q19 if path == "faithful" else BYTES_SCALAR_SWEEP
        kname = "k_mag_sweep<3>" if path == "faithful" else "k_mag_sweep_scalar<3>"
    else:
        per_launch_ms, bytes_cell, kname = cs_ms, BYTES_HCZ, "k_hcz_collide_stream_tma"
    achieved = bytes_cell * cells_rank / (per_launch_ms * 1e-3) / 1e9
    # DRAM traffic of the same kernel (dram__bytes_read.sum + dram__bytes_write.sum of one `ncu --set full` launch):
    # profiles/ncu_traffic_r02.json is measured AT 512^3, so for the default workload the figure is that launch's own
    # traffic; other grids scale the measured bytes per cell
    traffic, traffic_src = None, None
    for fn in ("ncu_traffic_r02.json", "ncu_traffic_r01.json"):
        try:
            with open(os.path.join(ROOT, "profiles", fn)) as fh:
                tj = json.load(fh)
            per_cell = tj["dram_bytes_per_cell"].get(kname)
            if per_cell:
                traffic = per_cell * cells_rank / 1e9
                traffic_src = f"profiles/{fn}: {per_cell:.2f} B/cell measured at {tj.get('res', '256^3')}"
                break
        except Exception:
            pass
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_unit": "GB per launch (ncu dram bytes)",
                "traffic_source": traffic_src,
                "algorithmic_gb_per_launch": bytes_cell * cells_rank / 1e9, "peak_source": peak_src,
                "algorithmic_bytes_per_cell": bytes_cell, "avg_launch_ms": per_launch_ms,
                "share_of_step": {"sweeps": sweeps_ms / ms_step, "collide_stream": cs_ms / ms_step},
                "collide_stream": {"kernel": "k_hcz_collide_stream_tma", "avg_launch_ms": cs_ms,
                                   "algorithmic_bytes_per_cell": BYTES_HCZ,
                                   "achieved": BYTES_HCZ * cells_rank / (cs_ms * 1e-3) / 1e9,
                                   "frac": BYTES_HCZ * cells_rank / (cs_ms * 1e-3) / 1e9 / peak}}

    launches_per_step = solver.launches_per_step
    exch = None
    if world > 1 and getattr(solver, "exchanger", None) is not None:
        ex = solver.exchanger
        done = max(1, solver.steps_done)
        exch = {"exchanges_per_step": ex.calls / done, "bytes_per_direction_per_step": ex.bytes_per_direction / done,
                "transport": getattr(ex, "transport", "nccl p2p (batch_isend_irecv)")}
        if hasattr(ex, "pushes"):     # sweeps whose halo exchange was fused into the boundary-plane kernels
            exch["fused_pushes_per_step"] = ex.pushes / done
    # ---- end to end through host buffers
    e2e = None
    if not args.no_e2e:
        faces = [torch.empty(t.shape, dtype=t.dtype).pin_memory() for t in solver.H_ext_mac]
        for h, d in zip(faces, solver.H_ext_mac):
            h.copy_(d)
        phi_host = torch.empty(solver.phi.shape, dtype=torch.float32).pin_memory()
        def run(n):
            for _ in range(n):
                solver.step_host(faces, phi_host)
            if hasattr(solver, "wait_host_copies"):
                solver.wait_host_copies()   # the timed region ends when the last phi has landed on the host
        run(min(2, args.warmup))
        ms_e2e = timed(run, args.steps)
        h2d = sum(t.numel() * 4 for t in faces)
        d2h = phi_host.numel() * 4
        e2e = {"value": cells * args.steps / (ms_e2e * 1e-3) / 1e6, "unit": "MLUPS", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / args.steps,
               "api": "FerroSolver.step_host(H_ext_mac faces from pinned host, phi to pinned host)"}
        del faces, phi_host

    # ---- tau == 1 scalar form of the sweeps (bit-identical psi), reported separately
    fast = None
    del solver
    torch.cuda.empty_cache()
    if not args.no_fast_path and path == "faithful":
        fs = build_solver(res, "scalar", device, world, rank, args.workload, **slab_kw)
        fs.step(args.warmup)
        ms_fast = timed(fs.step, args.steps)
        # the sweeps of a step run in psi-only form (psi read + written, rhs read: 12 B/cell/sweep, SURVEY 8d's own
        # 664 B/cell model), on one GPU and on slabs (whose per-sweep halo is then one psi plane per direction)
        own_bytes = BYTES_HCZ + N_SWEEPS * 12
        v = cells * args.steps / (ms_fast * 1e-3) / 1e6
        fast = {"magnetic_path": "scalar", "value": v, "unit": "MLUPS", "ms_per_step": ms_fast / args.steps,
                "model_bytes_per_cell": own_bytes,
                "frac_of_own_model_roofline": v / world / (peak * 1e9 / own_bytes / 1e6),
                "note": "tau == 1 collapse of the 30 sweeps; psi bit-identical to the faithful path; "
                        "never compared with the 4864 B/cell model"}
        del fs
        torch.cuda.empty_cache()

    # ---- the other BASELINE configurations, the CPU baseline and the eager reference (rank 0, N = 1 only)
    cfgs = cpu = eager = None
    if rank == 0 and world == 1:
        if not args.no_configs and args.workload == "rosensweig":
            cfgs = other_configs(device, peak)
        if not args.no_ref_eager:
            eager = reference_eager_on_gpu()
        if not args.no_cpu_baseline:
            cres = tuple(args.cpu_res)
            v, ms_cpu, cores, kind = cpu_rosensweig_mlups(cres, 3, 1)
            cpu = {"value": v, "unit": "MLUPS", "cores": cores, "kind": kind,
                   "sample": f"same workload at {cres[0]}x{cres[1]}x{cres[2]}, 1 warm-up + 3 timed steps, "
                             f"{ms_cpu:.0f} ms/step, torch CPU threads = {cores}, denormals flushed to zero"}

    if rank == 0:
        model_bytes = (2 + N_SWEEPS) * BYTES_D3Q19
        line = {
            "metric": METRIC, "value": mlups, "unit": "MLUPS", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": ("demo_3d_LBM_Rosensweig_instability D3Q19 HCZ+magnetic mg100, 30 sweeps/step"
                                    if args.workload == "rosensweig" else
                                    "demo_3d_LBM_two_droplets D3Q19 HCZ+magnetic (closed magnetic walls, kappa 0.5, "
                                    "H_ext 1), 30 sweeps/step, global grid "
                                    f"{res[0] * world}x{res[1]}x{res[2]}"),
                       "res_per_gpu": list(res), "cells": cells, "magnetic_path": path,
                       "parallelism": f"z-slab x{world}" if world > 1 else "single GPU",
                       "streaming": args.streaming if world > 1 else "ab", "state_gb_per_gpu": state_gb,
                       "l2": "state (tens of GB) far larger than the 126 MB L2; no flush needed"},
            "repetitions": {"n": len(reps_ms), "ms_per_rep": reps_ms, "reported": "median",
                            "steps_per_rep": args.steps},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "fast_path": fast,
            "gpu_launches": launches_per_step * args.steps * len(reps_ms), "clocks": clocks,
            "step_roofline": {"model_bytes_per_cell": model_bytes,
                              "mlups_at_peak": peak * 1e9 / model_bytes / 1e6,
                              "frac_of_model_roofline": mlups / world / (peak * 1e9 / model_bytes / 1e6)},
            "finite": finite,
        }
        if world > 1:
            line["slab_bit_identical"] = None if slab_verify is None else slab_verify["slab_bit_identical"]
            line["slab_verify"] = slab_verify
            line["exchange"] = exch
            line["numa_bound_cores"] = numa
        if cfgs is not None:
            line["configs"] = cfgs
        if eager is not None:
            line["reference_eager_b200"] = eager
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_JSON_FD = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line: whatever else writes to fd 1 during the run (NCCL's version banner,
    library chatter) is sent to stderr instead, and the line itself goes to the saved descriptor."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def main():
    args = parse()
    _claim_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
