"""z-slab domain decomposition of the fused 3-D step: one process per GPU, NCCL point-to-point halo
exchange over NVLink, boundary planes first so the exchange overlaps the interior compute.

The reference is single-device ("Parallelism": none), so there is no reference
interface here; the contract is: a P-slab run equals the one-GPU FerroSolver run BIT FOR BIT on every owned
cell.  That holds by construction -- the slab entries of the C ABI (include/lbmff.h, "slab tier") run the
same kernels with the same per-cell arithmetic, position-dependent rules use the global z, and halo planes
are copies of the neighbour's owned planes.

Layout of one slab (z is the slowest spatial axis, so one z plane of one channel is a contiguous H*W block):

    local plane   0 .. halo-1 | halo .. halo+Dl-1 | halo+Dl .. Dl+2*halo-1
                  lower halo  |      owned        |      upper halo
    global plane  (z + zoff) mod Dg,   zoff = rank*Dl - halo          (the global domain is periodic in z,
                  so rank 0 and rank P-1 are neighbours too: LBM_propagation_3d.py:31-37)

``SlabFerroSolver.step`` is written as a generator of halo requests so that the same code runs under
``torch.distributed`` (DistExchanger: NCCL on GPUs, gloo in the CPU tests) and as P "virtual ranks" inside
one process on one GPU (LoopbackExchanger: the bit-for-bit test).
"""
from __future__ import annotations

import ctypes
import math
import os
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch

from . import _host as H
from ._lib import Slab
from .types import CellType

FLUID, OBSTACLE = int(CellType.FLUID), int(CellType.OBSTACLE)
HALO = 2  # rho needs two planes (kappa rho grad(lap(rho))); everything else one


# ---------------------------------------------------------------------------------------------------
# layout
# ---------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class SlabLayout:
    """Which planes of the global [Dg, H, W] domain one rank holds."""
    Dg: int
    world: int
    rank: int
    halo: int = HALO

    def __post_init__(self):
        if self.world < 2:
            raise ValueError("slab decomposition needs world >= 2 (use FerroSolver on one GPU)")
        if self.Dg % self.world:
            raise ValueError(f"global depth {self.Dg} is not divisible by world size {self.world}")
        if self.Dl < 2 * self.halo:
            raise ValueError(f"{self.Dl} owned planes per rank < 2*halo = {2 * self.halo}")

    @property
    def Dl(self) -> int:
        return self.Dg // self.world

    @property
    def D(self) -> int:
        return self.Dl + 2 * self.halo

    @property
    def z0(self) -> int:
        """global z of the first owned plane"""
        return self.rank * self.Dl

    @property
    def zoff(self) -> int:
        return self.z0 - self.halo

    @property
    def own(self) -> slice:
        return slice(self.halo, self.halo + self.Dl)

    @property
    def up(self) -> int:
        return (self.rank + 1) % self.world

    @property
    def down(self) -> int:
        return (self.rank - 1) % self.world

    def global_planes(self) -> torch.Tensor:
        return (torch.arange(self.D) + self.zoff) % self.Dg

    def take(self, t: torch.Tensor) -> torch.Tensor:
        """Local array (halos filled) of a global [B, C, Dg, H, W] tensor."""
        return t.index_select(2, self.global_planes().to(t.device)).contiguous()

    def owned(self, t: torch.Tensor) -> torch.Tensor:
        return t[:, :, self.own]

    def slab(self, zb: int, zn: int) -> Slab:
        return Slab(Dg=self.Dg, zoff=self.zoff, zb=zb, zn=zn)

    # plane ranges (zb, zn) used by the step
    def planes_owned(self):
        return self.halo, self.Dl

    def planes_depth(self, d: int):
        """owned planes plus d halo planes on each side"""
        return self.halo - d, self.Dl + 2 * d


def slab_shell_flags(layout: SlabLayout, H_: int, W: int, open_axes: Sequence[int] = (), device="cuda"):
    """shell_flags (initializers.py) of the global domain, restricted to one slab -- built without ever
    materialising the global array."""
    gz = layout.global_planes().to(device)
    wall_z = ((gz == 0) | (gz == layout.Dg - 1)).reshape(-1, 1, 1)
    y = torch.arange(H_, device=device).reshape(1, -1, 1)
    x = torch.arange(W, device=device).reshape(1, 1, -1)
    wall = torch.zeros((layout.D, H_, W), dtype=torch.bool, device=device)
    if 2 not in open_axes:
        wall |= wall_z
    if 1 not in open_axes:
        wall |= (y == 0) | (y == H_ - 1)
    if 0 not in open_axes:
        wall |= (x == 0) | (x == W - 1)
    fl = torch.where(wall, OBSTACLE, FLUID).to(torch.uint8)
    return fl.reshape(1, 1, layout.D, H_, W).contiguous()


# ---------------------------------------------------------------------------------------------------
# halo exchange
# ---------------------------------------------------------------------------------------------------
def _send_slices(layout: SlabLayout, depth: int):
    """(to_up, to_down, from_down, from_up) plane slices for a halo of `depth` planes."""
    h, Dl = layout.halo, layout.Dl
    return (slice(h + Dl - depth, h + Dl),   # my top owned planes -> upper neighbour's lower halo
            slice(h, h + depth),             # my bottom owned planes -> lower neighbour's upper halo
            slice(h - depth, h),             # my lower halo  <- lower neighbour's top owned planes
            slice(h + Dl, h + Dl + depth))   # my upper halo  <- upper neighbour's bottom owned planes


class _Pending:
    def __init__(self, finish: Callable[[], None]):
        self.finish = finish


Z_UP = (9, 10, 11, 12, 13)       # D3Q19 populations with e_z = +1: pulled from plane z-1, i.e. needed in a LOWER halo
Z_DOWN = (14, 15, 16, 17, 18)    # e_z = -1: pulled from plane z+1, needed in an UPPER halo


class DistExchanger:
    """Ring halo exchange with torch.distributed point-to-point ops (NCCL over NVLink between GPUs; gloo in the
    CPU tests), IN PLACE: a z-plane range of one channel is a contiguous block of the reference layout, so every
    message is sent from and received into the state tensors themselves -- no staging buffers, no pack / unpack
    kernels.  For population lattices only the five z-moving populations per direction travel (``pops=True``).  All
    messages of one call form ONE grouped NCCL operation.

    Two schedules:
      * ``start`` / ``finish``: the exchange stream waits for the work queued so far, the caller queues work that does
        not touch the halos, ``finish`` makes the caller's stream wait for the halos;
      * ``pipelined`` (the magnetic sweeps): boundary planes AND their exchange run on the high-priority exchange
        stream while the interior planes run on the caller's stream; the only cross-stream edges are
        boundary(s) <- interior(s-1) and interior(s) <- boundary(s-1), so neither the two small boundary launches nor
        the NVLink round trip ever sit on the interior kernels' critical path."""

    transport = "nccl p2p, in place (5 z-moving populations per direction), boundary planes + exchange on a side stream"

    def __init__(self, layout: SlabLayout, device, group=None):
        import torch.distributed as dist

        self.dist = dist
        self.layout = layout
        self.device = torch.device(device)
        self.group = group
        self.stream = torch.cuda.Stream(device=self.device, priority=-1) if self.device.type == "cuda" else None
        self.calls = 0
        self.bytes_per_direction = 0
        self._pipe = None          # (event after the last interior launch, event after the last boundary launch)

    def _run(self, tensors: Sequence[torch.Tensor], depth: int, pops: bool = False):
        lay, dist = self.layout, self.dist
        to_up, to_down, from_down, from_up = _send_slices(lay, depth)
        s_up, s_down, r_down, r_up = [], [], [], []
        for t in tensors:
            if pops and t.shape[1] != 19:
                raise RuntimeError("pops=True is for D3Q19 population lattices")
            for b in range(t.shape[0]):
                for c in (Z_UP if pops else range(t.shape[1])):       # what the upper neighbour's lower halo needs
                    s_up.append(t[b, c, to_up])
                    r_down.append(t[b, c, from_down])
                for c in (Z_DOWN if pops else range(t.shape[1])):     # what the lower neighbour's upper halo needs
                    s_down.append(t[b, c, to_down])
                    r_up.append(t[b, c, from_up])
        # matching order between a pair of ranks: (to_up -> from_down) first, then (to_down -> from_up); with two
        # ranks both neighbours are the same peer and the order is what pairs the messages
        ops = [dist.P2POp(dist.isend, v, lay.up, self.group) for v in s_up] + \
              [dist.P2POp(dist.isend, v, lay.down, self.group) for v in s_down] + \
              [dist.P2POp(dist.irecv, v, lay.down, self.group) for v in r_down] + \
              [dist.P2POp(dist.irecv, v, lay.up, self.group) for v in r_up]
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        self.calls += 1
        self.bytes_per_direction += 4 * sum(v.numel() for v in s_up)

    def start(self, tensors, depth, pops=False) -> _Pending:
        if self.stream is None:
            self._run(tensors, depth, pops)
            return _Pending(lambda: None)
        cur = torch.cuda.current_stream(self.device)
        self.flush()
        self.stream.wait_stream(cur)
        with torch.cuda.stream(self.stream):
            self._run(tensors, depth, pops)
        return _Pending(lambda: cur.wait_stream(self.stream))

    def exchange(self, tensors, depth, pops=False):
        self.start(tensors, depth, pops).finish()

    def pipelined(self, req: "HaloRequest"):
        """One sweep of the pipelined schedule: req.pre (boundary planes) and the exchange on the side stream,
        req.overlap (interior planes) on the caller's stream."""
        if self.stream is None:                    # CPU (gloo) tests: plain order
            req.pre()
            self._run(req.tensors, req.depth, req.pops)
            req.overlap()
            return
        cur, side = torch.cuda.current_stream(self.device), self.stream
        if self._pipe is None:
            side.wait_stream(cur)                  # first sweep: everything queued so far
        else:
            side.wait_event(self._pipe[0])         # boundary(s) reads interior(s-1) and overwrites what it read
        with torch.cuda.stream(side):
            req.pre()
            e_bnd = torch.cuda.Event()
            e_bnd.record(side)
            self._run(req.tensors, req.depth, req.pops)
        if self._pipe is not None:
            cur.wait_event(self._pipe[1])          # interior(s) reads boundary(s-1) and overwrites what it read
        req.overlap()
        e_int = torch.cuda.Event()
        e_int.record(cur)
        self._pipe = (e_int, e_bnd)

    def flush(self):
        """End of a pipelined section: the caller's stream waits for the last boundary launch and its halos."""
        if self._pipe is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            self._pipe = None


class SymmPushExchanger(DistExchanger):
    """DistExchanger whose per-sweep halo exchange of the magnetic lattice is FUSED into the boundary-plane sweep
    kernels: the h buffers live in torch symmetric memory (every rank maps its neighbours' buffers over NVLink), the
    kernel that sweeps a boundary plane stores the five z-moving populations into the neighbour's halo plane as it
    stores its own (lbmff_slab_magnetic_sweep_push_d3q19), and a signal-pad handshake replaces the message: 30 x per
    step there is no NCCL kernel, no copy, no staging.  Everything else (the once-per-step halos of f, g, rho, ...)
    stays on DistExchanger's grouped NCCL point-to-point path.

    Ordering.  The neighbour may overwrite my halo plane of buffer X in its sweep s only after my boundary kernels of
    sweep s-1 (the last readers of that plane) are done: it starts sweep s only after it has seen my signal of sweep
    s-1, which I raise behind those kernels in stream order -- the data dependence already gives the WAR order.  Two
    signal channels (0: to the upper neighbour, 1: to the lower one) keep the two messages between the same pair of
    ranks apart when world == 2."""

    transport = ("fused boundary-plane sweep + halo push over NVLink peer memory (torch symmetric memory, signal pads) "
                 "for the 30 magnetic sweeps; nccl p2p, in place, for the once-per-step halos")
    SIGNAL_TIMEOUT_MS = 60000

    def __init__(self, layout: SlabLayout, device, group=None):
        super().__init__(layout, device, group)
        import torch.distributed._symmetric_memory as symm

        self._symm = symm
        self._peers = {}          # data_ptr of a local symmetric tensor -> (handle, ptr on `up`, ptr on `down`)
        self._handle = None
        self._signals_pending = False
        self.pushes = 0

    def alloc(self, shape) -> torch.Tensor:
        """Zero-filled fp32 tensor in symmetric memory, mapped on both neighbours (collective: every rank calls it
        with the same shape, in the same order)."""
        t = self._symm.empty(tuple(shape), dtype=torch.float32, device=self.device)
        t.zero_()
        hdl = self._symm.rendezvous(t, self.group if self.group is not None else self.dist.group.WORLD)
        lay = self.layout
        up = hdl.get_buffer(lay.up, tuple(shape), torch.float32)
        down = hdl.get_buffer(lay.down, tuple(shape), torch.float32)
        self._peers[t.data_ptr()] = (hdl, up, down)     # the views keep the mappings alive
        if self._handle is None:
            self._handle = hdl
        return t

    def _wait_signals(self):
        if self._signals_pending:
            lay, h = self.layout, self._handle
            h.wait_signal(lay.down, 0, self.SIGNAL_TIMEOUT_MS)   # the lower neighbour's "to up" message
            h.wait_signal(lay.up, 1, self.SIGNAL_TIMEOUT_MS)     # the upper neighbour's "to down" message
            self._signals_pending = False

    def pipelined(self, req: "HaloRequest"):
        ent = None
        if req.push is not None and len(req.tensors) == 1 and self.stream is not None:
            ent = self._peers.get(req.tensors[0].data_ptr())   # the solver's views start at element 0 of the buffer
        if ent is None:
            with torch.cuda.stream(self.stream) if self.stream is not None else _nullcontext():
                self._wait_signals()
            return super().pipelined(req)
        _, up, down = ent
        off = 0
        cur, side = torch.cuda.current_stream(self.device), self.stream
        if self._pipe is None:
            side.wait_stream(cur)
        else:
            side.wait_event(self._pipe[0])         # boundary(s) reads interior(s-1) and overwrites what it read
        with torch.cuda.stream(side):
            self._wait_signals()                   # my halos of sweep s-1 have landed
            req.push(up.data_ptr() + off, down.data_ptr() + off)
            e_bnd = torch.cuda.Event()
            e_bnd.record(side)
            self._handle.put_signal(self.layout.up, 0, self.SIGNAL_TIMEOUT_MS)
            self._handle.put_signal(self.layout.down, 1, self.SIGNAL_TIMEOUT_MS)
            self._signals_pending = True
        if self._pipe is not None:
            cur.wait_event(self._pipe[1])          # interior(s) reads boundary(s-1) and overwrites what it read
        req.overlap()
        e_int = torch.cuda.Event()
        e_int.record(cur)
        self._pipe = (e_int, e_bnd)
        self.calls += 1
        self.pushes += 1
        t = req.tensors[0]    # five z-moving populations of a D3Q19 lattice, or one plane of psi
        self.bytes_per_direction += 4 * (5 if t.shape[1] == 19 else t.shape[1]) * t.shape[0] * t.shape[3] * t.shape[4]

    def flush(self):
        if self.stream is not None and self._signals_pending:
            with torch.cuda.stream(self.stream):
                self._wait_signals()
        super().flush()


class _nullcontext:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


def make_exchanger(layout: SlabLayout, device, group=None):
    """The exchanger the benchmark and the hardware check use: the fused push exchanger on GPUs (LBMFF_SLAB_PUSH=0 or
    an unavailable symmetric-memory backend selects plain DistExchanger)."""
    dev = torch.device(device)
    if dev.type == "cuda" and os.environ.get("LBMFF_SLAB_PUSH", "1") != "0":
        try:
            return SymmPushExchanger(layout, dev, group)
        except Exception as e:   # the NCCL path is complete on its own
            import warnings
            warnings.warn(f"SymmPushExchanger unavailable ({type(e).__name__}: {e}); using NCCL point-to-point")
    return DistExchanger(layout, dev, group)


def loopback_exchange(layouts: Sequence[SlabLayout], tensor_sets: Sequence[Sequence[torch.Tensor]], depth: int,
                      pops: bool = False):
    """Halo exchange between P virtual ranks living in one process (tensor_sets[r] = rank r's tensors); with
    ``pops`` exactly the channels DistExchanger sends, so a population the real exchange leaves stale is stale here."""
    P = len(layouts)
    up_ch, down_ch = (list(Z_UP), list(Z_DOWN)) if pops else (slice(None), slice(None))
    for r in range(P):
        lay = layouts[r]
        to_up, to_down, from_down, from_up = _send_slices(lay, depth)
        for mine, below, above in zip(tensor_sets[r], tensor_sets[lay.down], tensor_sets[lay.up]):
            mine[:, up_ch, from_down] = below[:, up_ch, to_up]
            mine[:, down_ch, from_up] = above[:, down_ch, to_down]


@dataclass
class HaloRequest:
    """Yielded by SlabFerroSolver._step_gen: fill `depth` halo planes of `tensors`; `overlap` (optional)
    queues work that does not touch those halos and may run while they travel.  ``pops``: the tensors are D3Q19
    population lattices whose halos are only ever pulled from, so the five z-moving populations per direction
    suffice.  ``pre`` (with ``pipelined``): the boundary-plane launches that produce the planes to be sent; the
    exchanger may run them, and the exchange, on its own stream (DistExchanger.pipelined)."""
    tensors: List[torch.Tensor]
    depth: int
    overlap: Optional[Callable[[], None]] = None
    pops: bool = False
    pre: Optional[Callable[[], None]] = None
    pipelined: bool = False
    # fused alternative to `pre` + exchange: push(ptr_up, ptr_down) launches the boundary planes with kernels that store
    # the z-moving populations straight into the neighbours' arrays (raw device pointers of the SAME tensor on the
    # upper / lower neighbour, mapped into this process); only exchangers that own such mappings use it
    push: Optional[Callable[[int, int], None]] = None


# ---------------------------------------------------------------------------------------------------
# the slab solver
# ---------------------------------------------------------------------------------------------------
class SlabFerroSolver:
    """One z-slab of FerroSolver's 3-D HCZ (+ magnetic) step.  Same constructor vocabulary as FerroSolver;
    ``flags`` / ``magnetic_flags`` are the slab-LOCAL arrays [1, 1, D, H, W] (halo planes included, e.g. from
    slab_shell_flags or SlabLayout.take), H_ext is a uniform vector (all the reference's demos use one:
    demo_3d_LBM_Rosensweig_instability.py:147-148)."""

    def __init__(self, layout: SlabLayout, flags: torch.Tensor, *, exchanger=None, density_liquid=0.2508,
                 density_gas=0.02381, rho_liquid=0.2508, rho_gas=0.02381, kappa=0.01, tau_f=0.68, tau_g=0.68,
                 tau=1.0, contact_angle=0.5 * math.pi, gravity=1e-4, dx=1.0, dt=1.0,
                 magnetic_flags: Optional[torch.Tensor] = None, H_ext=None, n_sweeps: int = 30,
                 magnetic_path: str = "faithful", keep_force: bool = False, streaming: str = "ab", chunk: int = 32):
        self.layout = layout
        self.flags = H.need_cuda(flags, "flags", torch.uint8)
        self.device = self.flags.device
        dim, self.B, _, self.D, self.Hh, self.W = H.dims(self.flags)
        if dim != 3 or self.D != layout.D:
            raise RuntimeError(f"flags must be the slab-local [B,1,{layout.D},H,W] array")
        if layout.Dl < 4:
            raise RuntimeError("need at least 4 owned planes per slab")
        if ((self.flags != FLUID) & (self.flags != OBSTACLE)).any().item():
            raise RuntimeError("SlabFerroSolver supports FLUID / OBSTACLE flags only")
        self.dim, self.Q = 3, 19
        self.res = (self.D, self.Hh, self.W)
        self.dx, self.dt = float(dx), float(dt)
        self._density_liquid, self._density_gas = density_liquid, density_gas
        self._rho_liquid, self._rho_gas = rho_liquid, rho_gas
        self._kappa, self._tau_f, self._tau_g, self._tau = kappa, tau_f, tau_g, tau
        self.gravity = float(gravity)
        self.cot = H.cot_of_contact_angle(contact_angle)
        self.n_sweeps = int(n_sweeps)
        self.params = H.fluid_params(self, self.dx, self.dt, cot=self.cot, gravity=self.gravity)
        self.exchanger = exchanger
        opts = dict(dtype=torch.float32, device=self.device)
        # ---- streaming mode.  "ab": two buffers per lattice.  "compressed" (the compressed-grid scheme of Pohl et
        # al. 2003, in chunks): ONE buffer per lattice with S = chunk + 1 spare planes; a pass over the slab is issued
        # as launches of `chunk` planes in ascending z whose output goes to the same buffer displaced by -S planes (a
        # launch reads planes [a-1, a+chunk] and writes [a-S, a+chunk-S): nothing it or a later launch still needs),
        # the next step runs descending with +S.  Same kernels, same per-cell arithmetic and access pattern (so the
        # same roofline), (D + S) / 2D of the A/B memory -- what puts 1024^3 on two 180 GB GPUs.
        if streaming not in ("ab", "compressed"):
            raise ValueError("streaming must be 'ab' or 'compressed'")
        self.streaming, self.chunk = streaming, int(chunk)
        self.S = self.chunk + 1 if streaming == "compressed" else 0
        if streaming == "compressed" and (self.chunk < 1 or self.D + self.S > layout.Dg):
            raise RuntimeError(f"compressed streaming: chunk + 1 spare planes on top of the {self.D} local planes must "
                               f"not exceed the global depth {layout.Dg}")
        self.Dk = self.D + self.S       # planes per channel in memory = the channel stride the kernels are given
        sc = lambda c: torch.zeros((self.B, c, *self.res), **opts)  # noqa: E731
        # multi-channel arrays share the kernels' channel stride (Dk planes); the solver works on views of D planes
        mc = lambda c: torch.zeros((self.B, c, self.Dk, self.Hh, self.W), **opts)  # noqa: E731
        self._mc = mc
        self.rho, self.pressure, self.vel = sc(1), sc(1), mc(3)[:, :, :self.D]
        self.force = mc(3)[:, :, :self.D] if keep_force else None
        self._ws = mc(3)[:, :, :self.D]
        if streaming == "ab":
            self._f, self._g = [mc(19), mc(19)], [mc(19), mc(19)]
        else:
            self._f, self._g = [mc(19)], [mc(19)]
        self._cur = 0
        self._off = self.S              # compressed: plane offset of the current population view (S: home, 0: displaced)
        self.steps_done = 0
        self._timing = None
        self._hook_before_source = self._hook_after_source = self._hook_after_rho_final = None
        self._down_stream = self._up_stream = self._faces_read = None
        self.magnetic = magnetic_flags is not None
        if self.magnetic:
            self.mflags = H.need_cuda(magnetic_flags, "magnetic_flags", torch.uint8)
            if H_ext is None or torch.is_tensor(H_ext):
                raise NotImplementedError("slab mode takes a uniform applied field (3 floats)")
            vec = [float(v) for v in H_ext]
            self._H_uniform = (ctypes.c_double * 3)(*vec)
            D, Hh, W = self.res
            self.H_ext_mac = [torch.full((self.B, 1, D, Hh, W + 1), vec[0], **opts),
                              torch.full((self.B, 1, D, Hh + 1, W), vec[1], **opts),
                              torch.full((self.B, 1, D + 1, Hh, W), vec[2], **opts)]
            # same validity rule as FerroSolver: the scalar sweep kernels test neighbour flags only next to the
            # GLOBAL outer shell, so every OBSTACLE cell of magnetic_flags must lie on it (global z = 0 / Dg-1, or
            # the x / y rim), and only FLUID / OBSTACLE may occur
            gz = layout.global_planes().to(self.device)
            z_inner = ((gz != 0) & (gz != layout.Dg - 1)).reshape(1, 1, -1, 1, 1)
            inner = self.mflags[:, :, :, 1:-1, 1:-1]
            walls_on_shell = not ((inner == OBSTACLE) & z_inner).any().item()
            if ((self.mflags != FLUID) & (self.mflags != OBSTACLE)).any().item():
                raise RuntimeError("SlabFerroSolver supports FLUID / OBSTACLE magnetic_flags only")
            scalar_ok = float(tau) == 1.0 and walls_on_shell
            if magnetic_path == "auto":
                magnetic_path = "scalar" if scalar_ok else "faithful"
            if magnetic_path not in ("scalar", "faithful"):
                raise ValueError("magnetic_path must be 'auto', 'faithful' or 'scalar'")
            if magnetic_path == "scalar" and not scalar_ok:
                raise RuntimeError("magnetic_path='scalar' needs tau == 1 and every OBSTACLE cell of magnetic_flags "
                                   "on the outer shell of the global domain")
            self.magnetic_path = magnetic_path
            self.psi, self._rhs, self._chi = sc(1), sc(1), sc(1)
            # the magnetic lattice stays A/B in both modes (its sweeps alternate boundary-first on two streams)
            alloc = getattr(exchanger, "alloc", None)
            self._h = None
            if alloc is not None:
                # buffers the neighbours can store into (SymmPushExchanger): same shape and views as mc(19) / mc(2)
                try:
                    nch = 19 if magnetic_path == "faithful" else 2
                    self._h = [alloc((self.B, nch, self.Dk, self.Hh, self.W))[:, :, :self.D] for _ in range(2)]
                except Exception as e:   # no symmetric memory here: ordinary buffers, the exchanger's NCCL path
                    import warnings
                    warnings.warn(f"symmetric-memory allocation failed ({type(e).__name__}: {e}); "
                                  "the magnetic sweeps exchange their halos over NCCL")
                    self._h = None
            if self._h is None:
                self._h = [mc(19)[:, :, :self.D] for _ in range(2)] if magnetic_path == "faithful" else \
                          [mc(2)[:, :, :self.D] for _ in range(2)]
            self._mcur = 0
        else:
            self.magnetic_path, self.mflags = None, None

    # ---- state (slab-local views; *_owned drop the halo planes) -----------------------------------
    def _pop(self, bufs, nxt=False):
        """View (D planes) of the current (or, nxt, the next) state of a hydrodynamic lattice."""
        if self.streaming == "ab":
            return bufs[(1 - self._cur) if nxt else self._cur][:, :, :self.D]
        off = (self.S - self._off) if nxt else self._off
        return bufs[0][:, :, off:off + self.D]

    @property
    def f(self):
        return self._pop(self._f)

    @property
    def g(self):
        return self._pop(self._g)

    @property
    def state_bytes(self) -> int:
        """Device bytes held by this slab's state arrays."""
        seen, total = set(), 0
        for v in vars(self).values():
            for t in (v if isinstance(v, (list, tuple)) else [v]):
                if torch.is_tensor(t) and t.is_cuda:
                    st = t.untyped_storage()
                    if st.data_ptr() not in seen:
                        seen.add(st.data_ptr())
                        total += st.nbytes()
        return total

    @property
    def h(self):
        return self._h[self._mcur] if self.magnetic and self.magnetic_path == "faithful" else None

    def owned(self, name: str) -> torch.Tensor:
        return self.layout.owned(getattr(self, name)).contiguous()

    def initialize(self, rho: torch.Tensor):
        """rho: slab-local [B,1,D,H,W] with valid halo planes.  p = EOS(rho), f = feq(rho, 0), g = geq,
        h = 0 (demo_3d_LBM_Rosensweig_instability.py:128-145) -- pointwise, so halos come out valid too."""
        from .collision import LBMCollision3d
        from .macro_compute import LBMMacroCompute3d

        rho = H.need_cuda(rho, "rho")
        self.rho.copy_(rho)
        self.vel.zero_()
        kw = dict(density_liquid=self._density_liquid, density_gas=self._density_gas, rho_liquid=self._rho_liquid,
                  rho_gas=self._rho_gas, device=self.device)
        macro, col = LBMMacroCompute3d(**kw), LBMCollision3d(tau=1.0, **kw)
        self._cur, self._off = 0, self.S
        f, g = self.f, self.g
        # pointwise, so done in chunks of planes: no full-size [19, D, H, W] temporaries (at 1024^3 on two GPUs the
        # two of them would be 87 GB on top of the state)
        step = 16
        for z0 in range(0, self.D, step):
            sl = slice(z0, min(self.D, z0 + step))
            r = self.rho[:, :, sl].contiguous()
            zero = torch.zeros((self.B, 3, *r.shape[2:]), dtype=torch.float32, device=self.device)
            pr = macro.get_pressure(dx=self.dx, dt=self.dt, density=r)
            self.pressure[:, :, sl] = pr
            f0 = col.get_feq_(dx=self.dx, dt=self.dt, rho=r, vel=zero, force=zero)
            g[:, :, sl] = col.get_geq_(dx=self.dx, dt=self.dt, rho=r, vel=zero, density=r, pressure=pr, force=zero,
                                       feq=f0)
            f[:, :, sl] = f0
            del f0, pr, r, zero
        if self.magnetic:
            for t in self._h:
                t.zero_()
            self.psi.zero_()
            self._mcur = 0
        self.steps_done = 0
        return self

    # ---- timing hooks (same contract as FerroSolver) ----------------------------------------------
    def enable_kernel_timing(self, on: bool = True):
        self._timing = {"sweeps": [], "collide_stream": []} if on else None

    def kernel_times_ms(self):
        return {k: [a.elapsed_time(b) for a, b in v] for k, v in (self._timing or {}).items()}

    def _mark(self):
        if self._timing is None:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    # ---- launches ---------------------------------------------------------------------------------
    def _launch(self, name, *args, planes):
        zb, zn = planes
        if zn <= 0:
            return
        slab = self.layout.slab(zb, zn)
        H.launch(name, self.device, *args, ctypes.byref(slab), self.B, self.Dk, self.Hh, self.W)

    def _split(self):
        """(interior planes, [boundary plane ranges]) of the owned block"""
        h, Dl = self.layout.halo, self.layout.Dl
        return (h + 1, Dl - 2), [(h, 1), (h + Dl - 1, 1)]

    def _step_gen(self):
        L, p = self.layout, H.ptr
        own, d1 = L.planes_owned(), L.planes_depth(1)
        interior, edges = self._split()
        f, g = self.f, self.g
        fo, go = self._pop(self._f, nxt=True), self._pop(self._g, nxt=True)
        compressed = self.streaming == "compressed"
        ascending = self._off == self.S          # compressed: this pass moves the populations by -S planes
        chi = self._chi if self.magnetic else None

        def moments(planes):
            self._launch("lbmff_slab_hcz_moments_d3q19", p(f), p(self.flags), p(self.rho), p(chi), self.params,
                         planes=planes)

        # f, g halos travel while the interior moments are taken
        yield HaloRequest([f, g], 1, overlap=lambda: moments(interior), pops=True)
        for e in edges:
            moments(e)
        yield HaloRequest([self.rho, self.pressure] + ([chi] if self.magnetic else []), L.halo)
        if self.magnetic:
            if self._hook_before_source is not None:
                self._hook_before_source()
            # psi-only form of the tau == 1 sweeps: the intermediate states are psi alone, so a
            # sweep moves 12 instead of 20 B per cell and its halo is ONE plane of psi per direction instead of two
            # state planes.  rhs is then needed one plane beyond the owned block (chi is valid two planes out).
            psi_only = (self.magnetic_path == "scalar" and self.n_sweeps >= 3 and self.W >= 4 and self.Hh >= 4
                        and os.environ.get("LBMFF_MAG_PSI", "1") != "0")
            self._launch("lbmff_slab_magnetic_source_3d", p(chi), p(self.mflags), *[p(t) for t in self.H_ext_mac],
                         self.dx, p(self._rhs), planes=d1 if psi_only else own)
            if self._hook_after_source is not None:
                self._hook_after_source()
            t0 = self._mark()
            sweep = ("lbmff_slab_magnetic_sweep_d3q19" if self.magnetic_path == "faithful"
                     else "lbmff_slab_magnetic_sweep_scalar_3d")
            wm = self._h[1 - self._mcur][:, 1:]   # psi-only: second half of the array the first sweep writes
            for s in range(self.n_sweeps):
                m = self._mcur
                src, dst = self._h[m], self._h[1 - m]
                psi = p(self.psi if s == self.n_sweeps - 1 else None)
                tail = (float(self._tau), self.dx, self.dt) if self.magnetic_path == "faithful" else (self.dx, self.dt)

                if psi_only:
                    modes = (0, 0, 0) if s == 0 else ((1, 0, 1) if s == 1 else (1, 1, 1))

                    def run(planes, src=src, dst=dst, modes=modes, peer=None, shift=0):
                        self._launch("lbmff_slab_magnetic_sweep_psi_3d", p(src), p(dst), p(self.mflags), p(self._rhs),
                                     p(wm), *modes, self.dx, self.dt, ctypes.c_void_p(peer or 0), shift, planes=planes)
                else:
                    def run(planes, src=src, dst=dst, psi=psi, tail=tail):
                        self._launch(sweep, p(src), p(dst), p(self.mflags), p(self._rhs), psi, *tail, planes=planes)

                def boundary(run=run):
                    for e in edges:
                        run(e)

                push = None
                if self.magnetic_path == "faithful" and self.B == 1:
                    # bottom owned plane -> the lower neighbour's upper halo (e_z = -1 populations, Dl planes up in ITS
                    # array); top owned plane -> the upper neighbour's lower halo (e_z = +1, Dl planes down)
                    def push(ptr_up, ptr_down, src=src, dst=dst, psi=psi, tail=tail):
                        cells = L.Dl * self.Hh * self.W
                        for e, ptr, direction, shift in ((edges[0], ptr_down, -1, cells), (edges[1], ptr_up, 1, -cells)):
                            self._launch("lbmff_slab_magnetic_sweep_push_d3q19", p(src), p(dst), p(self.mflags),
                                         p(self._rhs), psi, *tail, ctypes.c_void_p(ptr), direction, shift, planes=e)

                if psi_only and self.B == 1:     # the psi plane of each boundary plane goes straight to the neighbour
                    def push(ptr_up, ptr_down, run=run):
                        cells = L.Dl * self.Hh * self.W
                        run(edges[0], peer=ptr_down, shift=cells)
                        run(edges[1], peer=ptr_up, shift=-cells)

                self._mcur = 1 - m
                yield HaloRequest([dst[:, :1] if psi_only else dst], 1, overlap=lambda run=run: run(interior),
                                  pre=boundary, pipelined=True, pops=self.magnetic_path == "faithful", push=push)
            yield HaloRequest([], 0)      # join: the caller's stream waits for the last boundary planes and halos
            if psi_only:   # back to the two-float state (owned planes + the halo plane the next step's first sweep reads)
                self._launch("lbmff_slab_magnetic_psi_expand_3d", p(self._h[self._mcur]), p(self._h[1 - self._mcur]),
                             p(self.mflags), p(self._rhs), p(self.psi), self.dx, self.dt, planes=d1)
            t1 = self._mark()
            if self._timing is not None:
                self._timing["sweeps"].append((t0, t1))
            self._launch("lbmff_slab_shell_rule_3d", p(self.psi), p(self.mflags), planes=own)
            self._launch("lbmff_slab_shell_rule_3d", p(chi), p(self.flags), planes=d1)
        # ordered wetting boundary: x faces one plane beyond the owned block (the y faces read them)
        self._launch("lbmff_slab_hcz_wetting_3d", p(self.rho), p(self.flags), self.params, 1, planes=d1)
        self._launch("lbmff_slab_hcz_wetting_3d", p(self.rho), p(self.flags), self.params, 2, planes=own)
        if self._hook_after_rho_final is not None:   # owned planes of rho have their end-of-step value
            self._hook_after_rho_final()
        yield HaloRequest([self.rho], L.halo)
        self._launch("lbmff_slab_hcz_fields_3d", p(self.rho), p(self.flags), p(self.pressure), self.params,
                     p(self._ws), planes=d1)
        hu = ctypes.cast(self._H_uniform, ctypes.c_void_p) if self.magnetic else ctypes.c_void_p(0)

        def collide(planes):
            self._launch("lbmff_slab_hcz_collide_stream_d3q19", p(f), p(g), p(fo), p(go),
                         p(self.rho), p(self.flags), p(self.mflags), p(self._ws), p(chi),
                         p(self.psi if self.magnetic else None), hu, self.params, p(self.vel), p(self.pressure),
                         p(self.force), planes=planes)

        t0 = self._mark()
        if not compressed:
            if self.magnetic:   # psi halos travel while the interior planes collide
                yield HaloRequest([self.psi], 1, overlap=lambda: collide(interior))
            else:
                collide(interior)
            for e in edges:
                collide(e)
        else:
            # one buffer: the launches must follow the pass direction (see __init__) -- first edge plane, interior in
            # chunks, last edge plane -- so the psi halos are fetched before the first of them (one exposed exchange
            # per step instead of an overlapped one)
            if self.magnetic:
                yield HaloRequest([self.psi], 1)
            zb, zn = interior
            starts = list(range(zb, zb + zn, self.chunk))
            order = [edges[0]] + [(a, min(self.chunk, zb + zn - a)) for a in starts] + [edges[1]]
            for planes in (order if ascending else order[::-1]):
                collide(planes)
        t1 = self._mark()
        if self._timing is not None:
            self._timing["collide_stream"].append((t0, t1))
        if compressed:
            self._off = self.S - self._off
        else:
            self._cur = 1 - self._cur
        self.steps_done += 1

    def step(self, n: int = 1):
        """Advance n outer steps (asynchronous on torch's current stream; halos via self.exchanger)."""
        if self.exchanger is None:
            raise RuntimeError("SlabFerroSolver.step needs an exchanger (DistExchanger); virtual ranks use "
                               "step_virtual_ranks")
        ex = self.exchanger
        for _ in range(int(n)):
            for req in self._step_gen():
                if req.pipelined and hasattr(ex, "pipelined"):
                    ex.pipelined(req)
                    continue
                if not req.tensors:               # join point after a pipelined section
                    if hasattr(ex, "flush"):
                        ex.flush()
                    continue
                if req.pre is not None:
                    req.pre()
                pending = ex.start(req.tensors, req.depth, req.pops)
                if req.overlap is not None:
                    req.overlap()
                pending.finish()
            if hasattr(ex, "flush"):
                ex.flush()
        return self

    @property
    def phi(self):
        """Order parameter of the owned planes (demo_3d...Rosensweig...py:162)."""
        rho = self.layout.owned(self.rho)          # a contiguous block of planes (B = 1, one channel)
        if not rho.is_contiguous():
            rho = rho.contiguous()
        out = torch.empty_like(rho)
        H.launch("lbmff_order_parameter", self.device, H.ptr(rho), float(self._rho_gas), float(self._rho_liquid),
                 float(self._density_gas), float(self._density_liquid), H.ptr(out), rho.numel())
        return out

    def step_host(self, faces_host, phi_host):
        """One outer step through host buffers (same contract and copy/compute overlap as FerroSolver.step_host):
        this slab's applied-field faces come from pinned memory on an upload stream that waits only for the previous
        step's source kernel, phi of the owned planes goes to pinned memory on a download stream as soon as rho is
        final, both under the magnetic sweeps."""
        cur = torch.cuda.current_stream(self.device)
        if self._down_stream is None:
            self._down_stream = torch.cuda.Stream(device=self.device)
            self._up_stream = torch.cuda.Stream(device=self.device)
        ds, us = self._down_stream, self._up_stream
        up = None
        if faces_host is not None:
            if self._faces_read is not None:
                us.wait_event(self._faces_read)
            else:
                us.wait_stream(cur)
            with torch.cuda.stream(us):
                for dst, src in zip(self.H_ext_mac, faces_host):
                    dst.copy_(src, non_blocking=True)
                up = torch.cuda.Event()
                up.record(us)
        phi_done = [None]

        def after_source():
            self._faces_read = torch.cuda.Event()
            self._faces_read.record(cur)

        def download():
            if phi_host is None:
                return
            ev = torch.cuda.Event()
            ev.record(cur)
            with torch.cuda.stream(ds):
                ds.wait_event(ev)
                phi = self.phi
                phi_done[0] = torch.cuda.Event()
                phi_done[0].record(ds)
                phi_host.copy_(phi, non_blocking=True)
                phi.record_stream(ds)

        self._hook_before_source = (lambda: cur.wait_event(up)) if up is not None else None
        self._hook_after_source = after_source
        self._hook_after_rho_final = download
        try:
            self.step(1)
        finally:
            self._hook_before_source = self._hook_after_source = self._hook_after_rho_final = None
        if phi_done[0] is not None:
            cur.wait_event(phi_done[0])
        return self

    def wait_host_copies(self):
        if self._down_stream is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_stream(self._down_stream)
            cur.wait_stream(self._up_stream)

    @property
    def launches_per_step(self) -> int:
        # moments x3, wetting 5, fields, collide x3, source, sweeps x3 each, shell rule (rim + z-shell planes) x2
        return 3 + 5 + 1 + 3 + ((1 + 3 * self.n_sweeps + 4) if self.magnetic else 0)

    def total_mass(self) -> float:
        return float(self.layout.owned(self.f).double().sum().item())

    # ---- benchmark constructor --------------------------------------------------------------------
    @classmethod
    def rosensweig(cls, res_per_rank, path, device, world, rank, mag=100.0, n_sweeps=30, **kw):
        """Weak-scaling Rosensweig case: every rank owns res_per_rank = (Dl, H, W); global depth world*Dl."""
        from . import initializers as I

        Dl, Hh, W = res_per_rank
        lay = SlabLayout(Dl * world, world, rank)
        flags = slab_shell_flags(lay, Hh, W, device=device)
        mflags = slab_shell_flags(lay, Hh, W, open_axes=(1,), device=device)
        rho = I.rosensweig_rho((lay.D, Hh, W), flags, global_depth=lay.Dg, z_offset=lay.zoff)
        s = cls(lay, flags, exchanger=make_exchanger(lay, device), magnetic_flags=mflags, H_ext=[0.0, mag, 0.0],
                n_sweeps=n_sweeps, magnetic_path=path, contact_angle=0.5 * math.pi, gravity=1e-4, **kw)
        return s.initialize(rho)


def _two_droplets(cls, res_per_rank, path, device, world, rank, mag=1.0, n_sweeps=30, **kw):
    """demo_3d_LBM_two_droplets.py scaled up (BASELINE.json configs[4]): two liquid spheres in gas, every wall closed in
    both flag arrays (:113-121), kappa 0.5, tau_f = tau_g = 0.68, gravity 1e-4, uniform H_ext = mag along y (:166-168).
    Every rank owns res_per_rank = (Dl, H, W) of the global (world * Dl, H, W) box."""
    from . import initializers as I

    Dl, Hh, W = res_per_rank
    lay = SlabLayout(Dl * world, world, rank)
    flags = slab_shell_flags(lay, Hh, W, device=device)
    mflags = slab_shell_flags(lay, Hh, W, device=device)
    rho = I.two_droplets_rho((lay.D, Hh, W), flags, global_depth=lay.Dg, z_offset=lay.zoff)
    s = cls(lay, flags, exchanger=make_exchanger(lay, device), magnetic_flags=mflags, H_ext=[0.0, mag, 0.0],
            n_sweeps=n_sweeps, magnetic_path=path, contact_angle=0.5 * math.pi, gravity=1e-4, kappa=0.5, **kw)
    return s.initialize(rho)


SlabFerroSolver.two_droplets = classmethod(_two_droplets)


def step_virtual_ranks(solvers: Sequence[SlabFerroSolver], n: int = 1, push: bool = False):
    """Advance P slab solvers living in ONE process (one GPU) in lockstep, halos by device copies.
    This is the bit-for-bit test vehicle of the decomposition.  ``push``: requests that offer the
    fused boundary-sweep + halo-push kernels use them, the "peer" arrays being the neighbour virtual ranks' tensors
    (what SymmPushExchanger does across GPUs over symmetric memory)."""
    layouts = [s.layout for s in solvers]
    for _ in range(int(n)):
        gens = [s._step_gen() for s in solvers]
        while True:
            reqs = [next(g, None) for g in gens]
            if all(r is None for r in reqs):
                break
            if any(r is None for r in reqs):
                raise RuntimeError("virtual ranks fell out of lockstep")
            if not reqs[0].tensors:
                continue
            if push and all(r.push is not None for r in reqs):
                for k, r in enumerate(reqs):
                    r.push(reqs[layouts[k].up].tensors[0].data_ptr(), reqs[layouts[k].down].tensors[0].data_ptr())
                for r in reqs:
                    if r.overlap is not None:
                        r.overlap()
                continue
            for r in reqs:
                if r.pre is not None:
                    r.pre()
            loopback_exchange(layouts, [r.tensors for r in reqs], reqs[0].depth, reqs[0].pops)
            for r in reqs:
                if r.overlap is not None:
                    r.overlap()
    return solvers


# ---------------------------------------------------------------------------------------------------
# hardware check of the N-rank data plane ("P ranks == 1 rank, bit for bit")
# ---------------------------------------------------------------------------------------------------
def verify_against_single_gpu(device, world: int, rank: int, res_global=(64, 48, 56), steps: int = 3,
                              paths: Sequence[str] = ("faithful", "scalar"), n_sweeps: int = 30, group=None,
                              exchanger_factory=None) -> dict:
    """Run a small perturbed-interface Rosensweig case on ``world`` ranks through the REAL exchanger (NCCL between
    processes, boundary planes first, halos travelling on the side stream under the interior kernels), gather the
    owned planes and compare them BIT FOR BIT with a one-GPU FerroSolver run of the global domain on rank 0.
    Collective: every rank of the process group must call it.  Returns (on every rank) a dict with
    ``slab_bit_identical`` and, per magnetic path, the first differing field if any.  bench.py calls this before its
    timed region whenever WORLD_SIZE > 1; tools/verify_slab.py is the torchrun entry point."""
    import torch.distributed as dist

    from . import initializers as I
    from .solver import FerroSolver

    Dg, Hh, W = res_global
    dev = torch.device(device)
    # identical global initial condition on every rank (seeded noise comes from the CPU generator)
    flags = I.shell_flags(res_global, device=dev)
    mflags = I.shell_flags(res_global, open_axes=(1,), device=dev)
    rho = I.rosensweig_rho(res_global, flags, noise=1e-3, seed=11)
    kw = dict(H_ext=[0.0, 100.0, 0.0], n_sweeps=n_sweeps, contact_angle=0.55 * math.pi, gravity=1e-4)
    names = ("f", "g", "rho", "pressure", "vel", "psi")
    report = {"res_global": list(res_global), "steps": steps, "world": world, "n_sweeps": n_sweeps, "paths": {}}
    ok_all = True
    for path in paths:
        lay = SlabLayout(Dg, world, rank)
        ex = (exchanger_factory or make_exchanger)(lay, dev) if group is None else DistExchanger(lay, dev, group)
        slab = SlabFerroSolver(lay, lay.take(flags), exchanger=ex, magnetic_flags=lay.take(mflags),
                               magnetic_path=path, **kw).initialize(lay.take(rho))
        slab.step(steps)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
        mine = torch.cat([slab.owned(n).reshape(-1) for n in names] +
                         ([slab.owned("h").reshape(-1)] if path == "faithful" else []))
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        verdict = torch.zeros(1, dtype=torch.int32, device=dev)
        first_bad = None
        if rank == 0:
            one = FerroSolver(flags, magnetic_flags=mflags, magnetic_path=path, keep_force=False,
                              **kw).initialize(rho).step(steps)
            fluid = (mflags != OBSTACLE)
            chans = [getattr(one, n).shape[1] for n in names] + ([19] if path == "faithful" else [])
            fields = list(names) + (["h"] if path == "faithful" else [])
            got = {n: [] for n in fields}
            for part in parts:
                o = 0
                for n, c in zip(fields, chans):
                    cnt = c * lay.Dl * Hh * W
                    got[n].append(part[o:o + cnt].reshape(1, c, lay.Dl, Hh, W))
                    o += cnt
            for n in fields:
                a, b = torch.cat(got[n], dim=2), getattr(one, n)
                if n == "psi":      # wall-cell psi is never observed (the shell rule overwrites it)
                    a, b = a * fluid, b * fluid
                if not torch.equal(a, b):
                    first_bad = first_bad or n
            verdict[0] = 0 if first_bad is None else 1
        dist.broadcast(verdict, src=0, group=group)
        ok = int(verdict.item()) == 0
        ok_all &= ok
        report["paths"][path] = {"bit_identical": ok, "first_differing_field": first_bad,
                                 "exchanges": ex.calls, "bytes_per_direction": ex.bytes_per_direction}
        del slab, ex
    report["slab_bit_identical"] = bool(ok_all)
    return report
