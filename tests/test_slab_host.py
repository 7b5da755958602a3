"""Host side of the z-slab decomposition on CPU: layout arithmetic, slab-local flags, and the ring halo
exchange through torch.distributed with the gloo backend at world sizes 2 and 3 (the NCCL path on the GPUs
runs the same DistExchanger code)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lbm_ferrofluid_b200.distributed import (DistExchanger, SlabLayout, loopback_exchange, slab_shell_flags)


def test_layout_planes_and_ownership():
    lays = [SlabLayout(24, 3, r) for r in range(3)]
    assert [l.Dl for l in lays] == [8, 8, 8] and lays[0].D == 12
    assert lays[0].global_planes().tolist() == [22, 23] + list(range(0, 8)) + [8, 9]
    assert lays[2].global_planes().tolist() == [14, 15] + list(range(16, 24)) + [0, 1]
    g = torch.arange(24.0).reshape(1, 1, 24, 1, 1).expand(1, 2, 24, 3, 4)
    parts = [l.owned(l.take(g)) for l in lays]
    assert torch.equal(torch.cat(parts, dim=2), g)
    with pytest.raises(ValueError):
        SlabLayout(25, 2, 0)
    with pytest.raises(ValueError):
        SlabLayout(24, 1, 0)
    with pytest.raises(ValueError):
        SlabLayout(6, 2, 0)


def test_slab_flags_match_global_flags():
    from oracle.make_golden import shell_flags

    res = (12, 7, 9)
    for open_axes in ((), (1,), (2,)):
        glob = shell_flags(res, open_axes=open_axes)
        for r in range(3):
            lay = SlabLayout(res[0], 3, r)
            assert torch.equal(slab_shell_flags(lay, res[1], res[2], open_axes, device="cpu"), lay.take(glob))


def test_loopback_exchange_fills_halos():
    P, Dg = 3, 18
    lays = [SlabLayout(Dg, P, r) for r in range(P)]
    glob = torch.randn(1, 5, Dg, 4, 6)
    sets = []
    for l in lays:
        t = torch.zeros(1, 5, l.D, 4, 6)
        t[:, :, l.own] = l.owned(l.take(glob))
        sets.append([t])
    loopback_exchange(lays, sets, 2)
    for l, (t,) in zip(lays, sets):
        assert torch.equal(t, l.take(glob))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, Dg, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        glob_a = torch.randn(1, 19, Dg, 5, 7)
        glob_b = torch.randn(1, 1, Dg, 5, 7)
        lay = SlabLayout(Dg, world, rank)
        ex = DistExchanger(lay, "cpu")
        a, b = torch.full((1, 19, lay.D, 5, 7), -7.0), torch.full((1, 1, lay.D, 5, 7), -7.0)
        a[:, :, lay.own] = lay.owned(lay.take(glob_a))
        b[:, :, lay.own] = lay.owned(lay.take(glob_b))
        ex.exchange([a, b], 1)          # depth 1: the outer halo plane must stay untouched
        want_a = lay.take(glob_a)
        ok = torch.equal(a[:, :, 1:-1], want_a[:, :, 1:-1]) and bool((a[:, :, 0] == -7).all()) \
            and bool((a[:, :, -1] == -7).all())
        pend = ex.start([b], 2)         # depth 2 through the start/finish (overlap) interface
        pend.finish()
        ok = ok and torch.equal(b, lay.take(glob_b)) and ex.calls == 2
        ok = ok and ex.bytes_per_direction == 4 * (20 * 35 + 2 * 35)
        # population lattices: only the five z-moving populations per direction travel, in place
        from lbm_ferrofluid_b200.distributed import HaloRequest, Z_DOWN, Z_UP
        c = torch.full((1, 19, lay.D, 5, 7), -7.0)
        c[:, :, lay.own] = lay.owned(want_a)
        ex.exchange([c], 1, pops=True)
        h = lay.halo
        ok = ok and torch.equal(c[:, list(Z_UP), h - 1], want_a[:, list(Z_UP), h - 1])
        ok = ok and torch.equal(c[:, list(Z_DOWN), h + lay.Dl], want_a[:, list(Z_DOWN), h + lay.Dl])
        ok = ok and bool((c[:, list(Z_DOWN), h - 1] == -7).all()) and bool((c[:, :9, h - 1] == -7).all())
        ok = ok and ex.bytes_per_direction == 4 * (20 * 35 + 2 * 35 + 5 * 35)
        # the pipelined schedule (CPU: plain order): pre -> exchange -> overlap, then a join
        order = []
        d = torch.full((1, 1, lay.D, 5, 7), -7.0)

        def pre():
            d[:, :, lay.own] = lay.owned(lay.take(glob_b))
            order.append("pre")

        ex.pipelined(HaloRequest([d], 1, overlap=lambda: order.append("overlap"), pre=pre, pipelined=True))
        ex.flush()
        ok = ok and order == ["pre", "overlap"] and torch.equal(d[:, :, 1:-1], lay.take(glob_b)[:, :, 1:-1])
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_dist_exchanger_gloo(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as m:
        out = m.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, 8 * world, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert dict(out) == {r: True for r in range(world)}


def test_make_exchanger_picks_the_plain_exchanger_off_gpu(monkeypatch):
    """distributed.make_exchanger: the fused push exchanger needs CUDA + symmetric memory; on the CPU (the gloo tests) and
    with LBMFF_SLAB_PUSH=0 it must hand back the plain point-to-point DistExchanger.  HaloRequest.push stays optional."""
    from lbm_ferrofluid_b200.distributed import DistExchanger, HaloRequest, SlabLayout, SymmPushExchanger, make_exchanger

    lay = SlabLayout(16, 2, 0)
    ex = make_exchanger(lay, "cpu")
    assert type(ex) is DistExchanger and not hasattr(ex, "alloc")
    monkeypatch.setenv("LBMFF_SLAB_PUSH", "0")
    assert type(make_exchanger(lay, "cpu")) is DistExchanger
    assert issubclass(SymmPushExchanger, DistExchanger)
    assert HaloRequest([], 0).push is None
