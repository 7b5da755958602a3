#!/usr/bin/env python
"""Probe of torch.distributed._symmetric_memory on this box (torchrun, one rank per GPU): large allocations, peer views,
stores into the peer's buffer from a kernel, signal pads.  Prints one JSON line from rank 0."""
import json
import os
import time

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    gb = float(os.environ.get("SYMM_GB", "10"))
    n = int(gb * 2**30 // 4)
    t0 = time.time()
    bufs = [symm.empty(n, dtype=torch.float32, device=dev) for _ in range(2)]
    hd = [symm.rendezvous(b, dist.group.WORLD) for b in bufs]
    torch.cuda.synchronize()
    out["alloc_rendezvous_s"] = round(time.time() - t0, 2)
    up, down = (rank + 1) % world, (rank - 1) % world
    for b in bufs:
        b.fill_(float(rank))
    torch.cuda.synchronize()
    dist.barrier()
    peer_up = hd[0].get_buffer(up, (n,), torch.float32)
    # store a plane-sized block into the upper neighbour's buffer 0 with an ordinary elementwise kernel, then signal
    plane = 512 * 512
    side = torch.cuda.Stream()
    h = hd[0]
    times = []
    for it in range(20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(side):
            a.record(side)
            peer_up[it * plane:(it + 5) * plane].copy_(bufs[1][:5 * plane] + 100.0 + it)
            h.put_signal(up, 0)
            h.wait_signal(down, 0)
            b.record(side)
        side.synchronize()
        times.append(a.elapsed_time(b))
    dist.barrier()
    torch.cuda.synchronize()
    got = bufs[0][19 * plane:20 * plane + 1].cpu()
    out_r = {"rank": rank, "first": float(got[0]), "expect": float(down + 100.0 + 19), "ms_median": sorted(times)[10]}
    rows = [None] * world
    dist.all_gather_object(rows, out_r)
    out["ranks"] = rows
    out["ok"] = all(abs(r["first"] - r["expect"]) < 1e-6 for r in rows)
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
