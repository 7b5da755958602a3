#!/usr/bin/env python
"""How much host<->device bandwidth do N ranks get when they copy at the same time?  (torchrun, one rank per GPU.)

Each rank binds to its GPU's NUMA node the way bench.py does, allocates pinned buffers of the e2e path's sizes
(1.6 GB up, 0.54 GB down) and times (a) its own copies with the other ranks idle, (b) the same copies with all ranks
copying at once.  Explains the e2e column at N = 8."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import bench  # noqa: E402


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    numa = bench.bind_to_gpu_numa_node(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    up_b, down_b = 1613758464, 536870912
    hu = torch.empty(up_b // 4, dtype=torch.float32).pin_memory()
    hd = torch.empty(down_b // 4, dtype=torch.float32).pin_memory()
    du = torch.empty(up_b // 4, dtype=torch.float32, device="cuda")
    dd = torch.empty(down_b // 4, dtype=torch.float32, device="cuda")
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()

    def both():
        with torch.cuda.stream(s_up):
            du.copy_(hu, non_blocking=True)
        with torch.cuda.stream(s_down):
            hd.copy_(dd, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s_up)
        torch.cuda.current_stream().wait_stream(s_down)

    alone = None
    for r in range(world):          # one rank at a time
        dist.barrier()
        if r == rank:
            alone = timed(both)
    dist.barrier()
    together = timed(both)          # all ranks at once
    dist.barrier()
    rows = [None] * world
    dist.all_gather_object(rows, dict(rank=rank, numa_cores=numa, alone_ms=alone, together_ms=together))
    if rank == 0:
        gb = (up_b + down_b) / 1e9
        out = dict(world=world, bytes_up=up_b, bytes_down=down_b,
                   per_rank=[dict(r, alone_gbs=gb / r["alone_ms"] * 1e3, together_gbs=gb / r["together_ms"] * 1e3)
                             for r in rows])
        out["aggregate_together_gbs"] = sum(r["together_gbs"] for r in out["per_rank"])
        out["slowest_together_ms"] = max(r["together_ms"] for r in rows)
        out["slowest_alone_ms"] = max(r["alone_ms"] for r in rows)
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
