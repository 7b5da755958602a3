#!/usr/bin/env python
"""torchrun entry point of the N-rank data-plane check (lbm_ferrofluid_b200.distributed.verify_against_single_gpu):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/verify_slab.py [--res 64 48 56] [--steps 3]

NCCL on GPUs (one rank per GPU); ``--backend gloo --device cpu`` is refused: the solver has no CPU path.  Rank 0
prints one JSON line; exit status 1 if any owned cell of any field differs from the one-GPU run."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, nargs=3, default=[64, 48, 56])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--n-sweeps", type=int, default=30)
    a = ap.parse_args()
    from lbm_ferrofluid_b200.distributed import verify_against_single_gpu

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    if world < 2:
        raise SystemExit("verify_slab.py needs WORLD_SIZE >= 2 (launch with torchrun)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("TORCH_NCCL_HIGH_PRIORITY", "1")
    dist.init_process_group("nccl", device_id=dev)
    rep = verify_against_single_gpu(dev, world, rank, tuple(a.res), a.steps, n_sweeps=a.n_sweeps)
    if rank == 0:
        print(json.dumps(rep))
    dist.destroy_process_group()
    sys.exit(0 if rep["slab_bit_identical"] else 1)


if __name__ == "__main__":
    main()
