#!/usr/bin/env python3
"""Bare N-way device-to-host bandwidth into ONE shared pinned host buffer (the destination of the
multi-GPU e2e path): the ceiling that path lives under.  Launch with torchrun:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/d2h_probe.py

Rank 0 prints one JSON line (bench.d2h_probe)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from spmm_b200 import multi  # noqa: E402


def main():
    world, rank, local = bench._env()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    multi._ARENA.ensure(world << 30, rank, register=True)
    res = bench.d2h_probe(torch, dist, dev, world, rank, multi)
    if rank == 0:
        res.update({"n_gpus": world, "build": bench.build_info()})
        print(json.dumps(res))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
