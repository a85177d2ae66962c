#!/usr/bin/env python
"""Standalone run of utils/fused_check over all ranks of a torchrun launch (degree = world size):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/fused_paths_bench.py [--small] [--out gpurun_out/fused_paths_nN.json]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.utils import fused_check  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--small", action="store_true")
ap.add_argument("--out", default=None)
a = ap.parse_args()
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
res = fused_check.run(dist.group.WORLD, dev, small=a.small)
if rank == 0:
    print(json.dumps(res), flush=True)
    if a.out:
        os.makedirs(os.path.dirname(a.out), exist_ok=True)
        json.dump(res, open(a.out, "w"), indent=1)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
