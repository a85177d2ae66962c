#!/usr/bin/env python
"""BASELINE.json config 5: GPipe vs 1F1B vs Interleaved-1F1B at WORLD_SIZE pipeline stages, micro-batches 4..64,
measured bubble vs the theoretical (p-1)/(v*m+p-1), tokens/s as max over ranks.  One torchrun launch sweeps all
configs in-process.  Writes gpurun_out/schedule_sweep_p{P}.json."""
import argparse
import gc
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.models import preset  # noqa: E402
from ppb200.parallel.mesh import init_mesh  # noqa: E402
from ppb200.parallel.partition import build_stage_model  # noqa: E402
from ppb200.parallel.runner import make_schedule  # noqa: E402
from ppb200.parallel.stage import PipelineStage  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--microbatches", default="4,8,16,32,64")
ap.add_argument("--schedules", default="GPipe,1F1B,Interleaved1F1B,ZeroBubble1F1B,ZeroBubbleInterleaved")
ap.add_argument("--partition", default="balanced", choices=["reference", "balanced", "fine"])
ap.add_argument("--global-batch", type=int, default=128)
ap.add_argument("--seq-len", type=int, default=1024)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--model", default="ref-gpt2-medium")
a = ap.parse_args()

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
from ppb200.parallel.mesh import configure_nccl_for_pipeline  # noqa: E402
configure_nccl_for_pipeline()
dist.init_process_group("nccl")
mesh = init_mesh(dp_size=1, pp_size=world)
margs = preset(a.model, dtype=torch.bfloat16)
rows = []
for sched_name in a.schedules.split(","):
    v = 1
    if "Interleaved" in sched_name:
        for cand in (2, 3, 4):
            if margs.n_layers % (world * cand) == 0:
                v = cand
                break
    for m in [int(x) for x in a.microbatches.split(",")]:
        if sched_name != "GPipe" and m < world:
            continue
        if "Interleaved" in sched_name and m % max(1, m // world) != 0:
            continue
        if a.global_batch % m != 0:
            continue
        try:
            stages = []
            for i in range(v):
                sidx = rank + world * i
                model = build_stage_model(margs, sidx, world * v, rule=a.partition, seq_len=a.seq_len, device=dev, seed=1)
                stages.append(PipelineStage(model, sidx, world * v, dev))
            sched = make_schedule(sched_name, stages, m, None, pp_group=mesh.pp_group, pp_group_bwd=mesh.pp_group_bwd,
                                  optimizer={"lr": 1e-4})
            first, last = 0 in sched.local, (world * v - 1) in sched.local
            g = torch.Generator().manual_seed(5)
            x = torch.randint(0, margs.vocab_size, (a.global_batch, a.seq_len), generator=g).to(dev)
            y = torch.randint(0, margs.vocab_size, (a.global_batch, a.seq_len), generator=g).to(dev)

            def step():
                sched.step(*((x,) if first else ()), target=y if last else None, losses=[] if last else None)

            for _ in range(4):
                step()
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                step()
            e1.record()
            dist.barrier(device_ids=[local])
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / a.steps, sched.measured().get("bubble_fraction", 0.0)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            mem = torch.tensor([torch.cuda.max_memory_allocated() / 2**30], device=dev)
            dist.all_reduce(mem, op=dist.ReduceOp.MAX)
            row = {"schedule": sched_name, "stages": world, "virtual": v, "n_microbatches": m, "ms_per_step": round(float(t[0]), 2),
                   "tokens_per_s": round(a.global_batch * a.seq_len / float(t[0]) * 1e3), "bubble_measured_max": round(float(t[1]), 4),
                   "bubble_theory": round(sched.theory_bubble, 4), "peak_mem_gib": round(float(mem[0]), 1)}
        except Exception as e:  # noqa: BLE001
            row = {"schedule": sched_name, "stages": world, "virtual": v, "n_microbatches": m, "error": str(e)[:200]}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        del stages, sched
        gc.collect()
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"schedule_sweep_p{world}.json"), "w"), indent=1)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
