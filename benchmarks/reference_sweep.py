#!/usr/bin/env python
"""The reference notebook's experiment grid (nb:345-394) on B200: layers x heads x schedules at WORLD_SIZE pipeline
ranks, batch 32 x seq 128, 4 micro-batches, dim 768, vocab 10000, dropout 0.1 -- the exact workload of
/root/reference/LLMsDistributedTrainingHelper.py:150-235 -- swept in-process under one torchrun launch.

Writes gpurun_out/reference_sweep_p{P}.json (rows with the notebook's columns + measured/theoretical bubble); merge
the P=2 and P=4 files with --report to get the notebook's speedup/efficiency table and its two figures.
"""
import argparse
import gc
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ap = argparse.ArgumentParser()
ap.add_argument("--layers", default="4,8,12")
ap.add_argument("--heads", default="4,8,12")
ap.add_argument("--schedules", default="GPipe,1F1B,Interleaved1F1B")
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--seq", type=int, default=128)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--no-optimizer", action="store_true", help="forward+backward only, like the reference")
ap.add_argument("--report", action="store_true", help="no GPU: merge gpurun_out/reference_sweep_p*.json into tables + plots")
a = ap.parse_args()

if a.report:
    import pandas as pd
    from ppb200.utils import analysis
    rows = []
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "reference_sweep_p*.json")) or
                    glob.glob(os.path.join(ROOT, "gpurun_out", "reference_sweep_p*.json"))):
        rows += [r for r in json.load(open(f)) if "error" not in r]
    df = pd.DataFrame(rows)
    sp = analysis.compute_speedup_and_efficiency(df)
    out = os.path.join(ROOT, "profiles")
    with open(os.path.join(out, "reference_sweep_report.md"), "w") as f:
        note = ""
        if "measured_in_round" in df.columns and df["measured_in_round"].nunique() > 1:
            r2 = df[df["measured_in_round"] == 2]
            note = ("Rows re-measured in round 2 (head dim 192 on the tcgen05 attention kernels; everything else of round 2 too): "
                    f"{len(r2)} cells -- num_processes {sorted(r2['num_processes'].unique().tolist())}, n_heads "
                    f"{sorted(r2['n_heads'].unique().tolist())}.  The other cells are the round-1 measurements (their 4-head rows "
                    "still ran attention on CUDA cores).\n\n")
        f.write("# Reference experiment grid on B200 (ours)\n\n" + note + "## Throughput (tokens/s)\n\n```\n" + analysis.throughput_pivot(df).round(0).to_string() +
                "\n```\n\n## Speedup vs GPipe / efficiency (nb:405-435)\n\n```\n" + sp.round(4).to_string(index=False) +
                "\n```\n\n## Measured vs theoretical bubble\n\n```\n" + analysis.bubble_table(df).round(4).to_string(index=False) + "\n```\n")
    analysis.plot_speedup_and_efficiency(sp, os.path.join(out, "reference_sweep_speedup.svg"))
    analysis.plot_throughput_grid(df, os.path.join(out, "reference_sweep_throughput.svg"))
    print("wrote profiles/reference_sweep_report.md and SVG figures;", len(df), "rows")
    sys.exit(0)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from ppb200.api import make_schedule, manual_model_split, run_train_iterations  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402
from ppb200.parallel.mesh import configure_nccl_for_pipeline, init_mesh  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
configure_nccl_for_pipeline()
dist.init_process_group("nccl")
mesh = init_mesh(dp_size=1, pp_size=world)
rows = []
for n_layers in [int(x) for x in a.layers.split(",")]:
    for n_heads in [int(x) for x in a.heads.split(",")]:
        for sched_name in a.schedules.split(","):
            v = 2 if (sched_name == "Interleaved1F1B" and n_layers % (world * 2) == 0) else 1   # helper:181-185
            row = {"n_layers": n_layers, "n_heads": n_heads, "num_processes": world, "schedule": sched_name, "virtual_stages": v}
            try:
                margs = ModelArgs(n_layers=n_layers, n_heads=n_heads, dtype=torch.bfloat16)
                stages = []
                for i in range(v):
                    torch.manual_seed(0)
                    st = manual_model_split(Transformer(margs), rank + world * i, world * v, dev)
                    stages.append(st)
                sched = make_schedule(sched_name, stages, 4, None, pp_group=mesh.pp_group, pp_group_bwd=mesh.pp_group_bwd,
                                      optimizer=None if a.no_optimizer else {"lr": 1e-4})
                x = torch.randint(0, margs.vocab_size, (a.batch, a.seq), device=dev)
                y = torch.randint(0, margs.vocab_size, (a.batch, a.seq), device=dev)
                res = run_train_iterations(sched, x, y, rank, world, a.iters, warmup=4)
                b = torch.tensor([res["bubble_fraction"]], device=dev)
                dist.all_reduce(b, op=dist.ReduceOp.MAX)
                row.update(throughput=res["throughput"], elapsed_time=res["elapsed_time"], tokens_processed=res["tokens_processed"],
                           step_time_ms=res["step_time_ms"], bubble_fraction=float(b.item()), bubble_theory=res["bubble_theory"],
                           n_microbatches=4)
            except Exception as e:  # noqa: BLE001  (same tolerance as nb:386-390)
                row["error"] = str(e)[:200]
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
            stages = sched = None
            gc.collect()
            torch.cuda.empty_cache()
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", f"reference_sweep_p{world}.json"), "w"), indent=1)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
