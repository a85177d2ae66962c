#!/usr/bin/env python
"""Schedule sweep against the reference arm, all cells of one GPU count inside ONE torchrun launch (start-up is paid once):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        benchmarks/sweep_vs_reference.py --out gpurun_out/sweep_nN.json [--microbatches 4,8,16,32,64] [--steps 3]

One arm per launch (`--impl ours` / `--impl reference`; `benchmarks/sweep_merge.py` joins the two files into the table).
For every schedule the reference selects (GPipe, 1F1B, Interleaved1F1B -- helper:215-220, nb:346-349) and every
micro-batch count it runs `bench.run_ours` and `bench.run_reference` (the UNMODIFIED helper on NCCL + cuBLAS) with the
same flags on the flagship model (BASELINE.json config 5) and records tokens/s of both, their ratio, the measured bubble
(critical rank and max over ranks) and the theoretical bound.  Interleaved uses two virtual stages per rank in our arm; the
reference arm follows the helper's own rule (two chunks iff n_layers % (2 * ranks) == 0, helper:181-185, else one).
Cells a schedule cannot run (1F1B needs m >= stages, interleaved needs m % ranks == 0) are skipped by both arms."""
import argparse
import copy
import gc
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "sweep.json"))
    ap.add_argument("--microbatches", default="4,8,16,32,64")
    ap.add_argument("--schedules", default="GPipe,1F1B,Interleaved1F1B")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"],
                    help="one arm per launch: the process group is created once (lazily for ours, eagerly with device_id for the "
                         "reference arm, exactly as bench.py does) and shared by all cells; merge the two JSON files afterwards")
    sa = ap.parse_args()
    rank, local_rank, world = bench.env_rank()
    sys.argv = [sys.argv[0]]
    base = bench.parse()
    base.gpus, base.steps, base.warmup, base.no_e2e, base.no_fused_check = world, sa.steps, sa.warmup, True, True
    base.keep_process_group = True
    import torch.distributed as dist
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if sa.impl == "ours":
        from ppb200.parallel.mesh import configure_nccl_for_pipeline
        configure_nccl_for_pipeline()
        dist.init_process_group("nccl")
    else:
        dist.init_process_group("nccl", device_id=dev)
    rows = []
    captured = []
    bench.emit = lambda obj: captured.append(obj)
    for sched in sa.schedules.split(","):
        for m in [int(x) for x in sa.microbatches.split(",")]:
            if base.global_batch % m:
                continue
            if sched == "1F1B" and m < world:
                continue
            if sched == "Interleaved1F1B" and (m % world or m < world):
                continue
            row = {"n_gpus": world, "schedule": sched, "n_microbatches": m}
            for impl in [sa.impl]:
                a = copy.copy(base)
                a.schedule, a.microbatches, a.impl = sched, m, impl
                a.virtual_stages = 2 if sched == "Interleaved1F1B" else 1
                if impl == "reference" and sched == "Interleaved1F1B" and 24 % (2 * world):
                    a.virtual_stages = 1      # the helper's own fallback rule (helper:181-185)
                captured.clear()
                try:
                    (bench.run_ours if impl == "ours" else bench.run_reference)(a, rank, local_rank, world)
                except Exception as e:  # noqa: BLE001
                    row[impl + "_error"] = f"{type(e).__name__}: {e}"[:300]
                    import traceback
                    traceback.print_exc()
                if rank == 0 and captured:
                    r = captured[-1]
                    if "value" in r:
                        row[impl + "_tok_s"] = round(r["value"], 1)
                        row[impl + "_ms"] = round(r["ms_per_step"], 3)
                        if impl == "ours":
                            row.update(bubble_critical=round(r["bubble_fraction_critical_rank"], 4),
                                       bubble_max=round(r["bubble_fraction_measured_max"], 4),
                                       bubble_theory=round(r["bubble_fraction_theory"], 4),
                                       partition=r["impl_notes"]["partition"], virtual_stages=a.virtual_stages)
                        else:
                            row["reference_virtual_stages"] = a.virtual_stages
                    else:
                        row[impl + "_unavailable"] = r.get("unavailable")
                gc.collect()
                torch.cuda.empty_cache()
            if rank == 0:
                rows.append(row)
                print(json.dumps(row), flush=True)
                os.makedirs(os.path.dirname(sa.out), exist_ok=True)
                json.dump(rows, open(sa.out, "w"), indent=1)


if __name__ == "__main__":
    main()
