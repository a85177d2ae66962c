"""Reference-compatible public API.

The names, positional signatures and result-dict keys are the ones the reference notebook imports and calls
(/root/reference/LLMsDistributedTrainingHelper.py:23-235; notebook cells at nb:306-394, nb:405-435):

    ModelArgs, Transformer, manual_model_split, run_train_iterations, worker_process,
    run_one_experiment, run_all_experiments, compute_speedup_and_efficiency

Behaviour that is fixed on purpose (SURVEY.md section 5.5 "measurement flaws to fix, not copy"): timing is
bracketed by a barrier + device sync and max-reduced over ranks instead of wall-clock on the last rank only; the
rendezvous uses an ephemeral port; the launcher has a timeout and kills all ranks on the first failure.
Everything else keeps the reference semantics, including the silent virtual-stage fallback of helper:181-185.
"""
from __future__ import annotations

import os
import time
import traceback
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from .models import ModelArgs, Transformer  # noqa: F401  (re-exported)
from .parallel.partition import prune_to_stage
from .parallel.runner import PipelineSchedule, Schedule1F1B, ScheduleGPipe, ScheduleInterleaved1F1B, make_schedule  # noqa: F401
from .parallel.stage import PipelineStage
from .utils.config import RunConfig  # noqa: F401
from .utils.analysis import compute_speedup_and_efficiency  # noqa: F401


def manual_model_split(model: Transformer, stage_index: int, num_stages: int, device: torch.device,
                       rule: str = "reference") -> PipelineStage:
    """split model across pipeline stages manually (same contract as helper:60-94): stage ``stage_index`` keeps
    ``n_layers // num_stages`` layers (the last stage takes the remainder), stage 0 keeps the embedding, the last
    stage keeps norm + output; the model is pruned IN PLACE and wrapped in a ``PipelineStage``."""
    prune_to_stage(model, stage_index, num_stages, rule=rule)
    return PipelineStage(model, stage_index, num_stages, device)


def run_train_iterations(schedule: PipelineSchedule, x: torch.Tensor, y: torch.Tensor, rank: int, world_size: int,
                         num_iterations: int = 10, warmup: int = 2) -> Dict[str, float]:
    """run training iterations and measure performance metrics (helper:98-143).

    Returns the reference's keys ``elapsed_time`` / ``throughput`` / ``tokens_processed`` plus
    ``step_time_ms``, ``bubble_fraction`` (measured on this rank), ``bubble_theory``, ``loss``."""
    total_toks = x.shape[0] * x.shape[1] * num_iterations
    first = 0 in schedule.local
    last = (schedule.num_stages - 1) in schedule.local
    losses: List[torch.Tensor] = []

    def one_step():
        losses.clear()
        schedule.step(*((x,) if first else ()), target=y if last else None, losses=losses if last else None)

    for _ in range(warmup):
        one_step()
    distributed = dist.is_available() and dist.is_initialized()
    cuda = schedule.cuda
    if distributed:
        dist.barrier()
    if cuda:
        torch.cuda.synchronize(schedule.device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    start_t = time.perf_counter()
    for _ in range(num_iterations):
        one_step()
    if cuda:
        e1.record()
        torch.cuda.synchronize(schedule.device)
        elapsed_t = e0.elapsed_time(e1) / 1e3
    else:
        elapsed_t = time.perf_counter() - start_t
    if distributed:  # every number is the max over ranks (BASELINE.json)
        t = torch.tensor([elapsed_t], dtype=torch.float64, device=schedule.device if cuda else "cpu")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed_t = float(t.item())
    meas = schedule.measured()
    out = {
        "elapsed_time": elapsed_t,
        "throughput": total_toks / elapsed_t,
        "tokens_processed": total_toks,
        "step_time_ms": elapsed_t / max(1, num_iterations) * 1e3,
        "bubble_fraction": meas.get("bubble_fraction", float("nan")),
        "bubble_theory": schedule.theory_bubble,
    }
    if last and losses:
        out["loss"] = float(torch.stack([l.reshape(()) for l in losses]).mean().item())
    return out


def worker_process(rank, world_size, n_layers, n_heads, schedule_type, batch_size, seq_length, num_iterations,
                   results_queue, *, backend: Optional[str] = None, device: Optional[str] = None,
                   n_microbatches: int = 4, master_port: Optional[int] = None, master_addr: str = "127.0.0.1",
                   dim: int = 768, vocab_size: int = 10000, dropout: float = 0.1, optimizer: Optional[dict] = None,
                   seed: Optional[int] = None, virtual_stages: Optional[int] = None, model_kwargs: Optional[dict] = None,
                   warmup: int = 2, all_ranks_report: bool = False, collective_timeout_s: Optional[float] = None):
    """Individual worker process for distributed training (same positional signature as helper:150-151).

    Keyword-only extras expose what the helper hard-codes: backend ('gloo' on CPU, 'nccl' on CUDA), device,
    ``n_microbatches`` (helper:214 fixes 4), rendezvous address/port (helper:168-169 fix localhost:29500), model
    width/vocab/dropout, an optional fused-AdamW config, and the seed."""
    try:
        use_cuda = (device or "").startswith("cuda") or (device is None and backend == "nccl")
        backend = backend or ("nccl" if use_cuda else "gloo")
        os.environ["MASTER_ADDR"] = master_addr
        os.environ["MASTER_PORT"] = str(master_port or os.environ.get("MASTER_PORT", 29500))
        os.environ["RANK"] = str(rank)
        os.environ["LOCAL_RANK"] = str(rank)
        os.environ["WORLD_SIZE"] = str(world_size)
        if use_cuda:
            torch.cuda.set_device(rank % torch.cuda.device_count())
            dev = torch.device("cuda", rank % torch.cuda.device_count())
        else:
            dev = torch.device("cpu")
            # One intra-op thread pool per rank, sized so that the ranks together do not oversubscribe the host.  The
            # helper leaves torch's default (every rank spins up a pool as large as the machine, helper:150-235), which
            # on an 8-core host makes a 2-rank step of a small model 20-30x slower than it has to be; an explicit
            # OMP_NUM_THREADS from the user wins.
            if "OMP_NUM_THREADS" not in os.environ:
                torch.set_num_threads(max(1, (os.cpu_count() or 1) // world_size))
        if backend == "nccl":
            from .parallel.mesh import configure_nccl_for_pipeline
            configure_nccl_for_pipeline()
        pg_kw = {}
        if collective_timeout_s:   # a peer that never shows up raises here instead of blocking forever (helper: no timeout)
            from datetime import timedelta
            pg_kw["timeout"] = timedelta(seconds=float(collective_timeout_s))
        dist.init_process_group(backend=backend, rank=rank, world_size=world_size, **pg_kw)

        # virtual-stage rule of helper:181-185 (silent fallback to one chunk per rank)
        stages_per_w = 1
        if schedule_type == "Interleaved1F1B":
            v = 2 if virtual_stages is None else virtual_stages
            if n_layers % (world_size * v) == 0:
                stages_per_w = v
        num_stages = world_size * stages_per_w

        model_args = ModelArgs(dim=dim, n_layers=n_layers, n_heads=n_heads, vocab_size=vocab_size, dropout=dropout,
                               **(model_kwargs or {}))
        if seed is not None:
            torch.manual_seed(seed)
        x = torch.randint(0, model_args.vocab_size, (batch_size, seq_length), dtype=torch.long).to(dev)
        y = torch.randint(0, model_args.vocab_size, (batch_size, seq_length), dtype=torch.long).to(dev)

        stages = []
        for i in range(stages_per_w):
            if seed is not None:
                torch.manual_seed(seed)  # every rank builds the same full model, then prunes (helper:207-209)
            model = Transformer(model_args)
            stage_idx = rank + (world_size * i)
            stage = manual_model_split(model, stage_idx, num_stages, dev)
            stage.base_seed = seed or 0
            stages.append(stage)

        schedule = make_schedule(schedule_type, stages, n_microbatches, None, optimizer=optimizer)
        metrics = run_train_iterations(schedule, x, y, rank, world_size, num_iterations, warmup=warmup)
        metrics["virtual_stages"] = stages_per_w

        # only last rank returns metrics (helper:226-227)
        if rank == world_size - 1 or all_ranks_report:
            results_queue.put(metrics)
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # same error-dict contract as helper:231-235
        print(f"Error in rank {rank}: {e}")
        traceback.print_exc()
        results_queue.put({"error": str(e)})


def run_one_experiment(n_layers, n_heads, num_processes, schedule_type, batch_size=32, seq_length=128,
                       num_iterations=10, **worker_kwargs) -> dict:
    """Spawn ``num_processes`` ranks and return the metrics dict of the last rank (nb:306-337)."""
    from .parallel.launch import launch_workers
    return launch_workers(worker_process, num_processes,
                          (n_layers, n_heads, schedule_type, batch_size, seq_length, num_iterations), worker_kwargs)


def run_all_experiments(layers=(4, 8, 12), heads=(4, 8, 12), processes=(2, 4),
                        schedules=("GPipe", "1F1B", "Interleaved1F1B"), num_iterations: int = 5, verbose: bool = True,
                        **worker_kwargs):
    """The reference's 54-config sweep (nb:345-394); returns a pandas DataFrame with the same columns."""
    import pandas as pd
    rows = []
    for n_layers in layers:
        for n_heads in heads:
            for p in processes:
                for sched in schedules:
                    res = run_one_experiment(n_layers, n_heads, p, sched, num_iterations=num_iterations, **worker_kwargs)
                    if "error" in res:
                        if verbose:
                            print(f"  L{n_layers} H{n_heads} P{p} {sched}: ERROR {res['error']}")
                        continue
                    rows.append({"n_layers": n_layers, "n_heads": n_heads, "num_processes": p, "schedule": sched,
                                 "throughput": res["throughput"], "elapsed_time": res["elapsed_time"],
                                 "tokens_processed": res["tokens_processed"],
                                 "bubble_fraction": res.get("bubble_fraction"), "bubble_theory": res.get("bubble_theory")})
                    if verbose:
                        print(f"  L{n_layers} H{n_heads} P{p} {sched}: {res['throughput']:.1f} tok/s")
    return pd.DataFrame(rows)
