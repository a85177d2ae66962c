#!/usr/bin/env python
"""Headline benchmark (driver contract): tokens/s of one full training step of the flagship pipeline config.

    python bench.py --gpus 1 --steps K --warmup W
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the UNMODIFIED reference (baseline/_ref) on NCCL + cuBLAS

Config (BASELINE.json config 2 without the DP replicas, so that it exists at every N): the helper's LM scaled to
GPT-2 medium (dim 1024, 24 layers, 16 heads, vocab 50257; FFN 2048 / post-LN / self+cross attention / dropout 0.1
are what the helper's nn.TransformerDecoderLayer fixes), sequence 1024, global batch 128 sequences in 32
micro-batches, 1F1B over N pipeline stages (N = 1: one stage, 32 accumulated micro-batches), bf16 compute with
fp32 master weights, fused AdamW INSIDE the timed step.  Strong scaling: the global batch is fixed as N grows.
Data is synthetic (random tokens) and weights are random-init -- there is no network for datasets/checkpoints.

Timing: W (>= 3) warm-up steps, then exactly K steps between barrier + torch.cuda.synchronize() on both sides,
CUDA events on the compute stream, max over ranks.  Every step touches > 2 GB of weights/activations per GPU, far
more than the 126 MB L2, so no explicit L2 flush is needed (stated in config.l2).  `e2e` repeats the measurement
through the public API with a pinned-host -> device copy of the inputs and a device -> host read of the loss in
every step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="ref-gpt2-medium")
    ap.add_argument("--schedule", default="1F1B", choices=["GPipe", "1F1B", "Interleaved1F1B", "ZeroBubble1F1B", "ZeroBubbleInterleaved"])
    ap.add_argument("--global-batch", type=int, default=128)
    ap.add_argument("--seq-len", type=int, default=1024)
    ap.add_argument("--microbatches", type=int, default=32)
    ap.add_argument("--virtual-stages", type=int, default=1)
    ap.add_argument("--dp", type=int, default=1)
    ap.add_argument("--tp", type=int, default=1, help="intra-stage tensor parallel degree (pre-LN presets, dropout 0)")
    ap.add_argument("--partition", default="auto", choices=["auto", "reference", "balanced", "fine"],
                    help="auto = measure the per-unit costs on the GPU, cut inside layers when that lowers the largest stage")
    ap.add_argument("--dropout", type=float, default=None)
    ap.add_argument("--n-layers", type=int, default=None)
    ap.add_argument("--extra-warmup", type=int, default=None,
                    help="1F1B: forwards kept in flight beyond torch's warm-up count (hides stage-to-stage transfer latency)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fused-check", action="store_true",
                    help="skip the K10/K11/K12 parity + roofline check that multi-GPU runs append (ranks 0-1, after the timed regions)")
    ap.add_argument("--fused-check-timeout", type=float, default=150.0,
                    help="seconds the appended K10/K11/K12 check may take before the line is printed without it")
    ap.add_argument("--trace", default=None, help="write a chrome trace of the measured timeline here")
    ap.add_argument("--small", action="store_true", help="tiny config for smoke tests")
    return ap.parse_args()


def env_rank():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def emit(obj):
    print(json.dumps(obj), flush=True)


def config_dict(a, dims: dict, parallelism: str) -> dict:
    """The SAME dict from both arms (driver check `same_config`): what is measured, nothing about how."""
    return {"model": {"name": a.model, "dim": dims["dim"], "n_layers": dims["n_layers"], "n_heads": dims["n_heads"],
                      "vocab": dims["vocab_size"], "ffn": dims.get("ffn_dim", 2048), "dropout": dims.get("dropout", 0.1),
                      "block": dims.get("block", "nn.TransformerDecoderLayer semantics: post-LN, self + cross attention over the layer input, ReLU")},
            "global_batch": a.global_batch, "seq_len": a.seq_len, "n_microbatches": a.microbatches,
            "schedule": a.schedule, "parallelism": parallelism,
            "optimizer": "AdamW lr 1e-4 betas (0.9, 0.95) wd 0.1, fused, inside the timed region",
            "l2": "inputs >> L2: >2 GB of weights+activations touched per step per GPU, no explicit flush"}


def max_over_ranks(val: float, device) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([val], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return val


def sum_over_ranks(val: float, device) -> float:
    if dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([val], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())
    return val


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier(device_ids=[torch.cuda.current_device()])


# ---------------------------------------------------------------------------------------------------------------
def run_ours(a, rank, local_rank, world):
    from ppb200.models import preset
    from ppb200.ops import _lib
    from ppb200.parallel.mesh import init_mesh
    from ppb200.parallel.partition import build_stage_model
    from ppb200.parallel.runner import make_schedule
    from ppb200.parallel.stage import PipelineStage
    from ppb200.utils.clocks import ClockSampler
    from ppb200.utils.trace import gather_trace, write_chrome_trace

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    _lib.load(required=True)
    if world > 1:
        # no device_id: lazy communicator creation keeps P2P on per-peer communicators/streams (eager init would
        # serialise every send/recv with every other op of the group)
        from ppb200.parallel.mesh import configure_nccl_for_pipeline
        configure_nccl_for_pipeline()
        if not dist.is_initialized():    # (benchmarks/sweep_vs_reference.py runs many cells on one process group)
            dist.init_process_group("nccl")
        mesh = init_mesh(dp_size=a.dp, pp_size=world // (a.dp * a.tp), tp_size=a.tp)
        pp_group, dp_group, pp_idx, dp_idx, P = mesh.pp_group, mesh.dp_group, mesh.pp_idx, mesh.dp_idx, mesh.pp_size
        pp_group_bwd = mesh.pp_group_bwd
        tp_group, tp_idx = mesh.tp_group, mesh.tp_idx
    else:
        pp_group = dp_group = pp_group_bwd = tp_group = None
        pp_idx = dp_idx = tp_idx = 0
        P = 1
    over = {"dtype": torch.bfloat16}
    if a.dropout is not None:
        over["dropout"] = a.dropout
    if a.n_layers is not None:
        over["n_layers"] = a.n_layers
    if a.small:
        over.update(dim=256, n_layers=max(2 * P * a.virtual_stages, 4), n_heads=4, vocab_size=4096)
    margs = preset(a.model, **over)
    v = a.virtual_stages
    num_stages = P * v
    from ppb200.parallel import partition as Pn
    unit_ms = None
    if a.partition in ("auto", "balanced", "fine") and num_stages > 1 and a.tp == 1:
        # stage boundaries are chosen from per-unit forward+backward times MEASURED here (about a second), not from constants
        unit_ms = Pn.calibrate(margs, a.global_batch // a.dp // a.microbatches, a.seq_len, dev)
    rule = Pn.resolve_rule(a.partition, margs.n_layers, num_stages, margs, a.seq_len, tp=a.tp)
    stages = []
    for i in range(v):
        sidx = pp_idx + P * i
        model = build_stage_model(margs, sidx, num_stages, rule=rule, seq_len=a.seq_len, device=dev, seed=1234)
        if a.tp > 1:
            from ppb200.models.tp import shard_model_for_tp
            shard_model_for_tp(model, tp_idx, a.tp)
        st = PipelineStage(model, sidx, num_stages, dev, group=pp_group, dp_group=dp_group, tp_group=tp_group)
        st.base_seed = 1234 + 0x9E37 * dp_idx   # replicas draw different dropout masks
        stages.append(st)
    opt = {"lr": 1e-4, "betas": (0.9, 0.95), "weight_decay": 0.1}
    sched = make_schedule(a.schedule, stages, a.microbatches, None, pp_group=pp_group, pp_group_bwd=pp_group_bwd,
                          dp_group=dp_group, optimizer=opt, extra_warmup=a.extra_warmup)

    B, S = a.global_batch // a.dp, a.seq_len
    first, last = 0 in sched.local, (num_stages - 1) in sched.local
    g = torch.Generator().manual_seed(1000 + dp_idx)
    x_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    y_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    x_dev, y_dev = x_host.to(dev), y_host.to(dev)
    losses = []

    def step_resident():
        losses.clear()
        sched.step(*((x_dev,) if first else ()), target=y_dev if last else None, losses=losses if last else None)

    def step_e2e():
        losses.clear()
        if first:
            x_dev.copy_(x_host, non_blocking=True)
        if last:
            y_dev.copy_(y_host, non_blocking=True)
        sched.step(*((x_dev,) if first else ()), target=y_dev if last else None, losses=losses if last else None)
        if last:
            return float(torch.stack([l.reshape(()) for l in losses]).mean().item())  # device -> host read
        return None

    def timed(fn, k):
        barrier()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        barrier()
        torch.cuda.synchronize(dev)
        return max_over_ranks(e0.elapsed_time(e1), dev)

    for _ in range(max(a.warmup, 3) if not a.small else max(a.warmup, 1)):
        step_resident()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    _lib.reset_launches()
    ms_total = timed(step_resident, a.steps)
    launches = sum_over_ranks(_lib.launches(), dev)
    clocks = sampler.stop() if sampler is not None else None
    meas = sched.measured()
    bubble = max_over_ranks(meas.get("bubble_fraction", 0.0), dev)
    bubble_min = -max_over_ranks(-meas.get("bubble_fraction", 0.0), dev)   # idle fraction of the busiest (critical) rank
    loss_val = None
    if last and losses:
        loss_val = float(torch.stack([l.reshape(()) for l in losses]).mean().item())
    e2e = None
    if not a.no_e2e:
        step_e2e()
        ms_e2e = timed(step_e2e, a.steps)
        tokens = a.global_batch * S
        e2e = {"value": tokens * a.steps / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / a.steps,
               "h2d_bytes_per_step": int(a.dp * a.tp * 2 * B * S * 8), "d2h_bytes_per_step": int(a.dp * 4)}
    if a.trace:
        evs = gather_trace(sched)
        if rank == 0 and evs is not None:
            write_chrome_trace(evs, a.trace)
    peak_mem = max_over_ranks(torch.cuda.max_memory_allocated(dev) / 2**30, dev)
    loss_t = torch.tensor([loss_val if loss_val is not None else -1.0], dtype=torch.float64, device=dev)
    summ = {k: round(v, 3) for k, v in sched.summary().items()}
    summ["layers"] = [[f"{k}:{getattr(l, 'part', 'full')}" if getattr(l, 'part', 'full') != 'full' else k
                       for k, l in st.submod.layers.items()] for st in stages]
    per_rank = [summ]
    if world > 1:
        dist.all_reduce(loss_t, op=dist.ReduceOp.MAX)
        per_rank = [None] * world
        dist.all_gather_object(per_rank, summ)
    result = None
    if rank == 0:   # the whole line is assembled BEFORE the appended check below: nothing that follows can cost it
        tokens = a.global_batch * S
        ms_step = ms_total / a.steps
        par = f"pp{P}" + (f"xdp{a.dp}" if a.dp > 1 else "") + (f"xtp{a.tp}" if a.tp > 1 else "") + (f"xv{v}" if v > 1 else "")
        result = {
            "metric": "tokens/sec (1F1B pipeline training step, max over ranks)" if a.schedule == "1F1B" else f"tokens/sec ({a.schedule})",
            "value": tokens / (ms_step / 1e3), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic random tokens, random-init weights", "impl": "ours",
            "config": config_dict(a, {"dim": margs.dim, "n_layers": margs.n_layers, "n_heads": margs.n_heads,
                                      "vocab_size": margs.vocab_size, "ffn_dim": margs.ffn_dim, "dropout": margs.dropout,
                                      **({} if a.model.startswith("ref-") else {"block": a.model})}, par),
            "impl_notes": {"partition": rule, "partition_requested": a.partition, "extra_warmup": getattr(sched, "extra_warmup", 0),
                           "nccl_p2p_channels": os.environ.get("NCCL_MAX_P2P_NCHANNELS"), "reserved_sms": os.environ.get("PPB_RESERVED_SMS"),
                           "partition_note": "the one real difference to the reference arm: stage boundaries are balanced on measured per-unit times (and may fall inside a layer); the reference uses helper:70-75 (n_layers // stages, remainder to the last stage)",
                           "unit_costs_ms": ({k: (round(v, 4) if isinstance(v, float) else v) for k, v in unit_ms.items()} if unit_ms else None),
                           "stack": "ppb200: explicit fused fwd/bwd stage program, libppb200.so sm_100a kernels, CUDA graphs per slot, NCCL p2p"},
            "bubble_fraction_measured_max": bubble, "bubble_fraction_critical_rank": bubble_min,
            "bubble_fraction_theory": sched.theory_bubble,
            "loss": float(loss_t.item()), "peak_mem_gib": peak_mem, "gpu_launches": int(launches), "clocks": clocks, "e2e": e2e,
            "model_tflops_per_gpu": margs.flops_per_token(S) * tokens / (ms_step / 1e3) / 1e12 / world,
            "fused_paths": None,
            "per_rank": per_rank,
        }
    # The fused compute+collective kernels (K10 wgrad+all-reduce, K11 GEMM->reduce-scatter, K12 all-gather->GEMM) are not
    # on the pipeline-only headline path, so every multi-GPU run appends their parity + roofline check over the real
    # NVLink peers (ranks 0 and 1, outside every timed region): the evidence travels with the driver's own records.
    # It runs under a deadline: if it has not come back in time, rank 0 prints the line it already holds (with the
    # time-out noted under `fused_paths`) and every rank leaves, so a stalled peer can never cost the measurement.
    if world > 1 and not a.no_fused_check and not a.small:
        from ppb200.utils.deadline import Deadline

        def on_timeout():
            if rank == 0:
                result["fused_paths"] = {"error": f"check did not finish within {a.fused_check_timeout:.0f} s; the headline numbers above were complete before it started"}
                emit(result)
            sys.stdout.flush()
            os._exit(0)

        with Deadline(a.fused_check_timeout, on_timeout) as dl:
            pair = dist.new_group([0, 1])
            fused_paths = None
            if rank < 2:
                try:
                    from ppb200.utils import fused_check
                    # (the NVLS multimem variant and the degree-4 run live in benchmarks/fused_paths_bench.py -> profiles/)
                    fused_paths = fused_check.run(pair, dev, nvls=False)
                except Exception as e:  # noqa: BLE001
                    fused_paths = {"error": f"{type(e).__name__}: {e}"[:300]}
            dist.barrier()
        if not dl.completed:   # the time-out handler owns the output from here (it ends the process)
            import time
            time.sleep(3600)
        if rank == 0:
            result["fused_paths"] = fused_paths
    if rank == 0:
        emit(result)
    if world > 1:
        dist.barrier()
        if not getattr(a, "keep_process_group", False):
            dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------------------
def reference_stage_with_replicas(ref, model, stage_index: int, num_stages: int, dev, pp_group, dp_group):
    """The helper's own `manual_model_split` (its partition rule and pruning, helper:60-94) for a pipeline that has
    data-parallel replicas.  The helper builds its PipelineStage on the WORLD group (helper:93), which torch rejects as
    soon as the world is larger than the pipeline ("group size cannot be larger than number of stages"), so for the call
    the `PipelineStage` name inside the helper module is bound to the stock composition SURVEY.md 6.4 names: the pruned
    module moved to its device, wrapped in DistributedDataParallel over the replica group (torch's stage runtime drives
    `no_sync` for all but the last micro-batch), torch's PipelineStage on the pipeline group.  The file on disk is
    untouched; tests/test_reference_arm.py runs this on CPU (gloo, 2 stages x 2 replicas)."""
    from torch.distributed.pipelining import PipelineStage as TorchStage
    from torch.nn.parallel import DistributedDataParallel as DDP

    def stage_on_pipeline_group(module, idx, n_stages, device):
        module.to(device)
        if any(True for _ in module.parameters()):
            module = DDP(module, device_ids=[device.index] if device.type == "cuda" else None, process_group=dp_group,
                         gradient_as_bucket_view=True)
        return TorchStage(module, idx, n_stages, device, group=pp_group)

    saved = ref.PipelineStage
    ref.PipelineStage = stage_on_pipeline_group
    try:
        return ref.manual_model_split(model, stage_index, num_stages, dev)
    finally:
        ref.PipelineStage = saved


def run_reference(a, rank, local_rank, world):
    """The unmodified helper (baseline/_ref/LLMsDistributedTrainingHelper.py) through its own public API:
    ``Transformer`` + ``manual_model_split`` + the torch.distributed.pipelining schedule classes it uses
    (helper:12, :215-220), on the NCCL backend with cuBLAS/cuDNN kernels, bf16 parameters, torch's fused AdamW."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "LLMsDistributedTrainingHelper.py")):
        emit({"impl": "reference", "unavailable": "baseline/_ref is missing (pip install --target baseline/_ref not run)"})
        return
    sys.path.insert(0, ref_dir)
    try:
        import LLMsDistributedTrainingHelper as ref
        from torch.distributed.pipelining import Schedule1F1B, ScheduleGPipe, ScheduleInterleaved1F1B
    except Exception as e:  # noqa: BLE001
        emit({"impl": "reference", "unavailable": f"import failed: {e}"[:200]})
        return
    from ppb200.utils.clocks import ClockSampler

    # what the helper cannot express is reported, not emulated: it has one block (post-LN, ReLU, FFN 2048), no data or
    # tensor parallelism (helper:178 creates an unused group) and only torch's three schedules (helper:215-220)
    why = None
    if a.tp > 1 and a.dp == 1 and a.schedule in ("GPipe", "1F1B") and not a.model.startswith("ref-"):
        # tensor parallelism is not in the helper (helper:178 creates a group it never uses) and its fixed block cannot be
        # sharded: the baseline for the tensor-parallel configurations is the composition a torch user would write --
        # a Megatron-style tensor + sequence parallel pre-LN block from torch modules (ncclAllGather / ncclReduceScatter,
        # cuBLAS, SDPA) inside torch.distributed.pipelining, benchmarks/tp_baseline.py (parity-tested on CPU)
        from benchmarks import tp_baseline
        tp_baseline.run(a, rank, local_rank, world, emit, config_dict, max_over_ranks)
        return
    if a.model not in ("ref-gpt2-medium", "ref-gpt2-xl", "ref-small"):
        why = f"the helper's Transformer cannot express model preset {a.model!r} (fixed nn.TransformerDecoderLayer block)"
    elif a.tp > 1:
        why = "the helper has no tensor parallelism (pipeline stages only) and its fixed block cannot be sharded"
    elif a.dp > 1 and a.schedule == "Interleaved1F1B":
        why = "DP baseline is built for one stage per rank (GPipe / 1F1B)"
    elif a.schedule not in ("GPipe", "1F1B", "Interleaved1F1B"):
        why = f"the helper selects among GPipe / 1F1B / Interleaved1F1B only (got {a.schedule})"
    if why is not None:
        if rank == 0:
            emit({"impl": "reference", "unavailable": why})
        return
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world == 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        os.environ.setdefault("RANK", "0")
        os.environ.setdefault("WORLD_SIZE", "1")
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    # data parallelism is not in the helper (helper:178 creates a group it never uses): the DP baseline is the stock
    # torch composition SURVEY.md 6.4 names -- the helper's pruned stage module wrapped in DistributedDataParallel
    # (torch's PipelineStage drives no_sync for all but the last micro-batch), NCCL all-reduce of the gradients
    dp = a.dp
    pp_world = world // dp
    pp_rank, dp_idx = rank % pp_world, rank // pp_world
    pp_group = dp_group = None
    if dp > 1:
        for d in range(dp):
            g_ = dist.new_group([d * pp_world + p_ for p_ in range(pp_world)])
            if d == dp_idx:
                pp_group = g_
        for p_ in range(pp_world):
            g_ = dist.new_group([d * pp_world + p_ for d in range(dp)])
            if p_ == pp_rank:
                dp_group = g_
    dims = {"ref-gpt2-medium": dict(dim=1024, n_layers=24, n_heads=16, vocab_size=50257),
            "ref-gpt2-xl": dict(dim=1600, n_layers=48, n_heads=25, vocab_size=50257),
            "ref-small": dict()}[a.model]
    if a.n_layers is not None:
        dims["n_layers"] = a.n_layers
    if a.small:
        dims = dict(dim=256, n_layers=max(2 * world * a.virtual_stages, 4), n_heads=4, vocab_size=4096)
    margs = ref.ModelArgs(**dims)
    v = a.virtual_stages if a.schedule == "Interleaved1F1B" else 1
    num_stages = pp_world * v
    B, S = a.global_batch // dp, a.seq_len
    g = torch.Generator().manual_seed(1000 + dp_idx)
    x_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    y_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    x, y = x_host.to(dev), y_host.to(dev)

    def tokenwise_loss_fn(outputs, targets):  # helper:197-201
        loss_fn = torch.nn.CrossEntropyLoss()
        return loss_fn(outputs.reshape(-1, margs.vocab_size), targets.reshape(-1))

    stages, params = [], []
    for i in range(v):
        torch.manual_seed(1234)
        model = ref.Transformer(margs).to(torch.bfloat16)
        if dp == 1:
            stage = ref.manual_model_split(model, rank + world * i, num_stages, dev)
            model.to(dev)
        else:
            stage = reference_stage_with_replicas(ref, model, pp_rank + pp_world * i, num_stages, dev, pp_group, dp_group)
        stages.append(stage)
        params += list(model.parameters())
    if a.schedule == "GPipe":
        schedule = ScheduleGPipe(stages[0], n_microbatches=a.microbatches, loss_fn=tokenwise_loss_fn)
    elif a.schedule == "1F1B":
        schedule = Schedule1F1B(stages[0], n_microbatches=a.microbatches, loss_fn=tokenwise_loss_fn)
    else:
        schedule = ScheduleInterleaved1F1B(stages, n_microbatches=a.microbatches, loss_fn=tokenwise_loss_fn)
    opt = torch.optim.AdamW(params, lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1, fused=True) if params else None
    losses = []

    def step(e2e=False):
        losses.clear()
        if e2e:
            if pp_rank == 0:
                x.copy_(x_host, non_blocking=True)
            if pp_rank == pp_world - 1:
                y.copy_(y_host, non_blocking=True)
        if pp_world == 1:
            schedule.step(x, target=y, losses=losses)
        elif pp_rank == 0:
            schedule.step(x)
        elif pp_rank == pp_world - 1:
            schedule.step(target=y, losses=losses)
        else:
            schedule.step()
        if opt is not None:
            opt.step()
            opt.zero_grad(set_to_none=False)
        if e2e and pp_rank == pp_world - 1:
            return float(torch.stack(losses).float().mean().item())
        return None

    def timed(fn, k):
        dist.barrier()
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        dist.barrier()
        torch.cuda.synchronize(dev)
        return max_over_ranks(e0.elapsed_time(e1), dev)

    try:
        for _ in range(max(a.warmup, 3) if not a.small else max(a.warmup, 1)):
            step()
        sampler = ClockSampler(local_rank).start() if rank == 0 else None
        ms_total = timed(step, a.steps)
        clocks = sampler.stop() if sampler is not None else None
        e2e = None
        if not a.no_e2e:
            step(True)
            ms_e2e = timed(lambda: step(True), a.steps)
            e2e = {"value": a.global_batch * S * a.steps / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / a.steps,
                   "h2d_bytes_per_step": int(dp * 2 * B * S * 8), "d2h_bytes_per_step": 4 * dp}
        loss_val = float(torch.stack(losses).float().mean().item()) if losses else -1.0
        loss_t = torch.tensor([loss_val], dtype=torch.float64, device=dev)
        dist.all_reduce(loss_t, op=dist.ReduceOp.MAX)
        peak_mem = max_over_ranks(torch.cuda.max_memory_allocated(dev) / 2**30, dev)
        if rank == 0:
            ms_step = ms_total / a.steps
            emit({
                "metric": "tokens/sec (1F1B pipeline training step, max over ranks)" if a.schedule == "1F1B" else f"tokens/sec ({a.schedule})",
                "value": a.global_batch * S / (ms_step / 1e3), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic random tokens, random-init weights", "impl": "reference",
                "config": config_dict(a, {"dim": margs.dim, "n_layers": margs.n_layers, "n_heads": margs.n_heads,
                                          "vocab_size": margs.vocab_size}, f"pp{pp_world}" + (f"xdp{dp}" if dp > 1 else "") + (f"xv{v}" if v > 1 else "")),
                "impl_notes": {"partition": "reference (helper:70-75)",
                               "data_parallel": (None if dp == 1 else "stock torch composition: DistributedDataParallel around the helper's pruned stage module inside torch's PipelineStage (no_sync on all but the last micro-batch), NCCL all-reduce"),
                               "stack": "unmodified helper Transformer/manual_model_split + torch.distributed.pipelining + NCCL p2p + cuBLAS/SDPA, bf16 parameters, torch.optim.AdamW(fused=True) after schedule.step"},
                "loss": float(loss_t.item()), "peak_mem_gib": peak_mem, "gpu_launches": 0, "clocks": clocks, "e2e": e2e,
            })
    except Exception as e:  # noqa: BLE001
        import traceback
        traceback.print_exc()
        if rank == 0:
            emit({"impl": "reference", "unavailable": f"reference run failed: {type(e).__name__}: {e}"[:300]})
    finally:
        try:
            dist.barrier()
            if not getattr(a, "keep_process_group", False):
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


def main():
    a = parse()
    rank, local_rank, world = env_rank()
    if world != a.gpus and world > 1:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}")
    if a.gpus > 1 and world == 1:
        # convenience: re-launch ourselves under torchrun
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + (os.getpid() % 1000)), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if not torch.cuda.is_available():
        emit({"impl": a.impl, "unavailable": "no CUDA device in this process"})
        return
    if a.impl == "reference":
        run_reference(a, rank, local_rank, world)
    else:
        run_ours(a, rank, local_rank, world)


if __name__ == "__main__":
    main()
