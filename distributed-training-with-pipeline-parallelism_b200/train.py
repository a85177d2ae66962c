"""Training entry point: one process per GPU (torchrun) or per CPU rank (gloo), driven by ``RunConfig``.

    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m ppb200.train --model llama-7b --pp 4 --tp 2 \\
        --schedule 1F1B --n-microbatches 16 --global-batch 64 --seq-len 2048 --steps 100 \\
        --lr 3e-4 --lr-warmup-steps 10 --lr-decay-steps 100 --max-grad-norm 1.0 \\
        --save-dir ckpt --save-every 50 --metrics-path metrics.jsonl

What the reference's ``run_train_iterations`` (helper:98-143) is to its notebook, this is to the framework: mesh +
stage construction, data (synthetic tokens like helper:191-192, or a flat token file), the schedule step with the fused
optimizer, learning-rate schedule, global-norm clipping, checkpoint save / resume across (pp, tp) layouts, JSONL
metrics (step time on the device, tokens/s, loss, gradient norm, measured bubble, peak memory) and an optional chrome
trace of the measured timeline.  Data-parallel replicas read disjoint slices of each global batch.
"""
from __future__ import annotations

import json
import os
import time
from typing import List, Optional

import torch
import torch.distributed as dist

from .models.tp import shard_model_for_tp
from .parallel.mesh import configure_nccl_for_pipeline, init_mesh
from .parallel.partition import build_stage_model, calibrate, resolve_rule
from .parallel.runner import make_schedule
from .parallel.stage import PipelineStage
from .utils import checkpoint
from .utils.config import RunConfig


class TokenData:
    """Batches of (input, target) token ids: next-token targets from a flat token file, or seeded synthetic tokens
    (deterministic per (step, dp rank), so a resumed run sees the batches the uninterrupted run would have seen)."""

    def __init__(self, cfg: RunConfig, vocab: int, dp_idx: int, dp_size: int):
        self.cfg, self.vocab, self.dp_idx, self.dp_size = cfg, vocab, dp_idx, dp_size
        self.tokens = None
        if cfg.data_path:
            import numpy as np
            if cfg.data_path.endswith(".npy"):
                arr = np.load(cfg.data_path, mmap_mode="r")
            else:
                arr = np.memmap(cfg.data_path, dtype=np.dtype(cfg.data_dtype), mode="r")
            if arr.size < cfg.seq_len + 2:
                raise ValueError(f"{cfg.data_path}: {arr.size} tokens is less than one sequence")
            self.tokens = arr

    def batch(self, step: int):
        cfg = self.cfg
        B = cfg.global_batch // self.dp_size
        g = torch.Generator().manual_seed(cfg.seed * 1_000_003 + step * 7919 + self.dp_idx)
        if self.tokens is None:
            x = torch.randint(0, self.vocab, (B, cfg.seq_len), generator=g)
            y = torch.randint(0, self.vocab, (B, cfg.seq_len), generator=g)
            return x, y
        import numpy as np
        starts = torch.randint(0, self.tokens.size - cfg.seq_len - 1, (B,), generator=g).tolist()
        rows = np.stack([np.asarray(self.tokens[s:s + cfg.seq_len + 1]).astype(np.int64) for s in starts])
        t = torch.from_numpy(rows).clamp_(0, self.vocab - 1)
        return t[:, :-1].contiguous(), t[:, 1:].contiguous()


def main(argv: Optional[List[str]] = None) -> dict:
    cfg = RunConfig.from_args(argv)
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    use_cuda = (cfg.device or ("cuda" if torch.cuda.is_available() else "cpu")).startswith("cuda")
    dev = torch.device("cuda", local) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    if cfg.dp * cfg.pp * cfg.tp != world:
        raise ValueError(f"dp*pp*tp = {cfg.dp * cfg.pp * cfg.tp} does not match WORLD_SIZE = {world}")
    own_pg = False
    if world > 1 and not dist.is_initialized():
        backend = cfg.backend or ("nccl" if use_cuda else "gloo")
        if backend == "nccl":
            configure_nccl_for_pipeline()
        os.environ.setdefault("MASTER_ADDR", cfg.master_addr)
        if cfg.master_port:
            os.environ.setdefault("MASTER_PORT", str(cfg.master_port))
        kw = {}
        if cfg.collective_timeout_s:   # dead-peer detection: the reference blocks forever in recv (nb:324-325)
            from datetime import timedelta
            kw["timeout"] = timedelta(seconds=float(cfg.collective_timeout_s))
        dist.init_process_group(backend, rank=rank, world_size=world, **kw)
        own_pg = True
    if world > 1:
        mesh = init_mesh(dp_size=cfg.dp, pp_size=cfg.pp, tp_size=cfg.tp)
        pp_idx, dp_idx, tp_idx = mesh.pp_idx, mesh.dp_idx, mesh.tp_idx
        groups = dict(pp_group=mesh.pp_group, pp_group_bwd=mesh.pp_group_bwd, dp_group=mesh.dp_group)
        tp_group = mesh.tp_group
    else:
        pp_idx = dp_idx = tp_idx = 0
        groups, tp_group = {}, None

    margs = cfg.model_args()
    margs.dtype = torch.bfloat16 if (use_cuda and cfg.dtype == "bf16") else torch.float32
    v = cfg.virtual_stages if "Interleaved" in cfg.schedule else 1
    num_stages = cfg.pp * v
    stages = []
    rule = cfg.partition
    if rule in ("auto", "balanced", "fine") and use_cuda and num_stages > 1 and cfg.tp == 1:
        # stage boundaries from unit times measured on this GPU (rank 0's numbers are used by every rank)
        calibrate(margs, max(1, cfg.global_batch // cfg.dp // cfg.n_microbatches), cfg.seq_len, dev)
    rule = resolve_rule(rule, margs.n_layers, num_stages, margs, cfg.seq_len, tp=cfg.tp)
    for i in range(v):
        sidx = pp_idx + cfg.pp * i
        model = build_stage_model(margs, sidx, num_stages, rule=rule, seq_len=cfg.seq_len, device=dev, seed=cfg.seed)
        if cfg.tp > 1:
            shard_model_for_tp(model, tp_idx, cfg.tp)
        st = PipelineStage(model, sidx, num_stages, dev, dp_group=groups.get("dp_group"), tp_group=tp_group)
        # replicas see different data and must draw different dropout masks; the tp ranks of one replica share the
        # base (models/tp.py mixes the tp rank in itself, so forward and backward seeds still match)
        st.base_seed = cfg.seed + 0x9E37 * dp_idx
        st.materialize()
        stages.append(st)
    start_step = 0
    if cfg.resume_from:
        start_step = max(checkpoint.load_stage(st, cfg.resume_from) for st in stages)
    opt = cfg.optimizer_dict()
    sched = make_schedule(cfg.schedule, stages, cfg.n_microbatches, None, optimizer=opt, recompute=cfg.recompute, **groups)
    first, last = 0 in sched.local, (num_stages - 1) in sched.local
    data = TokenData(cfg, margs.vocab_size, dp_idx, cfg.dp)
    tokens_per_step = cfg.global_batch * cfg.seq_len
    log = open(cfg.metrics_path, "a") if (cfg.metrics_path and rank == world - 1) else None
    pin = use_cuda

    def save(step: int) -> None:
        if cfg.save_dir and dp_idx == 0:
            for st in stages:
                checkpoint.save_stage(st, cfg.save_dir, step=step)
        if world > 1:
            dist.barrier()

    last_loss, history = None, []
    for step in range(start_step, cfg.steps):
        if opt is not None:
            sched.optimizer["lr"] = cfg.lr_at(step)
        x, y = data.batch(step)
        if pin:
            x, y = x.pin_memory().to(dev, non_blocking=True), y.pin_memory().to(dev, non_blocking=True)
        losses: list = []
        t0 = time.perf_counter()
        sched.step(*((x,) if first else ()), target=y if last else None, losses=losses if last else None)
        if last:
            last_loss = float(torch.stack([l.reshape(()) for l in losses]).mean().item())   # device -> host
            history.append(last_loss)
        elif use_cuda:
            torch.cuda.synchronize(dev)
        wall_ms = (time.perf_counter() - t0) * 1e3
        if log is not None and (step % max(1, cfg.log_every) == 0 or step == cfg.steps - 1):
            m = sched.measured()
            rec = {"step": step, "loss": last_loss, "lr": cfg.lr_at(step), "step_ms": m.get("step_ms", wall_ms),
                   "tokens_per_s": tokens_per_step / (m.get("step_ms", wall_ms) / 1e3),
                   "bubble_fraction": m.get("bubble_fraction"), "bubble_theory": sched.theory_bubble}
            if getattr(sched, "last_grad_norm_sq", None) is not None:
                gs = (1.0 / cfg.n_microbatches) / cfg.dp
                rec["grad_norm"] = float(sched.last_grad_norm_sq.sqrt()) * gs
            if use_cuda:
                rec["peak_mem_gib"] = torch.cuda.max_memory_allocated(dev) / 2**30
            log.write(json.dumps(rec) + "\n")
            log.flush()
        if cfg.save_dir and cfg.save_every and (step + 1) % cfg.save_every == 0 and step + 1 < cfg.steps:
            save(step + 1)
    if cfg.save_dir:
        save(cfg.steps)
    if cfg.trace_path:
        from .utils.trace import gather_trace, write_chrome_trace
        evs = gather_trace(sched)
        if rank == 0 and evs is not None:
            write_chrome_trace(evs, cfg.trace_path)
    if log is not None:
        log.close()
    out = {"rank": rank, "final_loss": last_loss, "losses": history, "steps": cfg.steps - start_step}
    if own_pg:
        dist.barrier()
        dist.destroy_process_group()
    return out


if __name__ == "__main__":
    res = main()
    if res.get("final_loss") is not None:
        print(json.dumps({k: res[k] for k in ("final_loss", "steps")}))
