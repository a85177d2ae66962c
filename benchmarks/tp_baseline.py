#!/usr/bin/env python
"""Stock-torch baseline for the tensor-parallel configurations (BASELINE.json configs 3 and 4: pre-LN GPT / Llama
blocks, pipeline stages x tensor-parallel ranks).

The reference has no tensor parallelism (its helper creates a group it never uses,
/root/reference/LLMsDistributedTrainingHelper.py:178) and its fixed block cannot be sharded, so `bench.py --impl
reference --tp N` measures the composition a torch user would write instead -- nothing from this repository's kernels
or engine is on this path:

  * Megatron-style tensor + sequence parallel block out of torch modules: activations stay sharded over the token rows
    between blocks, `all_gather_into_tensor` in front of every column-parallel linear and `reduce_scatter_tensor`
    behind every row-parallel one (NCCL on GPU, gloo in the CPU tests), `scaled_dot_product_attention` on the local
    heads, cuBLAS linears, torch LayerNorm / RMSNorm;
  * the pipeline around it is `torch.distributed.pipelining` (PipelineStage + Schedule1F1B / GPipe), exactly what the
    helper uses for its own model (helper:12, :215-220), one pipeline per tensor-parallel index;
  * parameters every tensor-parallel rank holds in full (norms, row-parallel biases, embedding, head) get one
    all-reduce of their gradients per step; torch's fused AdamW.

`train_steps` is the testable core (CPU, gloo: tests/test_tp_baseline.py checks it against the same model on one
rank); `run` is the benchmark arm `bench.py` calls."""
from __future__ import annotations

import math
import os
import sys
from typing import List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as TF

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


# ---------------------------------------------------------------------------------------------------------------
# sequence-parallel collectives with their transposes as backward
# ---------------------------------------------------------------------------------------------------------------
class _GatherRows(torch.autograd.Function):
    """[T/tp, d] -> [T, d]; backward: reduce-scatter of the gradient."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        tp = dist.get_world_size(group) if group is not None else 1
        if tp == 1:
            return x
        out = x.new_empty((x.shape[0] * tp,) + tuple(x.shape[1:]))
        dist.all_gather_into_tensor(out, x.contiguous(), group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        tp = dist.get_world_size(ctx.group) if ctx.group is not None else 1
        if tp == 1:
            return g, None
        out = g.new_empty((g.shape[0] // tp,) + tuple(g.shape[1:]))
        dist.reduce_scatter_tensor(out, g.contiguous(), group=ctx.group)
        return out, None


class _ScatterRows(torch.autograd.Function):
    """Partial sums [T, d] on every rank -> summed rows [T/tp, d]; backward: all-gather of the gradient."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        tp = dist.get_world_size(group) if group is not None else 1
        if tp == 1:
            return x
        out = x.new_empty((x.shape[0] // tp,) + tuple(x.shape[1:]))
        dist.reduce_scatter_tensor(out, x.contiguous(), group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        tp = dist.get_world_size(ctx.group) if ctx.group is not None else 1
        if tp == 1:
            return g, None
        out = g.new_empty((g.shape[0] * tp,) + tuple(g.shape[1:]))
        dist.all_gather_into_tensor(out, g.contiguous(), group=ctx.group)
        return out, None


def _replicated(p: nn.Parameter) -> nn.Parameter:
    p._tp_replicated = True   # every tensor-parallel rank holds it in full and sees only its own token rows
    return p


def _rope(x: torch.Tensor, theta: float) -> torch.Tensor:
    """Rotate-half rotary embedding on x [b, S, H, dh] (pairs (i, i + dh/2)), positions 0..S-1."""
    b, S, H, dh = x.shape
    half = dh // 2
    inv = theta ** (-torch.arange(0, half, device=x.device, dtype=torch.float32) / half)
    ang = torch.arange(S, device=x.device, dtype=torch.float32)[:, None] * inv[None, :]
    cos, sin = ang.cos()[None, :, None, :].to(x.dtype), ang.sin()[None, :, None, :].to(x.dtype)
    x1, x2 = x[..., :half], x[..., half:]
    return torch.cat((x1 * cos - x2 * sin, x2 * cos + x1 * sin), dim=-1)


class TPBlock(nn.Module):
    """Pre-LN block:  x += drop(attn(norm1(x)));  x += drop(ffn(norm2(x)))  on token-row shards [T/tp, d]."""

    def __init__(self, margs, tp_group, tp_idx: int, tp: int, layer_id: int, init: str, dtype, device):
        super().__init__()
        d, f, H = margs.dim, margs.ffn_dim, margs.n_heads
        if H % tp or f % tp:
            raise ValueError(f"heads ({H}) and ffn width ({f}) must be divisible by the tensor-parallel degree {tp}")
        self.m, self.group, self.tp, self.Hl = margs, tp_group, tp, H // tp
        up = 2 if margs.activation == "swiglu" else 1
        Norm = (lambda: nn.RMSNorm(d, eps=margs.layer_norm_eps)) if margs.norm == "rmsnorm" else (lambda: nn.LayerNorm(d, eps=margs.layer_norm_eps))
        self.norm1, self.norm2 = Norm(), Norm()
        # full tensors from a per-layer seed, then this rank's slice (init='sliced': bit-identical to the unsharded model,
        # used by the parity test), or the slice directly (init='sharded': no 7B-parameter host tensors in the benchmark)
        g = torch.Generator().manual_seed(1000 + layer_id)

        def make(rows: int, cols: int, std: float, shard_dim: Optional[int], blocks: int = 1) -> nn.Parameter:
            if init == "sliced" or shard_dim is None:
                w = torch.randn(rows, cols, generator=g) * std
                if shard_dim == 0:     # column-parallel: `blocks` stacked sub-matrices (q|k|v, gate|up), each sliced by rows
                    w = torch.cat([c.chunk(tp, 0)[tp_idx] for c in w.chunk(blocks, 0)], 0)
                elif shard_dim == 1:   # row-parallel: input features
                    w = w.chunk(tp, 1)[tp_idx]
            else:
                gs = torch.Generator().manual_seed(1000 + layer_id + 7919 * (tp_idx + 1))
                w = torch.randn(rows // tp if shard_dim == 0 else rows, cols // tp if shard_dim == 1 else cols, generator=gs) * std
            return nn.Parameter(w.to(dtype).contiguous())

        s = 0.02
        self.wqkv = make(3 * d, d, s, 0, blocks=3)
        self.bqkv = nn.Parameter(torch.zeros(3 * d // tp, dtype=dtype))
        self.wo = make(d, d, s / math.sqrt(2 * margs.n_layers), 1)
        self.bo = _replicated(nn.Parameter(torch.zeros(d, dtype=dtype)))
        self.w1 = make(up * f, d, s, 0, blocks=up)
        self.b1 = nn.Parameter(torch.zeros(up * f // tp, dtype=dtype))
        self.w2 = make(d, f, s / math.sqrt(2 * margs.n_layers), 1)
        self.b2 = _replicated(nn.Parameter(torch.zeros(d, dtype=dtype)))
        for n_ in (self.norm1, self.norm2):
            n_.to(dtype)
            for p in n_.parameters():
                _replicated(p)
        self.to(device)
        self.seq_len = 0   # set by the stage module before every forward

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        m, S, Hl = self.m, self.seq_len, self.Hl
        dh = m.dim // m.n_heads
        p = m.dropout if self.training else 0.0
        hg = _GatherRows.apply(self.norm1(x), self.group)                     # [T, d]
        qkv = TF.linear(hg, self.wqkv, self.bqkv).view(-1, S, 3, Hl, dh)     # [b, S, 3, Hl, dh]
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        if m.rope:
            q, k = _rope(q, m.rope_theta), _rope(k, m.rope_theta)
        o = TF.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), dropout_p=p, is_causal=m.causal)
        o = o.transpose(1, 2).reshape(-1, Hl * dh)                            # [T, d/tp]
        a = _ScatterRows.apply(TF.linear(o, self.wo), self.group) + self.bo   # [T/tp, d]
        x = x + TF.dropout(a, p, self.training)
        hg = _GatherRows.apply(self.norm2(x), self.group)
        u = TF.linear(hg, self.w1, self.b1)
        if m.activation == "swiglu":
            gate, up_ = u.chunk(2, dim=-1)
            u = TF.silu(gate) * up_
        elif m.activation == "gelu":
            u = TF.gelu(u)
        else:
            u = TF.dropout(TF.relu(u), p, self.training)
        y = _ScatterRows.apply(TF.linear(u, self.w2), self.group) + self.b2
        return x + TF.dropout(y, p, self.training)


class TPStage(nn.Module):
    """One pipeline stage: [embedding] + blocks + [final norm + head]; activations are [b * S / tp, d] row shards."""

    def __init__(self, margs, layer_ids: List[int], first: bool, last: bool, tp_group, tp_idx: int, tp: int, init: str,
                 dtype, device):
        super().__init__()
        self.m, self.tp, self.tp_idx, self.first, self.last = margs, tp, tp_idx, first, last
        self.blocks = nn.ModuleList([TPBlock(margs, tp_group, tp_idx, tp, i, init, dtype, device) for i in layer_ids])
        g = torch.Generator().manual_seed(77)
        emb = torch.randn(margs.vocab_size, margs.dim, generator=g) * 0.02
        head = torch.randn(margs.vocab_size, margs.dim, generator=g) * 0.02
        if first:
            self.tok_embeddings = nn.Embedding(margs.vocab_size, margs.dim, dtype=dtype, device=device)
            with torch.no_grad():
                self.tok_embeddings.weight.copy_(emb)
            _replicated(self.tok_embeddings.weight)
        if last:
            self.norm = (nn.RMSNorm(margs.dim, eps=margs.layer_norm_eps) if margs.norm == "rmsnorm"
                         else nn.LayerNorm(margs.dim, eps=margs.layer_norm_eps)).to(device=device, dtype=dtype)
            self.output = nn.Linear(margs.dim, margs.vocab_size, dtype=dtype, device=device)
            with torch.no_grad():
                self.output.weight.copy_(head)
                self.output.bias.zero_()
            for p in list(self.norm.parameters()) + list(self.output.parameters()):
                _replicated(p)

    def local_rows(self, flat: torch.Tensor) -> torch.Tensor:
        n = flat.shape[0] // self.tp
        return flat[self.tp_idx * n:(self.tp_idx + 1) * n]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.first:                       # tokens [b, S] -> this rank's rows of the embedded sequence
            S = x.shape[1]
            x = self.tok_embeddings(self.local_rows(x.reshape(-1)))
            self._S = S
        S = getattr(self, "_S", None) or self.seq_len
        for blk in self.blocks:
            blk.seq_len = S
            x = blk(x)
        if self.last:
            x = self.output(self.norm(x))    # logits of the local rows [T/tp, V]
        return x


def layer_split(n_layers: int, stages: int) -> List[List[int]]:
    """The helper's partition rule (helper:70-75): n_layers // stages per stage, remainder to the last one."""
    per = n_layers // stages
    return [list(range(s * per, (s + 1) * per if s < stages - 1 else n_layers)) for s in range(stages)]


def build(margs, pp_idx: int, pp: int, tp_group, tp_idx: int, tp: int, seq_len: int, dtype, device, init: str = "sharded") -> TPStage:
    st = TPStage(margs, layer_split(margs.n_layers, pp)[pp_idx], pp_idx == 0, pp_idx == pp - 1, tp_group, tp_idx, tp, init, dtype, device)
    st.seq_len = seq_len
    return st


def make_loss_fn(stage: TPStage):
    """Mean token cross-entropy of the WHOLE micro-batch, of which this rank holds T/tp rows: local sum / T, so that the
    per-rank losses (and every gradient that crosses a reduce-scatter) add up to the global mean."""
    def loss_fn(logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        tgt = stage.local_rows(target.reshape(-1))
        return TF.cross_entropy(logits.float(), tgt, reduction="sum") / target.numel()
    return loss_fn


def reduce_replicated_grads(params, tp_group) -> None:
    if tp_group is None or dist.get_world_size(tp_group) == 1:
        return
    gs = [p.grad for p in params if getattr(p, "_tp_replicated", False) and p.grad is not None]
    if not gs:
        return
    flat = torch.cat([g.reshape(-1) for g in gs])
    dist.all_reduce(flat, group=tp_group)
    o = 0
    for g in gs:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


def make_groups(world: int, tp: int, rank: int):
    """Rank layout: tensor-parallel index fastest (rank = pp_idx * tp + tp_idx).  Returns (pp_group, tp_group)."""
    pp = world // tp
    pp_group = tp_group = None
    for t in range(tp):
        g = dist.new_group([p * tp + t for p in range(pp)])
        if t == rank % tp:
            pp_group = g
    for p in range(pp):
        g = dist.new_group([p * tp + t for t in range(tp)])
        if p == rank // tp:
            tp_group = g
    return pp_group, tp_group


def make_step(margs, schedule_name: str, n_microbatches: int, tp: int, rank: int, world: int, seq_len: int, dtype, device,
              init: str = "sharded", lr: float = 1e-4, fused_adam: bool = True):
    """Build stage + torch schedule + optimizer; returns (step(x, y) -> list of micro-batch losses on the last stage, stage)."""
    from torch.distributed.pipelining import PipelineStage, Schedule1F1B, ScheduleGPipe
    pp = world // tp
    pp_idx, tp_idx = rank // tp, rank % tp
    pp_group, tp_group = make_groups(world, tp, rank) if world > 1 else (None, None)
    stage_mod = build(margs, pp_idx, pp, tp_group, tp_idx, tp, seq_len, dtype, device, init)
    stage = PipelineStage(stage_mod, pp_idx, pp, device, group=pp_group)
    cls = {"GPipe": ScheduleGPipe, "1F1B": Schedule1F1B}[schedule_name]
    sched = cls(stage, n_microbatches=n_microbatches, loss_fn=make_loss_fn(stage_mod))
    params = list(stage_mod.parameters())
    opt = torch.optim.AdamW(params, lr=lr, betas=(0.9, 0.95), weight_decay=0.1, fused=fused_adam)

    def step(x: Optional[torch.Tensor], y: Optional[torch.Tensor]) -> List[torch.Tensor]:
        losses: List[torch.Tensor] = []
        if pp == 1:
            sched.step(x, target=y, losses=losses)
        elif pp_idx == 0:
            sched.step(x)
        elif pp_idx == pp - 1:
            sched.step(target=y, losses=losses)
        else:
            sched.step()
        reduce_replicated_grads(params, tp_group)
        opt.step()
        opt.zero_grad(set_to_none=False)
        return losses

    step.tp_group, step.pp_idx, step.pp = tp_group, pp_idx, pp
    return step, stage_mod


def train_steps(margs, schedule_name: str, n_microbatches: int, tp: int, batch: int, seq_len: int, steps: int,
                device: torch.device, dtype=torch.float32) -> List[float]:
    """`steps` training steps on fixed synthetic tokens; returns the global mean loss per step (same on every rank)."""
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist.is_initialized() else (0, 1)
    step, _ = make_step(margs, schedule_name, n_microbatches, tp, rank, world, seq_len, dtype, device, init="sliced",
                        fused_adam=device.type == "cuda")
    g = torch.Generator().manual_seed(5)
    x = torch.randint(0, margs.vocab_size, (batch, seq_len), generator=g).to(device)
    y = torch.randint(0, margs.vocab_size, (batch, seq_len), generator=g).to(device)
    out = []
    for _ in range(steps):
        losses = step(x, y)
        # a rank's micro-batch loss is its share (local sum / T) of the micro-batch mean: sum over the tp ranks of the last stage
        t = torch.zeros(1, dtype=torch.float64, device=device)
        if losses:
            t += torch.stack([l.detach().double() for l in losses]).mean()
        if world > 1:
            dist.all_reduce(t)
        out.append(float(t.item()))
    return out


def run(a, rank: int, local_rank: int, world: int, emit, config_dict, max_over_ranks) -> None:
    """The `--impl reference --tp N` arm of bench.py: same metric, same JSON line, stock torch on the whole path."""
    from ppb200.models.args import preset
    from ppb200.utils.clocks import ClockSampler
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    over = {"n_layers": a.n_layers} if a.n_layers is not None else {}
    margs = preset(a.model, **over)
    if a.small:
        margs = preset(a.model, dim=256, n_layers=max(2 * (world // a.tp), 4), n_heads=4, vocab_size=4096, ffn_dim=512)
    B, S = a.global_batch, a.seq_len
    step, stage_mod = make_step(margs, a.schedule, a.microbatches, a.tp, rank, world, S, torch.bfloat16, dev)
    g = torch.Generator().manual_seed(1000)
    x_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    y_host = torch.randint(0, margs.vocab_size, (B, S), generator=g).pin_memory()
    x, y = x_host.to(dev), y_host.to(dev)

    def one(e2e: bool = False):
        if e2e:
            if step.pp_idx == 0:
                x.copy_(x_host, non_blocking=True)
            if step.pp_idx == step.pp - 1:
                y.copy_(y_host, non_blocking=True)
        losses = step(x, y)
        if e2e and losses:
            return float(torch.stack(losses).float().sum().item())
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

    for _ in range(max(a.warmup, 3) if not a.small else max(a.warmup, 1)):
        one()
    sampler = ClockSampler(local_rank).start() if rank == 0 else None
    ms_total = timed(one, a.steps)
    clocks = sampler.stop() if sampler is not None else None
    e2e = None
    if not a.no_e2e:
        one(True)
        ms_e2e = timed(lambda: one(True), a.steps)
        e2e = {"value": B * S * a.steps / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / a.steps,
               "h2d_bytes_per_step": int(a.tp * 2 * B * S * 8), "d2h_bytes_per_step": 4}
    peak = max_over_ranks(torch.cuda.max_memory_allocated(dev) / 2**30, dev)
    if rank == 0:
        ms = ms_total / a.steps
        par = f"pp{world // a.tp}xtp{a.tp}"
        emit({"metric": "tokens/sec (1F1B pipeline training step, max over ranks)" if a.schedule == "1F1B" else f"tokens/sec ({a.schedule})",
              "value": B * S / (ms / 1e3), "unit": "tokens/s", "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
              "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "bf16",
              "data": "synthetic random tokens, random-init weights", "impl": "reference",
              "config": config_dict(a, {"dim": margs.dim, "n_layers": margs.n_layers, "n_heads": margs.n_heads,
                                        "vocab_size": margs.vocab_size, "ffn_dim": margs.ffn_dim, "dropout": margs.dropout,
                                        "block": a.model}, par),
              "impl_notes": {"stack": "stock torch: tensor + sequence parallel block (all_gather_into_tensor / reduce_scatter_tensor over NCCL, "
                                      "cuBLAS linears, SDPA) inside torch.distributed.pipelining, fused AdamW; the helper itself has no "
                                      "tensor parallelism (helper:178) -- benchmarks/tp_baseline.py",
                             "partition": "helper rule (n_layers // stages, remainder to the last stage)"},
              "peak_mem_gib": peak, "clocks": clocks, "e2e": e2e, "gpu_launches": None})
    dist.barrier()
    if not getattr(a, "keep_process_group", False):
        dist.destroy_process_group()
