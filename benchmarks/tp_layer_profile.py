#!/usr/bin/env python
"""Per-op device-time breakdown of one tensor-parallel decoder layer (forward + backward), CUDA events around every
op of models/tp.py.  torchrun --nproc-per-node 2 benchmarks/tp_layer_profile.py [--model llama-7b] [--tokens 8192]"""
import argparse
import collections
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.models import preset  # noqa: E402
from ppb200.models.tp import shard_model_for_tp  # noqa: E402
from ppb200.models.transformer import Ctx, GradSink, Transformer, layer_backward, layer_forward  # noqa: E402
from ppb200.ops import functional as F  # noqa: E402
from ppb200.ops.collective import TPComm  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="llama-7b")
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--seq", type=int, default=2048)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
args = preset(a.model, n_layers=1, dtype=torch.bfloat16)
torch.manual_seed(0)
with torch.device(dev):
    model = Transformer(args)
model.tok_embeddings = model.norm = model.output = None
shard_model_for_tp(model, rank, world)
model.to(torch.bfloat16)
L = model.layers["0"]
G = GradSink()
for p in L.parameters():
    G.bind(p, torch.zeros(p.shape, dtype=torch.float32, device=dev))
T = a.batch * a.seq
tp = TPComm(dist.group.WORLD, T // world, args.dim, dev)
ctx = Ctx(args, a.batch, a.seq, 0, 0, True, 0)
ctx.tp = tp
x = torch.randn(T // world, args.dim, device=dev).bfloat16()
dy = torch.randn(T // world, args.dim, device=dev).bfloat16()

records = []


def timed(name, fn):
    def w(*args, **kw):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = fn(*args, **kw)
        e1.record()
        records.append((name, e0, e1))
        return r
    return w


for name in ("add_dropout_norm_fwd", "attention_fwd", "attention_bwd", "norm_bwd", "sum_dropout_bwd", "add_dropout_fwd",
             "swiglu_fwd", "swiglu_bwd", "gelu_fwd", "rope_", "linear_wgrad", "colsum_accum"):
    setattr(F, name, timed(name, getattr(F, name)))
tp.ag_linear = timed("ag_linear(K12)+gather", tp.ag_linear)
tp.linear_rs = timed("linear_rs(K11)+barrier", tp.linear_rs)

tot = collections.defaultdict(float)
phase_ms = {"fwd": 0.0, "bwd": 0.0}
for rep in range(a.reps + 2):
    records.clear()
    st = {}
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    with torch.no_grad():
        e[0].record()
        layer_forward(L, 0, x, ctx, st)
        e[1].record()
        layer_backward(L, 0, [dy], ctx, st, G)
        e[2].record()
    torch.cuda.synchronize()
    if rep >= 2:
        phase_ms["fwd"] += e[0].elapsed_time(e[1]) / a.reps
        phase_ms["bwd"] += e[1].elapsed_time(e[2]) / a.reps
        for name, s0, s1 in records:
            tot[name] += s0.elapsed_time(s1) / a.reps
if rank == 0:
    out = {"model": a.model, "tokens": T, "tp": world, "layer_fwd_ms": round(phase_ms["fwd"], 3), "layer_bwd_ms": round(phase_ms["bwd"], 3),
           "ops_ms": {k: round(v, 3) for k, v in sorted(tot.items(), key=lambda kv: -kv[1])}}
    out["accounted_ms"] = round(sum(tot.values()), 3)
    d, f = args.dim, args.ffn_dim
    gemm_flops = 2 * T * (4 * d * d + (3 if args.activation == "swiglu" else 2) * d * f) / world
    out["fwd_gemm_tflops_at_layer_time"] = round(gemm_flops / (phase_ms["fwd"] * 1e-3) / 1e12, 1)
    print(json.dumps(out))
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"tp_layer_profile_{a.model}.json"), "w"), indent=1)
dist.barrier(device_ids=[local])
dist.destroy_process_group()
