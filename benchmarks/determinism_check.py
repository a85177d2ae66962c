#!/usr/bin/env python
"""Kernel-library self-check on one GPU (no compute-sanitizer needed):
  1. every op, same inputs, buffers re-allocated at shifting addresses, many repetitions -> outputs must be bit-identical
     (ops that reduce in arrival order -- split-K weight gradients, dQ across kv tiles for S > 128 -- are reported with
     their fp32-rounding-level spread);
  2. one eager training step under guarded allocations (canary bands + NaN poison, utils/debug.py) -> no
     out-of-bounds write, no NaN.
Writes gpurun_out/determinism_check.json."""
import copy
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.api import make_schedule, manual_model_split  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402
from ppb200.ops import functional as F  # noqa: E402
from ppb200.utils import debug  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
out = {"ops": {}}


def run(name, fn, reps=30):
    r = debug.determinism_check(fn, reps=reps)
    out["ops"][name] = r
    print(f"{name:34s} mismatching={r['mismatching']:3d}/{reps} worst_abs={r['worst_abs']:.3e} nan={r['nan']}", flush=True)


for (B, H, S, dh) in [(2, 4, 128, 64), (4, 16, 1024, 64), (2, 8, 512, 128)]:
    d, T = H * dh, B * S
    qkv0 = torch.randn(T, 3 * d, device=dev).bfloat16()
    do0 = torch.randn(T, d, device=dev).bfloat16()
    for p in (0.0, 0.1):
        def fwd():
            qkv = qkv0.clone()
            return F.attention_fwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], B, H, True, None, p, 1234)
        run(f"attn_fwd B{B} H{H} S{S} dh{dh} p{p}", fwd)

        def bwd():
            qkv, do = qkv0.clone(), do0.clone()
            o, lse = F.attention_fwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], B, H, True, None, p, 1234)
            dqkv = torch.empty_like(qkv)
            F.attention_bwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], o, do, lse, B, H, True, None, p, 1234,
                            out=(dqkv[:, :d], dqkv[:, d:2 * d], dqkv[:, 2 * d:]))
            return (dqkv,)
        run(f"attn_bwd B{B} H{H} S{S} dh{dh} p{p}", bwd)

T, d, ff, V = 512, 512, 1024, 1000
x = torch.randn(T, d, device=dev).bfloat16()
a = torch.randn(T, d, device=dev).bfloat16()
g = torch.randn(d, device=dev).bfloat16()
b = torch.randn(d, device=dev).bfloat16()
w1 = (torch.randn(ff, d, device=dev) * 0.05).bfloat16()
b1 = torch.randn(ff, device=dev).bfloat16()
dy = torch.randn(T, d, device=dev).bfloat16()
dyf = torch.randn(T, ff, device=dev).bfloat16()
for p in (0.0, 0.1):
    run(f"add_dropout_norm_fwd p{p}", lambda: F.add_dropout_norm_fwd(x.clone(), a.clone(), g, b, 1e-5, p, 77, False))
    z, y, mean, rstd = F.add_dropout_norm_fwd(x, a, g, b, 1e-5, p, 77, False)

    def lnb():
        dg, db = torch.zeros(d, device=dev), torch.zeros(d, device=dev)
        dz, da = F.norm_bwd([dy.clone(), None, None], z.clone(), mean.clone(), rstd.clone(), g, dg, db, p, 77, False)
        return dz, da, dg, db
    run(f"norm_bwd p{p}", lnb)
    run(f"linear relu+dropout p{p}", lambda: (F.linear_fwd(x.clone(), w1, b1, "relu", p, 99),))
    run(f"sum_dropout_bwd p{p}", lambda: F.sum_dropout_bwd([dy.clone(), x.clone()], p, 55))


def wg():
    dw, dbb = torch.zeros(ff, d, device=dev), torch.zeros(ff, device=dev)
    F.linear_wgrad(dyf.clone(), x.clone(), dw, dbb)
    return dw, dbb


run("linear_wgrad (split-K) dW+db", wg)
F.DETERMINISTIC = True      # covers the weight gradient; the bias column-sum still lands with fp32 atomics
run("linear_wgrad (PPB_DETERMINISTIC) dW", lambda: wg()[:1])
run("linear_wgrad bias colsum (atomics)", lambda: wg()[1:])
F.DETERMINISTIC = False
run("linear_dgrad mask", lambda: (F.linear_dgrad(dyf.clone(), w1, mask_pos_aux=dyf, alpha=1.0 / 0.9),))


def xe():
    lg = torch.randn(T, V, device=dev, generator=torch.Generator(device=dev).manual_seed(3)).bfloat16()
    tg = torch.randint(0, V, (T,), device=dev, generator=torch.Generator(device=dev).manual_seed(4))
    return F.xent_fwd_bwd(lg, tg, 1.0 / T, True), lg


run("xent_fwd_bwd", xe)
gu = torch.randn(T, 2 * ff, device=dev).bfloat16()
run("swiglu fwd+bwd", lambda: (F.swiglu_fwd(gu.clone()), F.swiglu_bwd(dyf.clone(), gu.clone())))
run("rope", lambda: (F.rope_(torch.cat([x, a, dy], 1).clone(), 0, 128, 2 * 8, 64),))

# ---- one eager training step under guard bands + poison ------------------------------------------------------------
args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
full = Transformer(args)
xt, yt = torch.randint(0, 512, (8, 128)), torch.randint(0, 512, (8, 128))
with debug.guarded_allocations() as rep:
    st = manual_model_split(copy.deepcopy(full), 0, 1, dev)
    sched = make_schedule("1F1B", st, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0}, cuda_graphs=False)
    bad, losses = 0, []
    for it in range(2):
        losses = []
        sched.step(xt.cuda(), target=yt.cuda(), losses=losses)
        bad += rep.check(f"step {it}")
    nan_params = int(torch.isnan(st.master[:st.n_total]).sum())
out["guarded_step"] = {"buffers_checked": rep.checked, "violated": bad, "violations": rep.violations[:5], "nan_params": nan_params,
                       "loss": [round(float(l.item()), 4) for l in losses]}
print(json.dumps(out["guarded_step"]))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "determinism_check.json"), "w"), indent=1)
