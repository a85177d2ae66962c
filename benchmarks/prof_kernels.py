"""Tiny driver for ncu: a few launches of the GEMM (fwd/dgrad/wgrad shapes), LayerNorm fwd/bwd, cross-entropy and
AdamW kernels at the flagship shapes (T=4096 tokens, d=1024, V=50257)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.ops import functional as F

T, d, f, V = 4096, 1024, 2048, 50257
x = torch.randn(T, d, device="cuda", dtype=torch.bfloat16)
w = torch.randn(3 * d, d, device="cuda", dtype=torch.bfloat16) * 0.02
b = torch.zeros(3 * d, device="cuda", dtype=torch.bfloat16)
dy = torch.randn(T, 3 * d, device="cuda", dtype=torch.bfloat16)
dw = torch.zeros(3 * d, d, device="cuda")
g = torch.ones(d, device="cuda", dtype=torch.bfloat16)
be = torch.zeros(d, device="cuda", dtype=torch.bfloat16)
dg, db = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
logits = torch.randn(T, F.pad8(V), device="cuda", dtype=torch.bfloat16)[:, :V]
tgt = torch.randint(0, V, (T,), device="cuda")
n = 50_000_000
master, m, v, grad = (torch.zeros(n, device="cuda") for _ in range(4))
param = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    y = F.linear_fwd(x, w, b)                       # gemm fwd  4096x3072x1024 (+bias)
    dx = F.linear_dgrad(dy, w)                      # gemm dgrad 4096x1024x3072
    F.linear_wgrad(dy, x, dw)                       # gemm wgrad 3072x1024x4096 (split-K, TMA reduce-add)
    z, yn, mean, rstd = F.add_dropout_norm_fwd(x, dx, g, be, 1e-5, 0.1, 1)
    F.norm_bwd([x, dx], z, mean, rstd, g, dg, db, 0.1, 1)
    F.xent_fwd_bwd(logits, tgt, 1.0)
    F.adamw_step(param, master, m, v, grad, 1e-3, 0.9, 0.95, 1e-8, 0.1, 1)
torch.cuda.synchronize()
print("ok")
