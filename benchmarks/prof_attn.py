"""Tiny driver for ncu: a few launches of the attention forward/backward kernels at the flagship shape."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.ops import functional as F

B, H, S, dh = 4, 16, 1024, 64
D = H * dh
p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.1
qkv = torch.randn(B * S, 3 * D, device="cuda", dtype=torch.bfloat16)
q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
o, lse = F.attention_fwd(q, k, v, B, H, False, None, p, 1)
do = torch.randn_like(o)
dqkv = torch.empty_like(qkv)
outs = (dqkv[:, :D], dqkv[:, D:2 * D], dqkv[:, 2 * D:])
for _ in range(3):
    F.attention_fwd(q, k, v, B, H, False, None, p, 1, out=o, lse=lse)
    F.attention_bwd(q, k, v, o, do, lse, B, H, False, None, p, 1, out=outs)
torch.cuda.synchronize()
print("ok")
