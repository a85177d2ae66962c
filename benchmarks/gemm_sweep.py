"""Sweep (BN, splits) for the GEMM shapes of the flagship model; prints the best config per shape and what the
built-in heuristic picks (used to calibrate choose_config in csrc/gemm_tcgen05.cu)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.ops import functional as F
from benchmarks.kernel_bench import time_fn

T = 4096
shapes = [("wgrad 3072x1024", 3072, 1024, T, True, True, 2), ("wgrad 1024x1024", 1024, 1024, T, True, True, 2),
          ("wgrad 2048x1024", 2048, 1024, T, True, True, 2), ("wgrad 1024x2048", 1024, 2048, T, True, True, 2),
          ("wgrad 50257x1024", 50257, 1024, T, True, True, 2),
          ("fwd 4096x3072x1024", T, 3072, 1024, False, False, 0), ("fwd 4096x1024x1024", T, 1024, 1024, False, False, 0),
          ("fwd 4096x2048x1024", T, 2048, 1024, False, False, 0), ("fwd 4096x1024x2048", T, 1024, 2048, False, False, 0),
          ("dgrad 4096x1024x3072", T, 1024, 3072, False, True, 0), ("dgrad 4096x1024x1024", T, 1024, 1024, False, True, 0),
          ("dgrad 4096x2048x1024", T, 2048, 1024, False, True, 0), ("dgrad 4096x1024x2048", T, 1024, 2048, False, True, 0)]
out = []
for name, M, N, K, amn, bmn, om in shapes:
    a = torch.randn((K, F.pad8(M)) if amn else (M, K), device="cuda", dtype=torch.bfloat16)
    a = a[:, :M] if amn else a
    b = torch.randn((K, F.pad8(N)) if bmn else (N, K), device="cuda", dtype=torch.bfloat16)
    b = b[:, :N] if bmn else b
    c = torch.zeros(M, F.pad8(N), device="cuda", dtype=torch.float32 if om else torch.bfloat16)[:, :N]
    res = {}
    auto = time_fn(lambda: F.gemm(a, b, c, M, N, K, amn, bmn, om), iters=10, warmup=3)
    for bn in (64, 128, 256):
        for sp in ([1, 2, 3, 4, 6, 8, 12, 16] if om == 2 else [1]):
            try:
                res[(bn, sp)] = time_fn(lambda: F.gemm(a, b, c, M, N, K, amn, bmn, om, force_bn=bn, force_splits=sp), iters=10, warmup=3)
            except Exception as e:
                pass
    best = min(res, key=res.get)
    row = {"shape": name, "auto_ms": round(auto, 4), "best": best, "best_ms": round(res[best], 4),
           "all": {f"{k[0]}x{k[1]}": round(v, 4) for k, v in sorted(res.items())}}
    print(json.dumps(row), flush=True)
    out.append(row)
json.dump(out, open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out", "gemm_sweep.json"), "w"), indent=1)
