#!/usr/bin/env python
"""One small launch of every kernel family, for `compute-sanitizer`:

    compute-sanitizer --tool memcheck  python benchmarks/sanitize_smoke.py
    compute-sanitizer --tool racecheck python benchmarks/sanitize_smoke.py     (one tool per GPU call)

Shapes are small (the sanitizer slows kernels down 10-100x) but cover ragged edges (TMA out-of-bounds fill / clipping),
every GEMM operand layout, one-CTA and CTA-pair tiles, the fused gather / reduce-scatter / all-reduce kernels with the
"peers" emulated by separate buffers of this GPU, attention forward / backward with dropout and a causal mask, the
norm kernels, the loss, the optimizer and the embedding kernels.  Every result is also checked against torch, so a
clean sanitizer log belongs to a run that computed the right numbers."""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.ops import _lib, functional as F, reference as R  # noqa: E402
from ppb200.ops.collective import (GatherFlags, allgather_linear_fused, gemm_allreduce, linear_reducescatter, push_add)  # noqa: E402

DEV = "cuda"


def rel(a, b):
    return ((a.float() - b.float()).norm() / (b.float().norm() + 1e-12)).item()


def main():
    torch.manual_seed(0)
    lib = _lib.load(required=True)
    lib.ppb_set_gemm_cta_group.argtypes = [ctypes.c_int]
    checks = {}
    # ---- GEMM: fwd / dgrad / wgrad layouts, ragged shapes, one CTA per tile and CTA pairs ----
    for cg in (1, 2):
        lib.ppb_set_gemm_cta_group(cg)
        M, N, K = 392, 264, 136
        a, b = torch.randn(M, K, device=DEV).bfloat16(), torch.randn(N, K, device=DEV).bfloat16()
        c = torch.zeros(M, F.pad8(N), device=DEV, dtype=torch.bfloat16)[:, :N]
        bias = torch.randn(N, device=DEV).bfloat16()
        F.gemm(a, b, c, M, N, K, False, False, epi=F.EPI_BIAS_RELU, bias=bias)
        checks[f"gemm_fwd_cg{cg}"] = rel(c, torch.relu(a.float() @ b.float().t() + bias.float()))
        w = torch.randn(K, N, device=DEV).bfloat16()
        F.gemm(a, w, c, M, N, K, False, True)
        checks[f"gemm_dgrad_cg{cg}"] = rel(c, a.float() @ w.float())
        dy, x = torch.randn(K, M, device=DEV).bfloat16(), torch.randn(K, N, device=DEV).bfloat16()
        acc = torch.ones(M, N, device=DEV)
        F.gemm(dy, x, acc, M, N, K, True, True, F.OUT_F32_ACCUM)
        checks[f"gemm_wgrad_cg{cg}"] = rel(acc, 1 + dy.float().t() @ x.float())
    lib.ppb_set_gemm_cta_group(0)
    # ---- fused collectives with the peers emulated by buffers of this GPU ----
    tp, rows, N, K = 2, 256, 264, 128
    shards = [torch.randn(rows, K, device=DEV).bfloat16() for _ in range(tp)]
    w = (torch.randn(N, K, device=DEV) * 0.1).bfloat16()
    full = torch.empty(rows * tp, K, device=DEV, dtype=torch.bfloat16)
    gfl = GatherFlags(rows * tp, torch.device(DEV))
    for it in range(2):
        y = allgather_linear_fused([s.data_ptr() for s in shards], rows, K, full, gfl, w, 0)
    checks["k12_allgather_gemm"] = rel(y, torch.cat(shards).float() @ w.float().t())
    dsts = [torch.zeros(rows, N, device=DEV) for _ in range(tp)]
    xx = torch.randn(rows * tp, K, device=DEV).bfloat16()
    linear_reducescatter(xx, w, [d.data_ptr() for d in dsts], N)
    checks["k11_gemm_reducescatter"] = rel(torch.cat(dsts), xx.float() @ w.float().t())
    dy, x = torch.randn(520, 200, device=DEV).bfloat16(), torch.randn(520, 136, device=DEV).bfloat16()
    acc = torch.zeros(200, 136, device=DEV)
    outs = [torch.zeros(200, 136, device=DEV) for _ in range(2)]
    gemm_allreduce(dy, x, acc, [o.data_ptr() for o in outs])
    checks["k10_wgrad_allreduce"] = rel(outs[1], dy.float().t() @ x.float())
    src = torch.randn(1024, device=DEV)
    dst = [torch.zeros(1024, device=DEV)]
    push_add(src, [dst[0].data_ptr()], 2.0)
    checks["push_add"] = rel(dst[0], 2 * src)
    # ---- attention (dropout, causal, ragged sequence), fwd + bwd ----
    B, H, S, dh = 2, 2, 200, 64
    D = H * dh
    qkv = torch.randn(B * S, 3 * D, device=DEV).bfloat16()
    q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
    o, lse = F.attention_fwd(q, k, v, B, H, True, None, 0.1, 5)
    orf, _ = R.attention_fwd(q.float().cpu(), k.float().cpu(), v.float().cpu(), B, H, True, None, 0.1, 5)
    checks["attn_fwd"] = rel(o.cpu(), orf)
    do = torch.randn_like(o)
    dq, dk, dv = F.attention_bwd(q, k, v, o, do, lse, B, H, True, None, 0.1, 5)
    dqr, dkr, dvr = R.attention_bwd(q.float().cpu(), k.float().cpu(), v.float().cpu(), o.float().cpu(), do.float().cpu(),
                                    lse.cpu(), B, H, True, None, 0.1, 5)
    checks["attn_bwd"] = max(rel(dq.cpu(), dqr), rel(dk.cpu(), dkr), rel(dv.cpu(), dvr))
    # ---- norm fwd / bwd, loss, optimizer, embedding ----
    T, d = 300, 768
    resid, a = torch.randn(T, d, device=DEV).bfloat16(), torch.randn(T, d, device=DEV).bfloat16()
    g, be = torch.randn(d, device=DEV).bfloat16(), torch.randn(d, device=DEV).bfloat16()
    z, yv, mean, rstd = F.add_dropout_norm_fwd(resid, a, g, be, 1e-5, 0.1, 9)
    zr, yr, _, _ = R.add_dropout_norm_fwd(resid.cpu(), a.cpu(), g.cpu(), be.cpu(), 1e-5, 0.1, 9)
    checks["ln_fwd"] = rel(yv.cpu(), yr)
    dgam, dbet = torch.zeros(d, device=DEV), torch.zeros(d, device=DEV)
    dz, da = F.norm_bwd([a, resid], z, mean, rstd, g, dgam, dbet, 0.1, 9)
    checks["ln_bwd_finite"] = 0.0 if bool(torch.isfinite(dz.float()).all() and torch.isfinite(dgam).all()) else 1.0
    V = 1003
    logits = torch.randn(T, F.pad8(V), device=DEV).bfloat16()[:, :V]
    tgt = torch.randint(0, V, (T,), device=DEV)
    ref_loss = torch.nn.functional.cross_entropy(logits.float(), tgt, reduction="none")
    loss = F.xent_fwd_bwd(logits, tgt, 1.0)
    checks["xent"] = rel(loss, ref_loss)
    n = 4096
    master, m, vv, grad = torch.randn(n, device=DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV), torch.randn(n, device=DEV)
    param = master.bfloat16()
    F.adamw_step(param, master, m, vv, grad, 1e-3, 0.9, 0.95, 1e-8, 0.1, 1)
    checks["adamw_finite"] = 0.0 if bool(torch.isfinite(master).all()) else 1.0
    emb = torch.randn(V, d, device=DEV).bfloat16()
    e = F.embedding_fwd(tgt, emb)
    checks["embedding"] = rel(e, emb[tgt])
    gw = torch.zeros(V, d, device=DEV)
    F.embedding_bwd(tgt, [e], gw)
    torch.cuda.synchronize()
    bad = {k: v for k, v in checks.items() if not (v < 3e-2)}
    print("checks:", {k: round(v, 5) for k, v in checks.items()})
    if bad:
        raise SystemExit(f"numerical check failed: {bad}")
    print("sanitize_smoke ok")


if __name__ == "__main__":
    main()
