#!/usr/bin/env python
"""Per-kernel micro-benchmarks (CUDA events, L2 flushed between iterations) with roofline fractions against
MEASURED_PEAKS.json.  Writes gpurun_out/kernel_bench.json; summaries get copied to profiles/."""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.ops import functional as F  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


_flush = None


def time_fn(fn, iters=20, warmup=5, flush=True):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            _flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def bench_gemm(rows):
    hbm, tf, tag = peaks()
    shapes = [  # (name, M, N, K, a_mn, b_mn, out_mode)
        ("fwd qkv  T4096 d1024", 4096, 3072, 1024, False, False, 0),
        ("fwd proj T4096 d1024", 4096, 1024, 1024, False, False, 0),
        ("fwd ffn1 T4096", 4096, 2048, 1024, False, False, 0),
        ("fwd ffn2 T4096", 4096, 1024, 2048, False, False, 0),
        ("fwd head T4096 V50257", 4096, 50257, 1024, False, False, 0),
        ("dgrad qkv T4096", 4096, 1024, 3072, False, True, 0),
        ("dgrad head T4096", 4096, 1024, 50257, False, True, 0),
        ("wgrad qkv T4096", 3072, 1024, 4096, True, True, 2),
        ("wgrad proj T4096", 1024, 1024, 4096, True, True, 2),
        ("wgrad head T4096", 50257, 1024, 4096, True, True, 2),
        ("square 8192", 8192, 8192, 8192, False, False, 0),
        ("ref-small qkv T1024 d768", 1024, 2304, 768, False, False, 0),
    ]
    for name, M, N, K, amn, bmn, om in shapes:
        a = torch.randn((K, M) if amn else (M, K), device="cuda", dtype=torch.bfloat16)
        b = torch.randn((K, N) if bmn else (N, K), device="cuda", dtype=torch.bfloat16)
        ldc = F.pad8(N)
        c = torch.zeros(M, ldc, device="cuda", dtype=torch.float32 if om else torch.bfloat16)[:, :N]
        for bn in (0, 128, 256):
            try:
                ms = time_fn(lambda: F.gemm(a, b, c, M, N, K, amn, bmn, om, force_bn=bn))
            except Exception as ex:  # noqa: BLE001
                rows.append({"kernel": "gemm", "name": name, "bn": bn, "error": str(ex)})
                continue
            tfl = 2.0 * M * N * K / ms / 1e9
            af, bf = (a.float().t() if amn else a.float()), (b.float().t() if bmn else b.float())
            ta, tb = af.to(torch.bfloat16).contiguous(), bf.to(torch.bfloat16).contiguous()
            ms_ref = time_fn(lambda: torch.matmul(ta, tb.t())) if bn == 0 else None
            rows.append({"kernel": "gemm", "name": name, "bn": bn, "ms": round(ms, 4), "tflops": round(tfl, 1),
                         "frac_of_" + tag + "_peak": round(tfl / tf, 3),
                         "cublas_ms": None if ms_ref is None else round(ms_ref, 4)})
            print(rows[-1], flush=True)


def bench_bandwidth(rows):
    hbm, tf, tag = peaks()
    T, d, V = 4096, 1024, 50257
    x = torch.randn(T, d, device="cuda", dtype=torch.bfloat16)
    a = torch.randn(T, d, device="cuda", dtype=torch.bfloat16)
    g = torch.ones(d, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(d, device="cuda", dtype=torch.bfloat16)
    ms = time_fn(lambda: F.add_dropout_norm_fwd(x, a, g, b, 1e-5, 0.1, 1))
    rows.append({"kernel": "add_dropout_ln_fwd", "T": T, "d": d, "ms": round(ms, 4),
                 "gbs": round(4 * T * d * 2 / ms / 1e6, 1), "frac_of_" + tag + "_hbm": round(4 * T * d * 2 / ms / 1e6 / hbm, 3)})
    print(rows[-1], flush=True)
    z, y, mean, rstd = F.add_dropout_norm_fwd(x, a, g, b, 1e-5, 0.1, 1)
    dg, db = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
    ms = time_fn(lambda: F.norm_bwd([x, a], z, mean, rstd, g, dg, db, 0.1, 1))
    rows.append({"kernel": "ln_bwd(2 dy)", "T": T, "d": d, "ms": round(ms, 4), "gbs": round(5 * T * d * 2 / ms / 1e6, 1),
                 "frac_of_" + tag + "_hbm": round(5 * T * d * 2 / ms / 1e6 / hbm, 3)})
    print(rows[-1], flush=True)
    ld = F.pad8(V)
    logits = torch.randn(T, ld, device="cuda", dtype=torch.bfloat16)[:, :V]
    tgt = torch.randint(0, V, (T,), device="cuda")
    ms = time_fn(lambda: F.xent_fwd_bwd(logits, tgt, 1.0))
    rows.append({"kernel": "xent_fwd_bwd", "T": T, "V": V, "ms": round(ms, 4), "gbs": round(2 * T * V * 2 / ms / 1e6, 1),
                 "frac_of_" + tag + "_hbm": round(2 * T * V * 2 / ms / 1e6 / hbm, 3)})
    print(rows[-1], flush=True)
    n = 300_000_000
    master, m, v, grad = (torch.zeros(n, device="cuda") for _ in range(4))
    param = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    ms = time_fn(lambda: F.adamw_step(param, master, m, v, grad, 1e-3, 0.9, 0.95, 1e-8, 0.1, 1), flush=False)
    byts = n * (4 * 4 + 4 * 4 + 2)
    rows.append({"kernel": "adamw", "n": n, "ms": round(ms, 4), "gbs": round(byts / ms / 1e6, 1),
                 "frac_of_" + tag + "_hbm": round(byts / ms / 1e6 / hbm, 3)})
    print(rows[-1], flush=True)


ATTN_SHAPES = [(4, 16, 1024, 64, False, 0.0), (4, 16, 1024, 64, False, 0.1), (4, 16, 1024, 64, True, 0.0),
               (2, 32, 2048, 128, True, 0.0), (8, 8, 128, 96, False, 0.1)]
# head dim 192 = the 4-head rows of the reference grid (dim 768; one micro-batch there is 8 x 128 tokens); run once as is and
# once with PPB_ATTN_TC_MAX_DH=128 (CUDA-core path) for the A/B
ATTN_WIDE_SHAPES = [(8, 4, 128, 192, False, 0.1), (4, 4, 1024, 192, False, 0.1), (4, 4, 1024, 192, True, 0.0)]


def bench_attention(rows, shapes=ATTN_SHAPES):
    hbm, tf, tag = peaks()
    for (B, H, S, dh, causal, p) in shapes:
        D = H * dh
        qkv = torch.randn(B * S, 3 * D, device="cuda", dtype=torch.bfloat16)
        q, k, v = qkv[:, :D], qkv[:, D:2 * D], qkv[:, 2 * D:]
        try:
            o, lse = F.attention_fwd(q, k, v, B, H, causal, None, p, 1)
            ms = time_fn(lambda: F.attention_fwd(q, k, v, B, H, causal, None, p, 1, out=o, lse=lse))
            fl = 4.0 * B * H * S * S * dh * (0.5 if causal else 1.0)
            rows.append({"kernel": "attn_fwd", "B": B, "H": H, "S": S, "dh": dh, "causal": causal, "p": p,
                         "ms": round(ms, 4), "tflops": round(fl / ms / 1e9, 1), "frac_of_" + tag + "_peak": round(fl / ms / 1e9 / tf, 3)})
            print(rows[-1], flush=True)
            do = torch.randn_like(o)
            dqkv = torch.empty_like(qkv)
            outs = (dqkv[:, :D], dqkv[:, D:2 * D], dqkv[:, 2 * D:])
            ms = time_fn(lambda: F.attention_bwd(q, k, v, o, do, lse, B, H, causal, None, p, 1, out=outs))
            rows.append({"kernel": "attn_bwd", "B": B, "H": H, "S": S, "dh": dh, "causal": causal, "p": p,
                         "ms": round(ms, 4), "tflops": round(2.5 * fl / ms / 1e9, 1),
                         "frac_of_" + tag + "_peak": round(2.5 * fl / ms / 1e9 / tf, 3)})
            print(rows[-1], flush=True)
        except Exception as ex:  # noqa: BLE001
            rows.append({"kernel": "attn", "B": B, "H": H, "S": S, "dh": dh, "error": str(ex)[:300]})
            print(rows[-1], flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="gemm,bw,attn")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "kernel_bench.json"))
    a = ap.parse_args()
    rows = []
    if "gemm" in a.what.split(","):
        bench_gemm(rows)
    if "bw" in a.what.split(","):
        bench_bandwidth(rows)
    if "attn" in a.what.split(","):
        bench_attention(rows)
    if "attnwide" in a.what.split(","):
        bench_attention(rows, ATTN_WIDE_SHAPES)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)
