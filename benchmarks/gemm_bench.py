#!/usr/bin/env python
"""GEMM table for the flagship layer shapes (T = 4096 tokens per micro-batch, d = 1024, helper LM at GPT-2-medium):
our tcgen05 kernel with one CTA per tile (cta_group::1), with CTA pairs (cta_group::2) and with the automatic choice,
next to cuBLAS (torch.matmul) on the same operands, plus the maximum error against an fp32 product.

    python benchmarks/gemm_bench.py --out gpurun_out/gemm_bench.json

CUDA events, L2 flushed between iterations (a 256 MiB memset), median of 20.  Fractions are of the MEASURED cuBLAS
bf16 burst peak in MEASURED_PEAKS.json.  The LM-head shapes use the row-padded logits buffer the model uses
(leading dimension 50264 for V = 50257)."""
import argparse
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from benchmarks.kernel_bench import peaks, time_fn  # noqa: E402
from ppb200.ops import _lib, functional as F  # noqa: E402

T, D, FF, V = 4096, 1024, 2048, 50257
SHAPES = [  # name, M, N, K, a_mn, b_mn, out_mode
    ("fwd qkv", T, 3 * D, D, False, False, 0),
    ("fwd proj", T, D, D, False, False, 0),
    ("fwd ffn1", T, FF, D, False, False, 0),
    ("fwd ffn2", T, D, FF, False, False, 0),
    ("fwd head", T, V, D, False, False, 0),
    ("dgrad qkv", T, D, 3 * D, False, True, 0),
    ("dgrad proj", T, D, D, False, True, 0),
    ("dgrad ffn2", T, FF, D, False, True, 0),
    ("dgrad ffn1", T, D, FF, False, True, 0),
    ("dgrad head", T, D, V, False, True, 0),
    ("wgrad qkv", 3 * D, D, T, True, True, 2),
    ("wgrad proj", D, D, T, True, True, 2),
    ("wgrad ffn1", FF, D, T, True, True, 2),
    ("wgrad head", V, D, T, True, True, 2),
]


def operand(rows, cols, dev):
    """[rows, cols] bf16 view with a 16-byte aligned row stride."""
    return (torch.randn(rows, F.pad8(cols), device=dev, dtype=torch.float32) * 0.5).to(torch.bfloat16)[:, :cols]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "gemm_bench.json"))
    ap.add_argument("--extra", action="store_true", help="also 8192^3 and the reference-small shape")
    a = ap.parse_args()
    dev = torch.device("cuda")
    lib = _lib.load(required=True)
    lib.ppb_set_gemm_cta_group.argtypes = [ctypes.c_int]
    hbm, tf, tag = peaks()
    shapes = list(SHAPES)
    if a.extra:
        shapes += [("square 8192", 8192, 8192, 8192, False, False, 0), ("ref-small qkv", 1024, 2304, 768, False, False, 0)]
    rows = []
    for name, M, N, K, amn, bmn, om in shapes:
        A = operand(K, M, dev) if amn else operand(M, K, dev)
        B = operand(K, N, dev) if bmn else operand(N, K, dev)
        ldc = F.pad8(N)
        C = torch.zeros(M, ldc, device=dev, dtype=torch.float32 if om else torch.bfloat16)[:, :N]
        Am = (A.t() if amn else A)
        Bm = (B.t() if bmn else B)
        ref = Am[:512].float() @ Bm.float().t()                     # fp32 product of the first 512 rows
        row = {"name": name, "M": M, "N": N, "K": K}
        for cg, key in ((1, "cg1"), (2, "cg2"), (0, "auto")):
            _lib.check(lib.ppb_set_gemm_cta_group(cg), "ppb_set_gemm_cta_group")
            try:
                C.zero_()
                F.gemm(A, B, C, M, N, K, amn, bmn, om)
                torch.cuda.synchronize()
                err = ((C[:512].float() - ref).abs().max() / ref.abs().max()).item()
                ms = time_fn(lambda: F.gemm(A, B, C, M, N, K, amn, bmn, om))
                row[key + "_us"] = round(ms * 1e3, 2)
                row[key + "_max_rel_err"] = round(err, 5)
            except Exception as ex:  # noqa: BLE001
                row[key + "_error"] = str(ex)[:200]
        _lib.check(lib.ppb_set_gemm_cta_group(0), "ppb_set_gemm_cta_group")
        Ac, Bc = Am.contiguous(), Bm.contiguous()
        if om == 2:   # the stock counterpart of an accumulating wgrad: matmul + add into the gradient
            out32 = torch.zeros(M, N, device=dev, dtype=torch.float32)
            ms_ref = time_fn(lambda: out32.add_(torch.matmul(Ac, Bc.t())))
            row["cublas_matmul_only_us"] = round(time_fn(lambda: torch.matmul(Ac, Bc.t())) * 1e3, 2)
        else:
            ms_ref = time_fn(lambda: torch.matmul(Ac, Bc.t()))
        row["cublas_us"] = round(ms_ref * 1e3, 2)
        best = min(row.get(k + "_us", 1e9) for k in ("cg1", "cg2"))
        row["auto_vs_best"] = round(row.get("auto_us", 1e9) / best, 3)
        row["auto_vs_cublas"] = round(row.get("auto_us", 1e9) / row["cublas_us"], 3)
        row["auto_tflops"] = round(2.0 * M * N * K / row.get("auto_us", 1e9) / 1e6, 1)
        row["auto_frac_of_" + tag + "_peak"] = round(row["auto_tflops"] / tf, 3)
        rows.append(row)
        print(json.dumps(row), flush=True)
        del A, B, C
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
