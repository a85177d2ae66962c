"""Parity + roofline check of the three fused compute+collective kernels over real NVLink peers.

``run(group, device)`` is collective over a process group of >= 2 ranks on one NVSwitch box (``bench.py`` calls it on
every multi-GPU run after the headline measurement, so the driver's own runs carry the evidence; the multi-GPU tests
call it too).  For each path it reports

    ms            device time of the fused kernel path (CUDA events, max over the ranks of the group)
    nccl_ms       the same math with stock pieces: cuBLAS GEMM (torch.matmul) + the NCCL collective
    frac_roofline roofline_ms / ms, roofline = max(FLOPs / measured bf16 peak, bytes each GPU must send / 770 GB/s)
    max_rel_err   against an fp32 torch product combined with NCCL collectives

K10 = last-micro-batch weight-gradient GEMMs + data-parallel all-reduce (``ppb_gemm_allreduce_bf16``),
K11 = GEMM -> reduce-scatter (``ppb_gemm_reducescatter_bf16``), K12 = all-gather -> GEMM in one kernel
(``ppb_gemm_allgather_fused_bf16``).  Shapes: one Llama-7B layer's tensor-parallel linears at 4096 tokens for K11/K12,
one GPT-2-medium layer's weight matrices at 4096 tokens for K10 (BASELINE.json configs 4 and 2).
The reference has no collective of its own; the baselines are the stock torch / NCCL compositions.
"""
from __future__ import annotations

import json
import os
from typing import Dict

import torch
import torch.distributed as dist

NVLINK_GBS = 770.0   # measured peer-copy bandwidth per direction per GPU on this pool (B200_PROFILING.md)


def _peak_tflops() -> float:
    root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    p = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["bf16_tflops"])
    return 1590.0


def _timed(fn, group, dev, iters: int = 10, warm: int = 3) -> float:
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier(group=group, device_ids=[dev.index])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def _max_over(group, dev, v: float) -> float:
    t = torch.tensor([v], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def run(group, device, tokens: int = 4096, small: bool = False, nvls: bool = True) -> Dict[str, dict]:
    from ..ops import functional as F
    from ..ops.collective import (GatherFlags, SymmetricBuffer, allgather_linear_fused, gemm_allreduce,
                                  linear_reducescatter)
    dev = torch.device(device)
    tp = dist.get_world_size(group)
    rank = dist.get_rank(group)
    peak = _peak_tflops() * 1e12
    link = NVLINK_GBS * 1e9
    out: Dict[str, dict] = {}
    verbose = os.environ.get("PPB_FUSED_CHECK_VERBOSE", "0") == "1"

    def note(msg: str) -> None:
        if verbose and rank == 0:
            print(f"[fused_check tp={tp}] {msg}", flush=True)

    g = torch.Generator(device="cpu").manual_seed(1234)
    T = tokens if not small else 512
    rows = T // tp

    # ---------------- K12: all-gather -> GEMM (column-parallel linear1 of a Llama-7B layer, gate|up packed) ------------
    d, f = (4096, 11008) if not small else (512, 1024)
    n_loc = (2 * f // tp) // 8 * 8
    x = (torch.randn(T, d, generator=g) * 0.5).to(dev).to(torch.bfloat16)       # same on every rank (same seed)
    w1 = (torch.randn(n_loc, d, generator=g) * 0.02).to(dev).to(torch.bfloat16)
    xs = x[rank * rows:(rank + 1) * rows].contiguous()
    xsyms = [SymmetricBuffer((rows, d), torch.bfloat16, dev, group) for _ in range(2)]
    full = torch.empty(T, d, dtype=torch.bfloat16, device=dev)
    gfl = GatherFlags(T, dev)
    y = torch.empty(T, n_loc, dtype=torch.bfloat16, device=dev)
    kx = [0]

    def k12():   # as TPComm.ag_linear: double-buffered shards, so ONE device barrier per call orders everything
        xsym = xsyms[kx[0]]
        kx[0] ^= 1
        xsym.tensor.copy_(xs)
        xsym.barrier(0)
        allgather_linear_fused(xsym.ptrs, rows, d, full, gfl, w1, rank, out=y)

    xg = torch.empty(T, d, dtype=torch.bfloat16, device=dev)

    def k12_base():
        dist.all_gather_into_tensor(xg, xs, group=group)
        return torch.matmul(xg, w1.t())

    note("k12 first call")
    k12()
    torch.cuda.synchronize(dev)
    note("k12 done once")
    ref = x[:256].float() @ w1.float().t()
    err = ((y[:256].float() - ref).norm() / ref.norm()).item()
    gathered_ok = bool(torch.equal(full, x))
    ms, base = _timed(k12, group, dev), _timed(k12_base, group, dev)
    roof = max(2.0 * T * n_loc * d / peak, (tp - 1) * rows * d * 2 / link) * 1e3
    out["k12_allgather_gemm"] = {"ms": round(ms, 4), "nccl_ms": round(base, 4), "frac_roofline": round(roof / ms, 3),
                                 "max_rel_err": round(_max_over(group, dev, err), 5), "gathered_copy_exact": gathered_ok,
                                 "shape": [T, n_loc, d], "tp": tp}

    # ---------------- K11: GEMM -> reduce-scatter (row-parallel linear2 of the same layer) ------------------------------
    k_loc = (f // tp) // 8 * 8
    h = (torch.randn(T, k_loc, generator=torch.Generator(device="cpu").manual_seed(77 + rank)) * 0.5).to(dev).to(torch.bfloat16)
    w2 = (torch.randn(d, k_loc, generator=torch.Generator(device="cpu").manual_seed(99 + rank)) * 0.02).to(dev).to(torch.bfloat16)
    ysym = SymmetricBuffer((rows, d), torch.float32, dev, group)

    def k11():
        ysym.tensor.zero_()
        ysym.barrier(0)
        linear_reducescatter(h, w2, ysym.ptrs, d)
        ysym.barrier(1)

    yo = torch.empty(rows, d, dtype=torch.bfloat16, device=dev)

    def k11_base():
        dist.reduce_scatter_tensor(yo, torch.matmul(h, w2.t()), group=group)

    note("k11 first call")
    k11()
    torch.cuda.synchronize(dev)
    note("k11 done once")
    part = (h.float() @ w2.float().t())
    dist.all_reduce(part, group=group)
    ref = part[rank * rows:(rank + 1) * rows]
    err = ((ysym.tensor - ref).norm() / ref.norm()).item()
    ms, base = _timed(k11, group, dev), _timed(k11_base, group, dev)
    roof = max(2.0 * T * d * k_loc / peak, (tp - 1) * rows * d * 4 / link) * 1e3
    out["k11_gemm_reducescatter"] = {"ms": round(ms, 4), "nccl_ms": round(base, 4), "frac_roofline": round(roof / ms, 3),
                                     "max_rel_err": round(_max_over(group, dev, err), 5), "shape": [T, d, k_loc], "tp": tp,
                                     "wire": "fp32 partial sums (the NCCL baseline moves bf16)"}
    del part, h, w2, x, w1, y, xg, full

    # ---------------- K10: weight-gradient GEMMs + data-parallel all-reduce (one GPT-2-medium layer) ---------------------
    dm, fm = (1024, 2048) if not small else (256, 512)
    shapes = [(3 * dm, dm), (dm, dm), (dm, dm), (2 * dm, dm), (dm, dm), (fm, dm), (dm, fm)]   # (out, in) of a helper layer
    total = sum(o * i for o, i in shapes)
    accum = torch.zeros(total, device=dev)
    fin = SymmetricBuffer((total,), torch.float32, dev, group)
    gg = torch.Generator(device="cpu").manual_seed(500 + rank)
    dys = {o: (torch.randn(T, o, generator=gg) * 0.1).to(dev).to(torch.bfloat16) for o in sorted({s[0] for s in shapes})}
    xin = {i: (torch.randn(T, i, generator=gg) * 0.5).to(dev).to(torch.bfloat16) for i in sorted({s[1] for s in shapes})}

    def k10():
        off = 0
        for o, i in shapes:
            gemm_allreduce(dys[o], xin[i], accum[off:off + o * i].view(o, i), [p + 4 * off for p in fin.ptrs])
            off += o * i
        fin.barrier(0)

    def k10_base():
        off = 0
        for o, i in shapes:
            accum[off:off + o * i].view(o, i).add_(torch.matmul(dys[o].t(), xin[i]))
            off += o * i
        dist.all_reduce(accum, group=group)

    accum.zero_()
    fin.tensor.zero_()
    torch.cuda.synchronize(dev)
    dist.barrier(group=group, device_ids=[dev.index])
    note("k10 first call")
    k10()
    torch.cuda.synchronize(dev)
    note("k10 done once")
    o0, i0 = shapes[0]
    ref = dys[o0].float().t() @ xin[i0].float()
    dist.all_reduce(ref, group=group)
    err = ((fin.tensor[:o0 * i0].view(o0, i0) - ref).norm() / ref.norm()).item()
    ms, base = _timed(k10, group, dev), _timed(k10_base, group, dev)
    roof = max(2.0 * T * total / peak, (tp - 1) * total * 4 / link) * 1e3
    out["k10_wgrad_allreduce"] = {"ms": round(ms, 4), "nccl_ms": round(base, 4), "frac_roofline": round(roof / ms, 3),
                                  "max_rel_err": round(_max_over(group, dev, err), 5), "params": total, "tokens": T, "dp": tp,
                                  "baseline": "cuBLAS matmul + add into the fp32 gradient, then ncclAllReduce of the bucket"}
    # NVLS form of the same reduction: plain accumulating weight-gradient GEMMs, then ONE multimem.red stream over the bucket
    # (every GPU sends each gradient byte once; the NVSwitch adds it into every replica) -- the dp >= 4 path of the non-GEMM
    # gradients, timed here over the whole bucket
    mc = int(getattr(fin.hdl, "multicast_ptr", 0) or 0)
    if not nvls:
        pass
    elif mc:
        from ..ops.collective import push_add_multimem

        def k10_nvls():
            off = 0
            for o, i in shapes:
                F.linear_wgrad(dys[o], xin[i], accum[off:off + o * i].view(o, i))
                off += o * i
            push_add_multimem(accum, mc, 1.0)
            fin.barrier(0)

        accum.zero_()
        fin.tensor.zero_()
        torch.cuda.synchronize(dev)
        dist.barrier(group=group, device_ids=[dev.index])
        note("k10 nvls first call")
        k10_nvls()
        torch.cuda.synchronize(dev)
        note("k10 nvls done once")
        err = ((fin.tensor[:o0 * i0].view(o0, i0) - ref).norm() / ref.norm()).item()
        ms = _timed(k10_nvls, group, dev)
        roof_nvls = max(2.0 * T * total / peak, total * 4 / link) * 1e3
        out["k10_nvls_multimem_push"] = {"ms": round(ms, 4), "nccl_ms": round(base, 4), "frac_roofline": round(roof_nvls / ms, 3),
                                         "max_rel_err": round(_max_over(group, dev, err), 5), "dp": tp,
                                         "note": "accumulating wgrad GEMMs + one multimem.red pass over the bucket (in-switch reduction)"}
    else:
        out["k10_nvls_multimem_push"] = {"unavailable": "the symmetric-memory handle exposes no multicast address on this box"}
    out["denominators"] = {"bf16_tflops_measured": round(peak / 1e12, 1), "nvlink_gbs_per_direction": NVLINK_GBS}
    return out
