#!/usr/bin/env python
"""K10 micro-benchmark (torchrun, >= 2 GPUs): the weight-gradient GEMMs of one GPT-2-medium-shaped stage with the
data-parallel all-reduce (a) as cuBLAS-free baseline = our wgrad GEMMs + ncclAllReduce of the gradient arena and
(b) fused into the GEMM epilogues (TMA reduce-add into every replica over NVLink).  Reports device time (max over
ranks) and the achieved fraction of the roofline max(FLOPs / measured bf16 peak, NVLink bytes / 770 GB/s)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ppb200.ops import functional as F  # noqa: E402
from ppb200.ops.collective import gemm_allreduce  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
import torch.distributed._symmetric_memory as symm_mem  # noqa: E402

T, d, f, layers = 4096, 1024, 2048, int(os.environ.get("LAYERS", 6))
shapes = [(3 * d, d), (d, d), (d, d), (2 * d, d), (d, d), (f, d), (d, f)] * layers   # (out, in) per weight matrix
total = sum(o * i for o, i in shapes)
accum = torch.zeros(total, device=dev)
final = symm_mem.empty(total, dtype=torch.float32, device=dev)
final.zero_()
hdl = symm_mem.rendezvous(final, dist.group.WORLD)
ptrs = [int(p) for p in hdl.buffer_ptrs]
dys = {o: torch.randn(T, o, device=dev, dtype=torch.bfloat16) for o in {s[0] for s in shapes}}
xs = {i: torch.randn(T, i, device=dev, dtype=torch.bfloat16) for i in {s[1] for s in shapes}}
alpha = 1.0 / (32 * world)


def baseline():
    off = 0
    for o, i in shapes:
        F.linear_wgrad(dys[o], xs[i], accum[off:off + o * i].view(o, i))
        off += o * i
    dist.all_reduce(accum)


def fused():
    off = 0
    for o, i in shapes:
        gemm_allreduce(dys[o], xs[i], accum[off:off + o * i].view(o, i), [p + 4 * off for p in ptrs])
        off += o * i
    hdl.barrier(channel=0)


def timed(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier(device_ids=[local])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    dist.barrier(device_ids=[local])
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# correctness of the fused result against the baseline (both start from a zero accumulator)
accum.zero_()
baseline()
ref = accum.clone()
accum.zero_()
final.zero_()
torch.cuda.synchronize()
dist.barrier(device_ids=[local])
fused()
torch.cuda.synchronize()
err = ((final - ref).norm() / ref.norm()).item()

t_base = timed(baseline)
t_fused = timed(fused)
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0}
flops = 2.0 * T * total
nvlink_bytes = 4.0 * total * (world - 1)          # what every GPU must send (fp32 gradients to each peer)
t_compute = flops / (peaks["bf16_tflops"] * 1e12) * 1e3
t_link = nvlink_bytes / 770e9 * 1e3
roof = max(t_compute, t_link)
if rank == 0:
    print(json.dumps({"kernel": "wgrad+dp_allreduce", "n_gpus": world, "params": total, "rel_err_vs_baseline": err,
                      "baseline_gemm_plus_nccl_ms": round(t_base, 4), "fused_ms": round(t_fused, 4),
                      "speedup": round(t_base / t_fused, 3), "roofline_ms": round(roof, 4),
                      "roofline_bound": "nvlink" if t_link > t_compute else "tensor",
                      "fused_fraction_of_roofline": round(roof / t_fused, 3),
                      "baseline_fraction_of_roofline": round(roof / t_base, 3)}))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
