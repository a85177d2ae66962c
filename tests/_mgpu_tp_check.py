"""torchrun helper: tensor+sequence parallel MLP (K12 all-gather->GEMM, K11 GEMM->reduce-scatter) over real NVLink
peers vs the unsharded computation."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.ops.collective import TensorParallelMLP  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
T, d, f = 2048, 1024, 4096
g = torch.Generator(device="cpu").manual_seed(3)
x = torch.randn(T, d, generator=g).to(dev).to(torch.bfloat16)
w1 = (torch.randn(f, d, generator=g) * 0.03).to(dev).to(torch.bfloat16)
b1 = (torch.randn(f, generator=g) * 0.1).to(dev).to(torch.bfloat16)
w2 = (torch.randn(d, f, generator=g) * 0.03).to(dev).to(torch.bfloat16)
fs = f // world
mlp = TensorParallelMLP(w1[rank * fs:(rank + 1) * fs].contiguous(), b1[rank * fs:(rank + 1) * fs].contiguous(),
                        w2[:, rank * fs:(rank + 1) * fs].contiguous(), T, dist.group.WORLD)
rows = T // world
errs = []
for it in range(3):
    y = mlp.forward(x[rank * rows:(rank + 1) * rows])
    torch.cuda.synchronize()
    h = torch.relu(x.float() @ w1.float().t() + b1.float()).to(torch.bfloat16)
    ref = (h.float() @ w2.float().t())[rank * rows:(rank + 1) * rows]
    errs.append(((y - ref).norm() / ref.norm()).item())


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier(device_ids=[local])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


xs = x[rank * rows:(rank + 1) * rows].contiguous()
t_fused = timed(lambda: mlp.forward(xs))
xg = torch.empty(T, d, device=dev, dtype=torch.bfloat16)
w1s, b1s, w2s = mlp.w1, mlp.b1, mlp.w2
ys = torch.empty(rows, d, device=dev, dtype=torch.bfloat16)


def baseline():  # NCCL all-gather + cuBLAS + NCCL reduce-scatter
    dist.all_gather_into_tensor(xg, xs)
    hh = torch.relu(torch.nn.functional.linear(xg, w1s, b1s))
    yp = torch.nn.functional.linear(hh, w2s)
    dist.reduce_scatter_tensor(ys, yp)


t_base = timed(baseline)
e = torch.tensor([max(errs)], device=dev)
dist.all_reduce(e, op=dist.ReduceOp.MAX)
if rank == 0:
    flops = 2.0 * T * d * fs * 2
    link_bytes = (world - 1) * rows * d * 2 + (world - 1) * rows * d * 4
    peak = 1590.0
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["bf16_tflops"]
    roof = max(flops / (peak * 1e12), link_bytes / 770e9) * 1e3
    print(json.dumps({"kernel": "tp_mlp(allgather->gemm, gemm->reduce_scatter)", "tp": world, "T": T, "d": d, "f": f,
                      "max_rel_err": float(e.item()), "fused_ms": round(t_fused, 4), "nccl_cublas_ms": round(t_base, 4),
                      "roofline_ms": round(roof, 4), "fused_fraction_of_roofline": round(roof / t_fused, 3)}))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
