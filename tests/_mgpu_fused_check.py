"""torchrun helper: the three fused compute+collective kernels over real NVLink peers (utils/fused_check, small
shapes), plus a sentinel / epoch check of the fused all-gather: the gathered buffer is poisoned with NaN before every
call and the inputs change every call, so a tile that ran ahead of its panel counter (a missed signal) or a counter
left over from the previous call (a stale epoch) shows up as a NaN or as last call's data."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.ops.collective import GatherFlags, SymmetricBuffer, allgather_linear_fused  # noqa: E402
from ppb200.utils import fused_check  # noqa: E402

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
res = fused_check.run(dist.group.WORLD, dev, small=True)

rows, d, n = 512, 1024, 1536
bufs = [SymmetricBuffer((rows, d), torch.bfloat16, dev, dist.group.WORLD) for _ in range(2)]
gfl = GatherFlags(rows * world, dev)
w = (torch.randn(n, d, generator=torch.Generator().manual_seed(5)) * 0.05).to(dev).to(torch.bfloat16)
worst, exact = 0.0, True
for it in range(6):
    g = torch.Generator().manual_seed(100 + it)
    x = torch.randn(rows * world, d, generator=g).to(dev).to(torch.bfloat16)     # same on every rank
    buf = bufs[it & 1]
    buf.tensor.copy_(x[rank * rows:(rank + 1) * rows])
    buf.barrier(0)
    full = torch.full((rows * world, d), float("nan"), dtype=torch.bfloat16, device=dev)
    y = allgather_linear_fused(buf.ptrs, rows, d, full, gfl, w, rank)
    torch.cuda.synchronize()
    ref = x.float() @ w.float().t()
    worst = max(worst, ((y.float() - ref).norm() / ref.norm()).item())
    exact = exact and bool(torch.equal(full, x)) and bool(torch.isfinite(y.float()).all())
t = torch.tensor([worst, 0.0 if exact else 1.0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    res["sentinel"] = {"max_rel_err": float(t[0].item()), "gathered_exact_and_finite_on_all_ranks": bool(t[1].item() == 0.0), "calls": 6}
    print(json.dumps(res))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
