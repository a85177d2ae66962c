"""torchrun worker of tests/test_reference_arm.py: the reference arm's data-parallel composition (bench.py:
reference_stage_with_replicas) on CPU / gloo, 2 stages x 2 replicas of the UNMODIFIED helper model."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import LLMsDistributedTrainingHelper as ref  # noqa: E402
from torch.distributed.pipelining import Schedule1F1B  # noqa: E402

import bench  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
dp = 2
pp_world = world // dp
pp_rank, dp_idx = rank % pp_world, rank // pp_world
pp_group = dp_group = None
for d in range(dp):
    g_ = dist.new_group([d * pp_world + p_ for p_ in range(pp_world)])
    if d == dp_idx:
        pp_group = g_
for p_ in range(pp_world):
    g_ = dist.new_group([d * pp_world + p_ for d in range(dp)])
    if p_ == pp_rank:
        dp_group = g_
margs = ref.ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=128)
dev = torch.device("cpu")
torch.manual_seed(1234)
model = ref.Transformer(margs)
stage = bench.reference_stage_with_replicas(ref, model, pp_rank, pp_world, dev, pp_group, dp_group)


def loss_fn(o, t):
    return torch.nn.functional.cross_entropy(o.reshape(-1, margs.vocab_size), t.reshape(-1))


sched = Schedule1F1B(stage, n_microbatches=4, loss_fn=loss_fn)
g = torch.Generator().manual_seed(1000 + dp_idx)
x, y = torch.randint(0, 128, (8, 16), generator=g), torch.randint(0, 128, (8, 16), generator=g)
opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
hist = []
for it in range(3):
    losses = []
    if pp_rank == 0:
        sched.step(x)
    else:
        sched.step(target=y, losses=losses)
    opt.step()
    opt.zero_grad(set_to_none=False)
    if losses:
        hist.append(float(torch.stack(losses).detach().mean()))
p0 = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
ps = [torch.empty_like(p0) for _ in range(dp)]
dist.all_gather(ps, p0, group=dp_group)
same = torch.tensor([1.0 if torch.equal(ps[0], ps[1]) else 0.0])
dist.all_reduce(same, op=dist.ReduceOp.MIN)
out = [None] * world
dist.all_gather_object(out, hist)
if rank == 0:
    print(json.dumps({"replicas_equal": bool(same.item()), "losses": [h for h in out if h]}), flush=True)
dist.barrier()
dist.destroy_process_group()
