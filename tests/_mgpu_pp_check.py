"""torchrun helper: pp=2 over NCCL -- first-step loss equals the single-GPU loss (same weights, dropout 0) and
training reduces it."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.api import make_schedule, manual_model_split  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
torch.manual_seed(0)
args = ModelArgs(dim=256, n_layers=4, n_heads=4, vocab_size=1000, ffn_dim=512, dropout=0.0)
full = Transformer(args)
g = torch.Generator().manual_seed(7)
x = torch.randint(0, 1000, (8, 128), generator=g).to(dev)
y = torch.randint(0, 1000, (8, 128), generator=g).to(dev)

loss1 = None
if rank == 0:  # single-GPU oracle on rank 0's device (no process group involved: pp_size would be 2)
    st = manual_model_split(copy.deepcopy(full), 0, 1, dev)
    st.materialize()
    chunks_x, chunks_y = x.chunk(4), y.chunk(4)
    tot = 0.0
    for i in range(4):
        st.forward_one_chunk(i, chunks_x[i], chunks_y[i])
        tot += float(st.losses[i].item())
    loss1 = tot / 4

stage = manual_model_split(copy.deepcopy(full), rank, 2, dev)
sched = make_schedule("1F1B", stage, 4, None, optimizer={"lr": 2e-3, "weight_decay": 0.0})
hist = []
for _ in range(6):
    losses = []
    sched.step(*((x,) if rank == 0 else ()), target=y if rank == 1 else None, losses=losses if rank == 1 else None)
    if rank == 1:
        hist.append(float(torch.stack([l.reshape(()) for l in losses]).mean().item()))
t = torch.tensor(hist if rank == 1 else [0.0] * 6, device=dev)
dist.broadcast(t, src=1)
if rank == 0:
    print(json.dumps({"loss_pp1": loss1, "loss_pp2": float(t[0]), "loss_after": float(t[-1]), "bubble": sched.measured()}))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
