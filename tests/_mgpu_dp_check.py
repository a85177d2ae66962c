"""torchrun helper: dp=2, pp=1 -- parameters after 2 steps with the fused in-kernel all-reduce (K10) must match
the NCCL all-reduce path."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.api import make_schedule, manual_model_split  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402
from ppb200.parallel.mesh import init_mesh  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
mesh = init_mesh(dp_size=2, pp_size=1)
torch.manual_seed(0)
args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=1000, ffn_dim=512, dropout=0.0)
full = Transformer(args)
g = torch.Generator().manual_seed(100 + rank)
x = torch.randint(0, 1000, (8, 128), generator=g).to(dev)
y = torch.randint(0, 1000, (8, 128), generator=g).to(dev)


def run(fused: bool):
    os.environ["PPB_FUSED_DP"] = "1" if fused else "0"
    st = manual_model_split(copy.deepcopy(full), 0, 1, dev)
    st.dp_group = mesh.dp_group
    sched = make_schedule("1F1B", st, 4, None, pp_group=mesh.pp_group, dp_group=mesh.dp_group,
                          optimizer={"lr": 1e-3, "weight_decay": 0.0})
    last = None
    for _ in range(2):
        losses = []
        sched.step(x, target=y, losses=losses)
        last = torch.stack([l.reshape(()) for l in losses]).mean()
    torch.cuda.synchronize()
    return sched, st, float(last.item())


s_f, st_f, loss_f = run(True)
s_n, st_n, loss_n = run(False)
a, b = st_f.master[:st_f.n_total], st_n.master[:st_n.n_total]
diff = ((a - b).norm() / (b.norm() + 1e-12)).item()
# replicas must hold identical parameters after the fused path
other = a.clone()
dist.broadcast(other, src=0)
rep = ((a - other).norm() / (a.norm() + 1e-12)).item()
if rank == 0:
    print(json.dumps({"fused_enabled": bool(s_f.fused_dp), "nccl_path_fused": bool(s_n.fused_dp), "max_rel_param_diff": max(diff, rep),
                      "loss_diff": abs(loss_f - loss_n)}))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
