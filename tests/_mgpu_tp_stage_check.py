"""torchrun helper (2 GPUs): a tensor-parallel (tp=2, sequence-parallel activations, fused all-gather->GEMM and
GEMM->reduce-scatter kernels) pre-LN stage trains like the same model on one GPU: first-step loss, merged gradients,
loss after a few optimizer steps."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ppb200.api import make_schedule, manual_model_split  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402
from ppb200.models.tp import shard_model_for_tp, shard_tensor  # noqa: E402
from ppb200.parallel.mesh import init_mesh  # noqa: E402
from ppb200.parallel.stage import PipelineStage  # noqa: E402

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl")
mesh = init_mesh(dp_size=1, pp_size=1, tp_size=2)
torch.manual_seed(0)
args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=1000, ffn_dim=512, dropout=0.0, causal=True, cross_attn=False,
                 activation=os.environ.get("ACT", "gelu"), norm_first=True, norm="rmsnorm",
                 rope=os.environ.get("ACT", "gelu") == "swiglu")
full = Transformer(args)
g = torch.Generator().manual_seed(7)
x = torch.randint(0, 1000, (8, 128), generator=g).to(dev)
y = torch.randint(0, 1000, (8, 128), generator=g).to(dev)


def run(stage, steps, opt):
    sched = make_schedule("1F1B", stage, 4, None, pp_group=mesh.pp_group, optimizer=opt)
    hist = []
    for _ in range(steps):
        losses = []
        sched.step(x, target=y, losses=losses)
        hist.append(float(torch.stack([l.reshape(()) for l in losses]).mean().item()))
    return hist

# single-GPU oracle (no tensor parallelism), every rank computes it on its own device
ref = manual_model_split(copy.deepcopy(full), 0, 1, dev)
ref_hist = run(ref, 1, None)
ref_grads = ref.grad_views()
mdl = copy.deepcopy(full)
shard_model_for_tp(mdl, mesh.tp_idx, 2)
st = PipelineStage(mdl, 0, 1, dev, tp_group=mesh.tp_group)
hist = run(st, 1, None)
worst, worst_name = 0.0, ""
for n, gv in st.grad_views().items():
    want = shard_tensor(n, ref_grads[n], args.n_heads, mesh.tp_idx, 2, args.activation == "swiglu")
    err = ((gv - want).norm() / (want.norm() + 1e-12)).item()
    if err > worst:
        worst, worst_name = err, n
# a few optimizer steps on both
ref2 = manual_model_split(copy.deepcopy(full), 0, 1, dev)
h_ref = run(ref2, 5, {"lr": 2e-3, "weight_decay": 0.0})
mdl2 = copy.deepcopy(full)
shard_model_for_tp(mdl2, mesh.tp_idx, 2)
st2 = PipelineStage(mdl2, 0, 1, dev, tp_group=mesh.tp_group)
h_tp = run(st2, 5, {"lr": 2e-3, "weight_decay": 0.0})
t = torch.tensor([worst], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"loss_ref": ref_hist[0], "loss_tp": hist[0], "grad_rel_err_max": float(t.item()), "worst": worst_name,
                      "train_ref": h_ref, "train_tp": h_tp}))
dist.barrier(device_ids=[local])
dist.destroy_process_group()
