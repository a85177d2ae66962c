"""torchrun worker of tests/test_reference_arm.py::test_pipeline_losses_match_the_unmodified_helper.

SURVEY.md section 4, tier "Oracle vs. reference": the UNMODIFIED helper (its Transformer, its manual_model_split, torch's
pipeline runtime underneath, gloo) and this framework (own schedule IR + runner + explicit stage program) are given the
same weights (the state_dict format is shared) and the same batch; the per-micro-batch losses of the last stage and the
parameters after one SGD-free step of gradient accumulation must agree.  Dropout is switched off on the helper's module
instances (its code is not touched): the two sides cannot share random streams."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
import LLMsDistributedTrainingHelper as ref  # noqa: E402
from torch.distributed.pipelining import Schedule1F1B, ScheduleGPipe  # noqa: E402

from ppb200.api import make_schedule, manual_model_split  # noqa: E402
from ppb200.models import ModelArgs, Transformer  # noqa: E402

sched_name = sys.argv[1]
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
torch.set_num_threads(1)
dev = torch.device("cpu")
V, m = 128, 4
torch.manual_seed(1234)
ref_model = ref.Transformer(ref.ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=V))
for mod in ref_model.modules():                      # dropout off on the INSTANCES (train mode stays, like the helper)
    if isinstance(mod, torch.nn.Dropout):
        mod.p = 0.0
    if isinstance(mod, torch.nn.MultiheadAttention):
        mod.dropout = 0.0
full_sd = copy.deepcopy(ref_model.state_dict())
g = torch.Generator().manual_seed(99)
x, y = torch.randint(0, V, (8, 16), generator=g), torch.randint(0, V, (8, 16), generator=g)


def loss_fn(out, tgt):                               # helper:197-201
    return torch.nn.CrossEntropyLoss()(out.reshape(-1, V), tgt.reshape(-1))


# ---- the helper's pipeline (its worker builds one pruned copy per virtual stage, helper:186-194) ----
from torch.distributed.pipelining import ScheduleInterleaved1F1B  # noqa: E402
v = 2 if sched_name == "Interleaved1F1B" else 1
ref_models = [ref_model] + [copy.deepcopy(ref_model) for _ in range(v - 1)]
ref_stages = [ref.manual_model_split(mdl, rank + world * i, world * v, dev) for i, mdl in enumerate(ref_models)]
if v == 1:
    rs = {"GPipe": ScheduleGPipe, "1F1B": Schedule1F1B}[sched_name](ref_stages[0], n_microbatches=m, loss_fn=loss_fn)
else:
    rs = ScheduleInterleaved1F1B(ref_stages, n_microbatches=m, loss_fn=loss_fn)
ref_losses = []
if rank == 0:
    rs.step(x)
else:
    rs.step(target=y, losses=ref_losses)
ref_grads = {n: p.grad.clone() for mdl in ref_models for n, p in mdl.named_parameters() if p.grad is not None}

# ---- this framework, same weights ----
sts = []
for i in range(v):
    ours = Transformer(ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=V, dropout=0.0))
    ours.load_state_dict(full_sd)
    sts.append(manual_model_split(ours, rank + world * i, world * v, dev))
sched = make_schedule(sched_name, sts if v > 1 else sts[0], m, None, optimizer=None)
our_losses = []
sched.step(*((x,) if rank == 0 else ()), target=y if rank == world - 1 else None, losses=our_losses if rank == world - 1 else None)
views = {}
for st in sts:
    views.update(st.grad_views())
grad_err = 0.0
for n, gr in ref_grads.items():
    # torch's runtime scales the accumulated gradients by 1 / n_microbatches, like this runner does without an optimizer
    grad_err = max(grad_err, ((views[n] - gr).abs().max() / (gr.abs().max() + 1e-8)).item())
out = {"rank": rank, "grad_rel_err": grad_err, "n_grads": len(ref_grads),
       "ref_losses": [float(l) for l in ref_losses], "our_losses": [float(l.reshape(())) for l in our_losses]}
box = [None] * world
dist.all_gather_object(box, out)
if rank == 0:
    print(json.dumps(box), flush=True)
dist.barrier()
dist.destroy_process_group()
