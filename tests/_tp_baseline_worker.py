"""torchrun worker of tests/test_tp_baseline.py: a few training steps of the stock-torch tensor-parallel baseline on
CPU / gloo; prints the global mean loss per step as one JSON line (rank 0)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from benchmarks import tp_baseline  # noqa: E402
from ppb200.models.args import ModelArgs  # noqa: E402

tp, block, sched = int(sys.argv[1]), sys.argv[2], sys.argv[3]
dist.init_process_group("gloo")
torch.manual_seed(0)
kw = dict(dim=64, n_layers=4, n_heads=4, vocab_size=128, ffn_dim=128, causal=True, cross_attn=False, norm_first=True, dropout=0.0)
if block == "llama":
    kw.update(activation="swiglu", norm="rmsnorm", rope=True)
else:
    kw.update(activation="gelu")
losses = tp_baseline.train_steps(ModelArgs(**kw), sched, 4, tp, batch=8, seq_len=16, steps=3, device=torch.device("cpu"))
if dist.get_rank() == 0:
    print(json.dumps({"losses": losses}), flush=True)
dist.barrier()
dist.destroy_process_group()
