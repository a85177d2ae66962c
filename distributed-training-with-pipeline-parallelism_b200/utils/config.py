"""One config object + CLI for everything the reference hard-codes or hides in positional arguments
(SURVEY.md section 5.6): micro-batch count (helper:214), model width/vocab (helper:188), backend/device
(helper:175-177), port (helper:169), warm-up count (helper:113), virtual chunks (helper:183)."""
from __future__ import annotations

import argparse
from dataclasses import asdict, dataclass, field, fields
from typing import Optional


@dataclass
class RunConfig:
    model: str = "ref-small"              # preset name in ppb200.models.PRESETS
    n_layers: Optional[int] = None        # overrides the preset
    n_heads: Optional[int] = None
    dim: Optional[int] = None
    vocab_size: Optional[int] = None
    dropout: Optional[float] = None
    schedule: str = "1F1B"                # 'GPipe' | '1F1B' | 'Interleaved1F1B'
    pp: int = 1
    dp: int = 1
    tp: int = 1
    virtual_stages: int = 1
    n_microbatches: int = 4
    global_batch: int = 32
    seq_len: int = 128
    steps: int = 10
    warmup: int = 2
    backend: Optional[str] = None         # default: nccl on CUDA, gloo on CPU
    device: Optional[str] = None          # 'cuda' | 'cpu'
    dtype: str = "bf16"                   # 'bf16' (CUDA) | 'fp32'
    partition: str = "reference"          # 'reference' (helper:70-75) | 'balanced' | 'fine' (sub-layer cuts) | 'auto' (measured unit times)
    lr: float = 1e-4
    weight_decay: float = 0.1
    beta1: float = 0.9
    beta2: float = 0.95
    max_grad_norm: Optional[float] = None
    optimizer: bool = True                # the reference has none; BASELINE.json requires one in the timed step
    seed: int = 0
    master_addr: str = "127.0.0.1"
    master_port: Optional[int] = None
    collective_timeout_s: Optional[float] = None   # a collective / P2P that waits longer aborts the rank (dead-peer detection; torch default 10-30 min)
    trace_path: Optional[str] = None      # chrome trace of the measured timeline
    metrics_path: Optional[str] = None    # JSONL
    recompute: bool = False               # activation recomputation (memory for one extra forward per micro-batch)
    # ---- training loop (ppb200.train) ----
    log_every: int = 1
    save_dir: Optional[str] = None        # per-(stage, tp rank) shards in the reference state_dict layout
    save_every: int = 0                   # 0 = only at the end (when save_dir is set)
    resume_from: Optional[str] = None     # directory written by save_dir; may have a different (pp, tp) layout
    data_path: Optional[str] = None       # flat token file (.bin uint16/int32 via --data-dtype, or .npy); default synthetic
    data_dtype: str = "uint16"
    lr_warmup_steps: int = 0
    lr_decay_steps: int = 0               # 0 = constant after warm-up; else cosine down to lr * lr_min_ratio
    lr_min_ratio: float = 0.1

    def optimizer_dict(self) -> Optional[dict]:
        if not self.optimizer:
            return None
        d = {"lr": self.lr, "betas": (self.beta1, self.beta2), "weight_decay": self.weight_decay}
        if self.max_grad_norm:
            d["max_grad_norm"] = self.max_grad_norm
        return d

    def lr_at(self, step: int) -> float:
        """Linear warm-up to ``lr``, then cosine decay to ``lr * lr_min_ratio`` over ``lr_decay_steps`` (0 = constant)."""
        import math
        if self.lr_warmup_steps and step < self.lr_warmup_steps:
            return self.lr * (step + 1) / self.lr_warmup_steps
        if not self.lr_decay_steps:
            return self.lr
        t = min(1.0, (step - self.lr_warmup_steps) / max(1, self.lr_decay_steps - self.lr_warmup_steps))
        return self.lr * (self.lr_min_ratio + (1.0 - self.lr_min_ratio) * 0.5 * (1.0 + math.cos(math.pi * t)))

    def model_args(self):
        from ..models import preset
        over = {k: getattr(self, k) for k in ("n_layers", "n_heads", "dim", "vocab_size", "dropout") if getattr(self, k) is not None}
        return preset(self.model, **over)

    def to_dict(self) -> dict:
        return asdict(self)

    @classmethod
    def from_args(cls, argv=None) -> "RunConfig":
        ap = argparse.ArgumentParser(description="ppb200 run configuration")
        for f in fields(cls):
            name = "--" + f.name.replace("_", "-")
            if f.type in ("bool", bool):
                ap.add_argument(name, type=lambda s: s.lower() in ("1", "true", "yes"), default=f.default)
            else:
                base = {"int": int, "float": float, "str": str, "Optional[int]": int, "Optional[float]": float,
                        "Optional[str]": str}.get(str(f.type), str)
                ap.add_argument(name, type=base, default=f.default)
        ns = ap.parse_args(argv)
        return cls(**vars(ns))
