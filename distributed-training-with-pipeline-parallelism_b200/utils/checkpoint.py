"""Checkpoint / resume (the reference has none).  Format = the reference ``state_dict``
layout: each pipeline stage writes ``stage_{i:03d}.pt`` holding a strict subset of the full model's keys (global
layer indices), so merging is a dict union and re-sharding to a different stage count is a dict filter."""
from __future__ import annotations

import glob
import os
from typing import Dict, Iterable, Optional

import torch


def save_stage(stage, directory: str, step: Optional[int] = None, with_optimizer: bool = True) -> str:
    os.makedirs(directory, exist_ok=True)
    payload = {"model": {k: v.cpu() for k, v in stage.state_dict().items()}, "stage_index": stage.stage_index,
               "num_stages": stage.num_stages, "step": step if step is not None else stage.step_count}
    if with_optimizer and getattr(stage, "_materialized", False):
        osd = stage.optimizer_state_dict()
        payload["optimizer"] = {"step": osd["step"], "state": {k: {kk: vv.cpu() for kk, vv in v.items()} for k, v in osd["state"].items()}}
    path = os.path.join(directory, f"stage_{stage.stage_index:03d}.pt")
    tmp = path + ".tmp"
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return path


def merge_shards(directory: str) -> Dict[str, torch.Tensor]:
    """Union of all stage shards = the full-model state_dict (loadable by the unmodified reference Transformer)."""
    full: Dict[str, torch.Tensor] = {}
    for p in sorted(glob.glob(os.path.join(directory, "stage_*.pt"))):
        shard = torch.load(p, map_location="cpu", weights_only=False)
        dup = set(full) & set(shard["model"])
        if dup:
            raise ValueError(f"{p}: keys present in two shards: {sorted(dup)[:3]}")
        full.update(shard["model"])
    return full


def load_stage(stage, directory: str, with_optimizer: bool = True) -> int:
    """Load whatever this stage owns from the shards in ``directory`` (works across different stage counts
    because keys carry global layer indices).  Returns the saved step."""
    full = merge_shards(directory)
    stage.load_state_dict(full, strict=True)
    step = 0
    if with_optimizer and getattr(stage, "_materialized", False):
        state, steps = {}, []
        for p in sorted(glob.glob(os.path.join(directory, "stage_*.pt"))):
            shard = torch.load(p, map_location="cpu", weights_only=False)
            if "optimizer" in shard:
                state.update(shard["optimizer"]["state"])
                steps.append(shard["optimizer"]["step"])
        if state:
            step = max(steps)
            stage.load_optimizer_state_dict({"state": state, "step": step})
    return step


def filter_for_stage(full: Dict[str, torch.Tensor], keys: Iterable[str]) -> Dict[str, torch.Tensor]:
    keys = set(keys)
    return {k: v for k, v in full.items() if k in keys}
