"""Checkpoint / resume (the reference has none).  Format = the reference ``state_dict``
layout: each pipeline stage writes ``stage_{i:03d}.pt`` holding a strict subset of the full model's keys (global
layer indices), so merging is a dict union and re-sharding to a different stage count is a dict filter.

Tensor-parallel stages write one file per (stage, tp rank), ``stage_{i:03d}_tp{r}.pt``, holding that rank's slices under
the same key names plus the facts needed to undo the split (``models/tp.py``: packed QKV is split per head within each
of q, k, v; SwiGLU's packed gate|up per half; out_proj / linear2 on dim 1).  ``merge_shards`` re-assembles the
reference layout first, so a checkpoint written at (pp, tp) = (4, 2) loads at (8, 1), (2, 4), or into the unmodified
reference model.  Optimizer moments have parameter shapes and travel the same way."""
from __future__ import annotations

import glob
import os
from typing import Dict, Iterable, Optional

import torch


def save_stage(stage, directory: str, step: Optional[int] = None, with_optimizer: bool = True) -> str:
    os.makedirs(directory, exist_ok=True)
    payload = {"model": {k: v.cpu() for k, v in stage.state_dict().items()}, "stage_index": stage.stage_index,
               "num_stages": stage.num_stages, "step": step if step is not None else stage.step_count}
    tp_size = int(getattr(stage, "tp_size", 1))
    if tp_size > 1:
        payload["tp"] = {"rank": int(stage.submod.tp_rank), "size": tp_size, "n_heads": stage.args.n_heads,
                         "swiglu": stage.args.activation == "swiglu"}
    if with_optimizer and getattr(stage, "_materialized", False):
        osd = stage.optimizer_state_dict()
        payload["optimizer"] = {"step": osd["step"], "state": {k: {kk: vv.cpu() for kk, vv in v.items()} for k, v in osd["state"].items()}}
    suffix = f"_tp{payload['tp']['rank']}" if tp_size > 1 else ""
    path = os.path.join(directory, f"stage_{stage.stage_index:03d}{suffix}.pt")
    tmp = path + ".tmp"
    torch.save(payload, tmp)
    os.replace(tmp, path)
    return path


def _load_stage_files(directory: str):
    """[(payload with tensor-parallel slices merged back), ...] one entry per pipeline stage, in stage order."""
    from ..models.tp import merge_tp_shards
    by_stage: Dict[int, list] = {}
    for p in sorted(glob.glob(os.path.join(directory, "stage_*.pt"))):
        shard = torch.load(p, map_location="cpu", weights_only=False)
        by_stage.setdefault(int(shard["stage_index"]), []).append((p, shard))
    out = []
    for idx in sorted(by_stage):
        files = by_stage[idx]
        if len(files) == 1 and "tp" not in files[0][1]:
            out.append(files[0])
            continue
        files.sort(key=lambda ps: ps[1]["tp"]["rank"])
        meta = files[0][1]["tp"]
        if [f[1]["tp"]["rank"] for f in files] != list(range(meta["size"])):
            raise ValueError(f"stage {idx}: need tensor-parallel shards 0..{meta['size'] - 1}, found "
                             f"{[f[1]['tp']['rank'] for f in files]}")
        merged = dict(files[0][1])
        merged["model"] = merge_tp_shards([f[1]["model"] for f in files], meta["n_heads"], meta["swiglu"])
        if all("optimizer" in f[1] for f in files):
            names = files[0][1]["optimizer"]["state"].keys()
            slots = files[0][1]["optimizer"]["state"][next(iter(names))].keys() if names else []
            state = {n: {} for n in names}
            for slot in slots:
                per_rank = [{n: f[1]["optimizer"]["state"][n][slot] for n in names} for f in files]
                for n, t in merge_tp_shards(per_rank, meta["n_heads"], meta["swiglu"]).items():
                    state[n][slot] = t
            merged["optimizer"] = {"step": files[0][1]["optimizer"]["step"], "state": state}
        merged.pop("tp")
        out.append((files[0][0], merged))
    return out


def merge_shards(directory: str) -> Dict[str, torch.Tensor]:
    """Union of all stage shards = the full-model state_dict (loadable by the unmodified reference Transformer)."""
    full: Dict[str, torch.Tensor] = {}
    for p, shard in _load_stage_files(directory):
        dup = set(full) & set(shard["model"])
        if dup:
            raise ValueError(f"{p}: keys present in two shards: {sorted(dup)[:3]}")
        full.update(shard["model"])
    return full


def _for_tp_rank(stage, sd: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    tp_size = int(getattr(stage, "tp_size", 1))
    if tp_size == 1:
        return sd
    from ..models.tp import shard_tensor
    return {k: shard_tensor(k, v, stage.args.n_heads, stage.submod.tp_rank, tp_size, stage.args.activation == "swiglu")
            for k, v in sd.items()}


def load_stage(stage, directory: str, with_optimizer: bool = True) -> int:
    """Load whatever this stage owns from the shards in ``directory`` (works across different stage counts and
    tensor-parallel degrees because keys carry global layer indices and slices are re-cut from the merged tensors).
    Returns the saved step."""
    files = _load_stage_files(directory)
    full: Dict[str, torch.Tensor] = {}
    for _, shard in files:
        full.update(shard["model"])
    own = set(stage.state_dict().keys())
    stage.load_state_dict(_for_tp_rank(stage, {k: v for k, v in full.items() if k in own}), strict=True)
    step = 0
    if with_optimizer and getattr(stage, "_materialized", False):
        state, steps = {}, []
        for _, shard in files:
            if "optimizer" in shard:
                state.update(shard["optimizer"]["state"])
                steps.append(shard["optimizer"]["step"])
        if state:
            step = max(steps)
            mine = {}
            for slot in next(iter(state.values())).keys():
                cut = _for_tp_rank(stage, {n: st[slot] for n, st in state.items() if n in own})
                for n, t in cut.items():
                    mine.setdefault(n, {})[slot] = t
            stage.load_optimizer_state_dict({"state": mine, "step": step})
    if not step:
        # no optimizer state (saved or requested without it): the step recorded beside the parameters still tells the
        # training loop where the LR schedule / data stream were, and seeds the dropout streams (stage.step_count)
        step = max([int(shard.get("step", 0) or 0) for _, shard in files] or [0])
        if step and hasattr(stage, "step_count"):
            stage.step_count = step
    return step


def filter_for_stage(full: Dict[str, torch.Tensor], keys: Iterable[str]) -> Dict[str, torch.Tensor]:
    keys = set(keys)
    return {k: v for k, v in full.items() if k in keys}
