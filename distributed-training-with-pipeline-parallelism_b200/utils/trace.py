"""Chrome-trace export of MEASURED per-action timelines, merged over ranks.  The
reference has no tracing (two time.time() calls, helper:121/:133); torch only dumps the PLANNED schedule with
unit durations (schedules.py:3388-3421)."""
from __future__ import annotations

import json
from typing import List, Optional

import torch.distributed as dist


def gather_trace(schedule, dst: int = 0) -> Optional[List[dict]]:
    """Collect every rank's trace events on rank ``dst`` (gather_object; call on all ranks)."""
    evs = schedule.trace_events()
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return evs
    out = [None] * dist.get_world_size() if dist.get_rank() == dst else None
    dist.gather_object(evs, out, dst=dst)
    if dist.get_rank() != dst:
        return None
    return [e for part in out for e in part]


def write_chrome_trace(events: List[dict], path: str) -> None:
    with open(path, "w") as f:
        json.dump({"traceEvents": events, "displayTimeUnit": "ms"}, f)


def bubble_from_events(events: List[dict]) -> dict:
    """Per-rank measured bubble = 1 - busy/span over one step, and the max over ranks (BASELINE.json metric)."""
    per = {}
    for e in events:
        r = per.setdefault(e["pid"], {"busy": 0.0, "t0": float("inf"), "t1": 0.0})
        r["busy"] += e["dur"]
        r["t0"] = min(r["t0"], e["ts"])
        r["t1"] = max(r["t1"], e["ts"] + e["dur"])
    if not per:
        return {}
    t0 = min(r["t0"] for r in per.values())
    t1 = max(r["t1"] for r in per.values())
    span = t1 - t0
    out = {f"rank{k}": 1.0 - v["busy"] / span for k, v in per.items()}
    out["max"] = max(out.values())
    out["span_us"] = span
    return out
