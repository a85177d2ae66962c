"""Schedule diagrams: render a (simulated or measured) pipeline timeline as SVG or ASCII.

Counterpart of the four concept figures of the reference notebook (bubble.png, gpipe.png, 1f1b.png,
interleave.png -- referenced at nb:63/102/131/160 but not shipped) and of torch's
matplotlib schedule visualiser (_schedule_visualizer.py); here the pictures are generated from this framework's own
IR + simulator (or from a measured chrome-trace event list), with no plotting dependency.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

from ..parallel.schedule import Action, ActionType as T, build_schedule, simulate, theoretical_bubble

_F_COLORS = ["#4c8eda", "#6aa5e8", "#8fbdf2", "#b5d3f7"]
_B_COLORS = ["#3aa655", "#5fbf77", "#8ad59b", "#b6e6c0"]
_OTHER = {"OPTIM": "#b07cc6", "REDUCE_GRAD": "#999999", "ALLREDUCE": "#e39b3d", "I": "#3aa655", "W": "#d0b13c"}


def timeline_of(name: str, pp_size: int, n_microbatches: int, n_virtual: int = 1, f_cost: float = 1.0,
                b_cost: float = 2.0, comm_latency: float = 0.0):
    """(per-rank [(action, start, end)], makespan, per-rank bubble) of a schedule under the simulator."""
    order = build_schedule(name, pp_size, n_microbatches, n_virtual=n_virtual, with_optimizer=False)
    res = simulate(order, pp_size * n_virtual, comm_latency=comm_latency, f_cost=f_cost / n_virtual, b_cost=b_cost / n_virtual)
    return res.timeline, res.makespan, res.bubble


def _color(a: Action, pp_size: int) -> str:
    chunk = a.stage // max(1, pp_size)
    if a.type == T.F:
        return _F_COLORS[chunk % len(_F_COLORS)]
    if a.type in (T.B, T.I):
        return _B_COLORS[chunk % len(_B_COLORS)]
    return _OTHER.get(a.type.value, "#cccccc")


def svg_timeline(timeline: Dict[int, Sequence[Tuple[Action, float, float]]], makespan: float, title: str = "",
                 pp_size: Optional[int] = None, width: int = 980, row_h: int = 30) -> str:
    ranks = sorted(timeline)
    P = pp_size or len(ranks)
    left, top = 70, 36
    W = width - left - 20
    H = top + row_h * len(ranks) + 34
    sx = W / makespan if makespan > 0 else 1.0
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{H}" font-family="sans-serif" font-size="10">',
           '<rect width="100%" height="100%" fill="white"/>',
           f'<text x="{left}" y="18" font-size="13">{title}</text>']
    for i, r in enumerate(ranks):
        y = top + i * row_h
        out.append(f'<text x="{left - 8}" y="{y + row_h * 0.62:.0f}" text-anchor="end">rank {r}</text>')
        out.append(f'<rect x="{left}" y="{y + 2}" width="{W}" height="{row_h - 4}" fill="#f0f0f0" stroke="#ccc"/>')  # grey = bubble
        for a, s, e in timeline[r]:
            w = max(0.5, (e - s) * sx)
            out.append(f'<rect x="{left + s * sx:.1f}" y="{y + 2}" width="{w:.1f}" height="{row_h - 4}" fill="{_color(a, P)}" stroke="white" stroke-width="0.5"/>')
            if w > 16 and a.mb is not None:
                out.append(f'<text x="{left + (s + e) / 2 * sx:.1f}" y="{y + row_h * 0.62:.0f}" text-anchor="middle" fill="white">{a.type.value}{a.mb}</text>')
    y = top + row_h * len(ranks) + 16
    out.append(f'<rect x="{left}" y="{y - 9}" width="10" height="10" fill="{_F_COLORS[0]}"/><text x="{left + 14}" y="{y}">forward</text>')
    out.append(f'<rect x="{left + 70}" y="{y - 9}" width="10" height="10" fill="{_B_COLORS[0]}"/><text x="{left + 84}" y="{y}">backward</text>')
    out.append(f'<rect x="{left + 150}" y="{y - 9}" width="10" height="10" fill="#f0f0f0" stroke="#ccc"/><text x="{left + 164}" y="{y}">idle (bubble)</text>')
    out.append(f'<rect x="{left + 250}" y="{y - 9}" width="10" height="10" fill="{_OTHER["W"]}"/><text x="{left + 264}" y="{y}">weight gradient (W)</text>')
    out.append("</svg>")
    return "\n".join(out)


def ascii_timeline(timeline: Dict[int, Sequence[Tuple[Action, float, float]]], makespan: float, unit: float = 1.0) -> str:
    """One character per ``unit`` of time: F/B digits = micro-batch id (mod 10), '.' = idle."""
    n = int(round(makespan / unit))
    rows = []
    for r in sorted(timeline):
        cells = ["."] * n
        for a, s, e in timeline[r]:
            for k in range(int(round(s / unit)), min(n, int(round(e / unit)))):
                ch = str((a.mb or 0) % 10)
                cells[k] = ch if a.type == T.F else (chr(ord("a") + (a.mb or 0) % 26) if a.type == T.B else "*")
        rows.append(f"rank{r}: " + "".join(cells))
    return "\n".join(rows)


def timeline_from_trace(events: List[dict]):
    """Measured chrome-trace events (runner.trace_events / utils.trace.gather_trace) -> the same timeline structure."""
    tl: Dict[int, List[Tuple[Action, float, float]]] = {}
    t0 = min((e["ts"] for e in events), default=0.0)
    end = 0.0
    for e in events:
        a = Action.from_str(e["name"])
        s = (e["ts"] - t0) / 1e3
        tl.setdefault(e["pid"], []).append((a, s, s + e["dur"] / 1e3))
        end = max(end, s + e["dur"] / 1e3)
    return tl, end


def write_concept_figures(out_dir: str, pp_size: int = 4, n_microbatches: int = 8) -> List[str]:
    """bubble / GPipe / 1F1B / interleaved figures for docs/CONCEPTS.md."""
    import os
    os.makedirs(out_dir, exist_ok=True)
    paths = []
    specs = [("bubble", "GPipe", pp_size, 1, 1, "Naive model parallelism (1 micro-batch): only one stage works at a time"),
             ("gpipe", "GPipe", pp_size, n_microbatches, 1, f"GPipe, {pp_size} stages x {n_microbatches} micro-batches"),
             ("1f1b", "1F1B", pp_size, n_microbatches, 1, f"1F1B, {pp_size} stages x {n_microbatches} micro-batches (same bubble, <= P micro-batches in flight)"),
             ("interleave", "Interleaved1F1B", pp_size, n_microbatches, 2, f"Interleaved 1F1B, {pp_size} ranks x 2 virtual stages x {n_microbatches} micro-batches"),
             ("zerobubble", "ZeroBubble1F1B", pp_size, n_microbatches, 1, f"ZeroBubble1F1B, {pp_size} stages x {n_microbatches} micro-batches: backward = I (green) + W (yellow)")]
    for fname, sched, P, m, v, title in specs:
        tl, span, bub = timeline_of(sched, P, m, v)
        title = f"{title} -- bubble {max(bub.values()):.3f} (theory {theoretical_bubble(P, m, v):.3f})"
        path = os.path.join(out_dir, fname + ".svg")
        with open(path, "w") as f:
            f.write(svg_timeline(tl, span, title, P))
        paths.append(path)
    return paths
