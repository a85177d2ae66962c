"""Analysis / reporting layer (L5 of SURVEY.md section 1): speedup vs GPipe, the reference's 'efficiency',
pivot tables, bubble-vs-theory tables, and the two figures of the notebook (nb:405-435, nb:766-783, nb:842-849,
nb:880-941, nb:972-1002).  matplotlib runs headless (Agg)."""
from __future__ import annotations

import os
from typing import Optional


def compute_speedup_and_efficiency(df):
    """speedup = throughput / throughput_GPipe within a (n_layers, n_heads, num_processes) group (nb:421);
    efficiency = speedup / num_processes * 100 (nb:422 -- about 100/P by construction, not a scaling efficiency).
    Returns one row per non-GPipe schedule, like the notebook."""
    import pandas as pd
    rows = []
    for (nl, nh, npr), g in df.groupby(["n_layers", "n_heads", "num_processes"]):
        base = g[g["schedule"] == "GPipe"]["throughput"]
        if base.empty:
            continue
        gp = float(base.iloc[0])
        for _, r in g.iterrows():
            if r["schedule"] == "GPipe":
                continue
            sp = float(r["throughput"]) / gp
            rows.append({"n_layers": nl, "n_heads": nh, "num_processes": npr, "schedule": r["schedule"], "speedup": sp,
                         "efficiency": sp / npr * 100.0})
    return pd.DataFrame(rows)


def throughput_pivot(df):
    """Throughput by (n_layers, n_heads) x (schedule, num_processes) (nb:776-781)."""
    return df.pivot_table(values="throughput", index=["n_layers", "n_heads"], columns=["schedule", "num_processes"])


def scaling_ratio(df):
    """throughput(P=4) / throughput(P=2) per (model, schedule) -- SURVEY.md section 6.2 (ideal 2.0)."""
    import pandas as pd
    rows = []
    for (nl, nh, s), g in df.groupby(["n_layers", "n_heads", "schedule"]):
        t = dict(zip(g["num_processes"], g["throughput"]))
        if 2 in t and 4 in t:
            rows.append({"n_layers": nl, "n_heads": nh, "schedule": s, "ratio_4_over_2": t[4] / t[2]})
    return pd.DataFrame(rows)


def bubble_table(df):
    """Measured vs theoretical bubble per row (new: the reference never measures the bubble)."""
    cols = [c for c in ("n_layers", "n_heads", "num_processes", "schedule", "n_microbatches", "bubble_fraction",
                        "bubble_theory") if c in df.columns]
    out = df[cols].copy()
    if "bubble_fraction" in out and "bubble_theory" in out:
        out["bubble_minus_theory"] = out["bubble_fraction"] - out["bubble_theory"]
    return out


_COLORS = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f"]


def _svg_panels(panels, path: str, cols: int = 2, w: int = 420, h: int = 300) -> str:
    """Dependency-free line-plot renderer (matplotlib is not part of this image): ``panels`` is a list of
    (title, x_labels, {series label: [y...]}, hline or None); writes one SVG with a grid of panels."""
    rows = (len(panels) + cols - 1) // cols
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{cols * w}" height="{rows * h}" font-family="sans-serif" font-size="10">',
           '<rect width="100%" height="100%" fill="white"/>']
    for i, (title, xl, series, hline) in enumerate(panels):
        ox, oy = (i % cols) * w, (i // cols) * h
        l, r, t, b = ox + 50, ox + w - 110, oy + 24, oy + h - 50
        ys = [v for vals in series.values() for v in vals if v == v]
        if hline is not None:
            ys.append(hline)
        lo, hi = (min(ys), max(ys)) if ys else (0.0, 1.0)
        if hi - lo < 1e-12:
            hi = lo + 1.0
        pad = 0.05 * (hi - lo)
        lo, hi = lo - pad, hi + pad
        n = max(1, len(xl) - 1)
        X = lambda k: l + (r - l) * (k / n if len(xl) > 1 else 0.5)   # noqa: E731
        Y = lambda v: b - (b - t) * (v - lo) / (hi - lo)              # noqa: E731
        out.append(f'<text x="{(l + r) / 2}" y="{oy + 14}" text-anchor="middle" font-size="12">{title}</text>')
        out.append(f'<rect x="{l}" y="{t}" width="{r - l}" height="{b - t}" fill="none" stroke="#444"/>')
        for g in range(5):
            v = lo + (hi - lo) * g / 4
            out.append(f'<line x1="{l}" x2="{r}" y1="{Y(v):.1f}" y2="{Y(v):.1f}" stroke="#ddd"/>')
            out.append(f'<text x="{l - 4}" y="{Y(v) + 3:.1f}" text-anchor="end">{v:.3g}</text>')
        for k, lab in enumerate(xl):
            out.append(f'<text x="{X(k):.1f}" y="{b + 12}" text-anchor="end" transform="rotate(-40 {X(k):.1f} {b + 12})">{lab}</text>')
        if hline is not None:
            out.append(f'<line x1="{l}" x2="{r}" y1="{Y(hline):.1f}" y2="{Y(hline):.1f}" stroke="#000" stroke-dasharray="3,3"/>')
        for j, (lab, vals) in enumerate(series.items()):
            c = _COLORS[j % len(_COLORS)]
            pts = " ".join(f"{X(k):.1f},{Y(v):.1f}" for k, v in enumerate(vals) if v == v)
            out.append(f'<polyline points="{pts}" fill="none" stroke="{c}" stroke-width="1.5"/>')
            for k, v in enumerate(vals):
                if v == v:
                    out.append(f'<circle cx="{X(k):.1f}" cy="{Y(v):.1f}" r="2.2" fill="{c}"/>')
            out.append(f'<rect x="{r + 6}" y="{t + 12 * j}" width="8" height="8" fill="{c}"/>')
            out.append(f'<text x="{r + 18}" y="{t + 12 * j + 8}">{lab}</text>')
    out.append("</svg>")
    if not path.endswith(".svg"):
        path = path + ".svg" if "." not in os.path.basename(path) else path
    with open(path, "w") as f:
        f.write("\n".join(out))
    return path


def _have_matplotlib() -> bool:
    try:
        import matplotlib  # noqa: F401
        return True
    except Exception:  # noqa: BLE001
        return False


def plot_speedup_and_efficiency(speedup_df, path: Optional[str] = None):
    """1x2 figure: speedup and efficiency vs model config 'L{layers}_H{heads}' (nb:880-941).  Uses matplotlib (Agg)
    when it is installed, otherwise the built-in SVG renderer; returns the path written (or None)."""
    df = speedup_df.copy()
    df["config"] = df.apply(lambda r: f"L{int(r.n_layers)}_H{int(r.n_heads)}", axis=1)
    configs = sorted(df["config"].unique(), key=lambda c: tuple(int(x[1:]) for x in c.split("_")))
    sp, ef = {}, {}
    for (sched, p), g in df.groupby(["schedule", "num_processes"]):
        by = dict(zip(g["config"], zip(g["speedup"], g["efficiency"])))
        sp[f"{sched} P={p}"] = [by.get(c, (float("nan"),) * 2)[0] for c in configs]
        ef[f"{sched} P={p}"] = [by.get(c, (float("nan"),) * 2)[1] for c in configs]
    if _have_matplotlib():
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig, axes = plt.subplots(1, 2, figsize=(14, 5))
        for lab in sp:
            axes[0].plot(configs, sp[lab], marker="o", label=lab)
            axes[1].plot(configs, ef[lab], marker="o", label=lab)
        axes[0].axhline(1.0, color="k", lw=0.5)
        axes[0].set_title("Speedup vs GPipe")
        axes[1].set_title("Efficiency (speedup / P * 100)")
        for ax in axes:
            ax.tick_params(axis="x", rotation=45)
            ax.legend(fontsize=7)
        fig.tight_layout()
        if path:
            fig.savefig(path, dpi=110)
        plt.close(fig)
        return path
    if path is None:
        return None
    return _svg_panels([("Speedup vs GPipe", configs, sp, 1.0), ("Efficiency (speedup / P * 100)", configs, ef, None)], path)


def plot_throughput_grid(df, path: Optional[str] = None):
    """Grid of panels: throughput vs num_processes for every (layers, heads) (nb:972-1002)."""
    layers = sorted(df["n_layers"].unique())
    heads = sorted(df["n_heads"].unique())
    panels = []
    for nl in layers:
        for nh in heads:
            sub = df[(df.n_layers == nl) & (df.n_heads == nh)]
            procs = sorted(sub["num_processes"].unique())
            series = {}
            for sched, g in sub.groupby("schedule"):
                by = dict(zip(g["num_processes"], g["throughput"]))
                series[sched] = [by.get(p, float("nan")) for p in procs]
            panels.append((f"L{nl} H{nh} (tokens/s vs processes)", [str(p) for p in procs], series, None))
    if _have_matplotlib():
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig, axes = plt.subplots(len(layers), len(heads), figsize=(4 * len(heads), 3.2 * len(layers)), squeeze=False)
        for k, (title, xl, series, _) in enumerate(panels):
            ax = axes[k // len(heads)][k % len(heads)]
            for lab, vals in series.items():
                ax.plot(xl, vals, marker="o", label=lab)
            ax.set_title(title, fontsize=9)
            ax.legend(fontsize=6)
        fig.tight_layout()
        if path:
            fig.savefig(path, dpi=110)
        plt.close(fig)
        return path
    if path is None:
        return None
    return _svg_panels(panels, path, cols=len(heads))


def environment_report() -> dict:
    """The notebook's environment probe (nb:256-259), plus what matters on the GPU box."""
    import os
    import torch
    info = {"torch": torch.__version__, "cuda_available": torch.cuda.is_available(), "cpu_count": os.cpu_count()}
    if torch.cuda.is_available():
        info["gpu"] = torch.cuda.get_device_name(0)
        info["gpu_count"] = torch.cuda.device_count()
        info["capability"] = ".".join(map(str, torch.cuda.get_device_capability(0)))
    return info
