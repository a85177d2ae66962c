"""Analysis / reporting layer (L5 of SURVEY.md section 1): speedup vs GPipe, the reference's 'efficiency',
pivot tables, bubble-vs-theory tables, and the two figures of the notebook (nb:405-435, nb:766-783, nb:842-849,
nb:880-941, nb:972-1002).  matplotlib runs headless (Agg)."""
from __future__ import annotations

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


def plot_speedup_and_efficiency(speedup_df, path: Optional[str] = None):
    """1x2 figure: speedup and efficiency vs model config 'L{layers}_H{heads}' (nb:880-941)."""
    import matplotlib
    matplotlib.use("Agg")
    import matplotlib.pyplot as plt
    fig, axes = plt.subplots(1, 2, figsize=(14, 5))
    df = speedup_df.copy()
    df["config"] = df.apply(lambda r: f"L{int(r.n_layers)}_H{int(r.n_heads)}", axis=1)
    for (sched, p), g in df.groupby(["schedule", "num_processes"]):
        axes[0].plot(g["config"], g["speedup"], marker="o", label=f"{sched} P={p}")
        axes[1].plot(g["config"], g["efficiency"], marker="o", label=f"{sched} P={p}")
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
    return fig


def plot_throughput_grid(df, path: Optional[str] = None):
    """3x3 grid: throughput vs num_processes for every (layers, heads) (nb:972-1002)."""
    import matplotlib
    matplotlib.use("Agg")
    import matplotlib.pyplot as plt
    layers = sorted(df["n_layers"].unique())
    heads = sorted(df["n_heads"].unique())
    fig, axes = plt.subplots(len(layers), len(heads), figsize=(4 * len(heads), 3.2 * len(layers)), squeeze=False)
    for i, nl in enumerate(layers):
        for j, nh in enumerate(heads):
            ax = axes[i][j]
            sub = df[(df.n_layers == nl) & (df.n_heads == nh)]
            for sched, g in sub.groupby("schedule"):
                g = g.sort_values("num_processes")
                ax.plot(g["num_processes"], g["throughput"], marker="o", label=sched)
            ax.set_title(f"L{nl} H{nh}", fontsize=9)
            ax.set_xlabel("processes")
            ax.set_ylabel("tokens/s")
            ax.legend(fontsize=6)
    fig.tight_layout()
    if path:
        fig.savefig(path, dpi=110)
    plt.close(fig)
    return fig


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
