#!/usr/bin/env python
"""The reference notebook, cell by cell, on this framework's API (CPU / gloo by default, like the notebook; the same
calls run on B200 -- one process per GPU over NCCL -- when CUDA is present).

    python examples/notebook_flow.py                     # small grid, a minute on a laptop-class CPU
    python examples/notebook_flow.py --full              # the notebook's 54 configurations (nb:345-349)
    python examples/notebook_flow.py --out /tmp/report   # tables (markdown) + figures (SVG) are written there

Notebook section -> call:
    environment cell (nb:24-44)            ppb200.utils.analysis.environment_report()
    "Running experiments" (nb:345-394)     ppb200.run_all_experiments(...)      spawns ranks, error-dict contract, timeouts
    results table / pivots (nb:640-700)    analysis.throughput_pivot(df)
    speedup & efficiency (nb:405-435)      analysis.compute_speedup_and_efficiency(df)
    plots (nb:460-560)                     analysis.plot_speedup_and_efficiency / plot_throughput_grid
    concept figures (nb:60-280)            docs/CONCEPTS.md; regenerated here from the schedule IR + simulator
plus what the notebook cannot show: the MEASURED bubble of every run next to (p-1)/(v m + p-1)."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="the notebook's grid: layers x heads in {4, 8, 12}, 2 and 4 processes")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "notebook_flow"))
    ap.add_argument("--iterations", type=int, default=None)
    a = ap.parse_args()
    import ppb200
    from ppb200.utils import analysis, diagram

    os.makedirs(a.out, exist_ok=True)
    print("== environment ==")
    for k, v in analysis.environment_report().items():
        print(f"  {k}: {v}")

    print("== experiments ==")
    if a.full:
        grid = dict(layers=(4, 8, 12), heads=(4, 8, 12), processes=(2, 4), num_iterations=a.iterations or 5)
        extra = {}
    else:   # same code path, a model small enough for a quick look
        grid = dict(layers=(4, 8), heads=(4,), processes=(2,), num_iterations=a.iterations or 2)
        extra = dict(batch_size=8, seq_length=32, dim=64, vocab_size=256, warmup=1)
    df = ppb200.run_all_experiments(**grid, **extra)
    if df.empty:
        raise SystemExit("no experiment returned a result")

    print("== throughput (tokens/s) ==")
    pivot = analysis.throughput_pivot(df)
    print(pivot.round(1).to_string())
    print("== speedup vs GPipe / efficiency ==")
    sp = analysis.compute_speedup_and_efficiency(df)
    print(sp.round(4).to_string(index=False))
    print("== measured vs theoretical bubble ==")
    bub = analysis.bubble_table(df)
    print(bub.round(4).to_string(index=False))

    with open(os.path.join(a.out, "report.md"), "w") as f:
        f.write("# Pipeline-parallel experiment grid\n\n## Throughput (tokens/s)\n\n```\n" + pivot.round(1).to_string() +
                "\n```\n\n## Speedup vs GPipe / efficiency\n\n```\n" + sp.round(4).to_string(index=False) +
                "\n```\n\n## Measured vs theoretical bubble\n\n```\n" + bub.round(4).to_string(index=False) + "\n```\n")
    analysis.plot_speedup_and_efficiency(sp, os.path.join(a.out, "speedup_efficiency.svg"))
    analysis.plot_throughput_grid(df, os.path.join(a.out, "throughput.svg"))
    diagram.write_concept_figures(os.path.join(a.out, "figures"))   # bubble / GPipe / 1F1B / interleaved / zero-bubble timelines
    print(f"wrote {a.out}/report.md and figures")


if __name__ == "__main__":
    main()
