#!/usr/bin/env python
"""Join the per-arm outputs of benchmarks/sweep_vs_reference.py into one table (markdown to stdout, JSON with --out):

    python benchmarks/sweep_merge.py profiles/sweep_n2_ours.json profiles/sweep_n2_reference.json ... [--out profiles/sweep_r2.json]
"""
import argparse
import json

ap = argparse.ArgumentParser()
ap.add_argument("files", nargs="+")
ap.add_argument("--out", default=None)
a = ap.parse_args()
cells = {}
for f in a.files:
    for r in json.load(open(f)):
        key = (r["n_gpus"], r["schedule"], r["n_microbatches"])
        cells.setdefault(key, {}).update(r)
rows = []
for key in sorted(cells, key=lambda k: (k[0], ["GPipe", "1F1B", "Interleaved1F1B"].index(k[1]) if k[1] in ("GPipe", "1F1B", "Interleaved1F1B") else 9, k[2])):
    r = cells[key]
    if "ours_tok_s" in r and "reference_tok_s" in r:
        r["ratio"] = round(r["ours_tok_s"] / r["reference_tok_s"], 3)
    rows.append(r)
print("| GPUs | schedule | m | ours tok/s | reference tok/s | ratio | bubble (critical rank) | bubble (max) | theory | partition | v ours / ref |")
print("|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows:
    print(f"| {r['n_gpus']} | {r['schedule']} | {r['n_microbatches']} | {r.get('ours_tok_s', r.get('ours_error', '-'))} | "
          f"{r.get('reference_tok_s', r.get('reference_error', r.get('reference_unavailable', '-')))} | {r.get('ratio', '-')} | "
          f"{r.get('bubble_critical', '-')} | {r.get('bubble_max', '-')} | {r.get('bubble_theory', '-')} | {r.get('partition', '-')} | "
          f"{r.get('virtual_stages', '-')} / {r.get('reference_virtual_stages', '-')} |")
if a.out:
    json.dump(rows, open(a.out, "w"), indent=1)
