"""Pick (schedule, virtual stages, micro-batches, partition rule) for a model and a GPU count WITHOUT running it:
the lowered schedule is executed by the discrete-event simulator with per-action costs derived from the calibrated
cost model of ``parallel/partition.py`` (or from measured per-layer times), and the candidates are ranked by predicted
step time.  The prediction tracks the measured B200 steps within a few per cent (profiles/schedule_sweep_predicted.md).

    python -m ppb200.utils.autotune --model ref-gpt2-medium --gpus 8 --global-batch 128 --seq-len 1024

The reference picks its schedule by hand and fixes 4 micro-batches (helper:214); this is the tool that answers the
notebook's question ("which schedule is fastest for this model on this many ranks?") before spending GPU time.
"""
from __future__ import annotations

import argparse
import json
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

from ..models import ModelArgs, preset
from ..parallel.partition import layer_units, part_units, stage_layer_parts, unit_costs
from ..parallel.schedule import ActionType as T, build_schedule, simulate, theoretical_bubble


@dataclass
class UnitCosts:
    """Milliseconds per 4096-token micro-batch on one B200 (flagship measurements, profiles/bench_r1_n4_1f1b_v2.json)."""
    layer_f: float = 0.333
    layer_b: float = 0.667
    head_f: float = 0.67          # embedding + final norm + LM head + loss
    head_b: float = 0.79
    optim: float = 0.7
    i_frac: float = 0.78          # input-gradient share of a backward
    w_frac: float = 0.23
    attn_share: float = 0.8       # share of a layer's cost in its attention sub-blocks (rule 'fine')
    # Per-token cost relative to a 4096-token micro-batch: kernels are not linear in the micro-batch size (launch / prologue
    # costs, wave quantisation of the attention grid).  Calibrated on the measured 2-rank cells of the round-2 sweep and
    # validated out of sample on the 4- and 8-rank cells (profiles/schedule_sweep_predicted_r2.md); log-linear in between.
    token_factor: tuple = ((2048, 1.33), (4096, 1.0), (8192, 0.875), (16384, 0.83))

    def per_token(self, tokens: float) -> float:
        import math
        pts = self.token_factor
        if tokens <= pts[0][0]:
            return pts[0][1] * (pts[0][0] / max(tokens, 1.0)) ** 0.4      # smaller still: fixed costs take over
        if tokens >= pts[-1][0]:
            return pts[-1][1]
        for (t0, f0), (t1, f1) in zip(pts, pts[1:]):
            if t0 <= tokens <= t1:
                w = (math.log(tokens) - math.log(t0)) / (math.log(t1) - math.log(t0))
                return f0 + w * (f1 - f0)
        return 1.0

    @staticmethod
    def from_model(args: ModelArgs, seq_len: int, ref: Optional["UnitCosts"] = None) -> "UnitCosts":
        """Scale the measured flagship costs to another model with the FLOP / efficiency model of the partitioner."""
        ref = ref or UnitCosts()
        base = preset("ref-gpt2-medium")

        def layer(a: ModelArgs, s: int) -> float:
            blocks = 2 if a.cross_attn else 1
            return ((blocks * 8 * a.dim * a.dim + (6 if a.activation == "swiglu" else 4) * a.dim * a.ffn_dim) / 0.40
                    + blocks * 4 * s * a.dim * (0.5 if a.causal else 1.0) / 0.20)

        def head(a: ModelArgs) -> float:
            return 2 * a.dim * a.vocab_size / 0.9 + 650.0 * a.vocab_size

        kl, kh = layer(args, seq_len) / layer(base, 1024), head(args) / head(base)
        blocks = 2 if args.cross_attn else 1
        attn = blocks * 8 * args.dim ** 2 / 0.40 + blocks * 4 * seq_len * args.dim * (0.5 if args.causal else 1.0) / 0.20
        return UnitCosts(ref.layer_f * kl, ref.layer_b * kl, ref.head_f * kh, ref.head_b * kh, ref.optim, ref.i_frac,
                         ref.w_frac, attn / layer(args, seq_len), ref.token_factor)


def predict(args: ModelArgs, pp: int, schedule: str, n_microbatches: int, global_batch: int, seq_len: int,
            n_virtual: int = 1, rule: str = "balanced", costs: Optional[UnitCosts] = None) -> Dict[str, float]:
    """Simulated step of one configuration: {'ms', 'idle_critical', 'idle_max', 'theory', 'tokens_per_s'}."""
    c = costs or UnitCosts.from_model(args, seq_len)
    S = pp * n_virtual
    # layer-equivalents per stage: unit shares from the partitioner's cost table (measured when calibrated, else its model)
    uc, units = unit_costs(args, seq_len), layer_units(args)
    layer_total = sum(uc[u] for u in units)
    weights = []
    for s in range(S):
        w = 0.0
        for _, part in stage_layer_parts(args.n_layers, S, s, rule, args, seq_len):
            w += sum(uc[u] for u in part_units(part, units)) / layer_total
        weights.append(w)
    tokens = (global_batch / n_microbatches) * seq_len
    scale = tokens / 4096.0 * c.per_token(tokens)

    def cost(a) -> float:
        f = weights[a.stage] * c.layer_f + (c.head_f if a.stage == S - 1 else 0.0)
        b = weights[a.stage] * c.layer_b + (c.head_b if a.stage == S - 1 else 0.0)
        return {T.F: f * scale, T.B: b * scale, T.I: b * scale * c.i_frac, T.W: b * scale * c.w_frac,
                T.OPTIM: c.optim / n_virtual}.get(a.type, 0.0)

    order = build_schedule(schedule, pp, n_microbatches, n_virtual=n_virtual)
    res = simulate(order, S, cost=cost)
    return {"ms": res.makespan, "idle_critical": min(res.bubble.values()), "idle_max": max(res.bubble.values()),
            "theory": theoretical_bubble(pp, n_microbatches, n_virtual),
            "tokens_per_s": global_batch * seq_len / (res.makespan / 1e3)}


def recommend(args: ModelArgs, pp: int, global_batch: int, seq_len: int, microbatches: Optional[Sequence[int]] = None,
              max_virtual: int = 3, costs: Optional[UnitCosts] = None, min_tokens_per_microbatch: int = 4096) -> List[dict]:
    """Every valid (schedule, v, m, rule) for ``pp`` pipeline ranks, best predicted step first.  Micro-batches smaller than
    ``min_tokens_per_microbatch`` are skipped: below that the kernels lose efficiency, which the simulator does not model."""
    if microbatches is None:
        microbatches = [m for m in (4, 8, 16, 32, 64, 128) if global_batch % m == 0]
    rows: List[dict] = []
    rules = ("balanced", "fine")
    for sched in ("GPipe", "1F1B", "Interleaved1F1B", "ZeroBubble1F1B", "ZeroBubbleInterleaved"):
        for v in (range(2, max_virtual + 1) if "Interleaved" in sched else (1,)):
            if args.n_layers < pp * v:
                continue
            for m in microbatches:
                if (global_batch // m) * seq_len < min_tokens_per_microbatch:
                    continue
                for rule in rules:
                    try:
                        p = predict(args, pp, sched, m, global_batch, seq_len, v, rule, costs)
                    except (ValueError, RuntimeError):
                        continue        # the schedule's own constraints (m % p, m >= p, ...) reject the combination
                    rows.append({"schedule": sched, "virtual_stages": v, "n_microbatches": m, "partition": rule, **p})
    rows.sort(key=lambda r: r["ms"])
    return rows


def main(argv=None) -> List[dict]:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--model", default="ref-gpt2-medium")
    ap.add_argument("--gpus", type=int, default=8, help="pipeline ranks")
    ap.add_argument("--global-batch", type=int, default=128)
    ap.add_argument("--seq-len", type=int, default=1024)
    ap.add_argument("--top", type=int, default=10)
    ap.add_argument("--json", action="store_true")
    a = ap.parse_args(argv)
    rows = recommend(preset(a.model), a.gpus, a.global_batch, a.seq_len)
    if a.json:
        print(json.dumps(rows[:a.top]))
    else:
        print(f"{'schedule':18s} {'v':>2s} {'m':>4s} {'partition':10s} {'pred ms':>9s} {'tokens/s':>11s} {'idle crit':>9s} {'theory':>7s}")
        for r in rows[:a.top]:
            print(f"{r['schedule']:18s} {r['virtual_stages']:2d} {r['n_microbatches']:4d} {r['partition']:10s} {r['ms']:9.1f} "
                  f"{r['tokens_per_s']:11.0f} {r['idle_critical']:9.3f} {r['theory']:7.3f}")
    return rows


if __name__ == "__main__":
    main()
