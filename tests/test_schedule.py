"""Unit tests of the schedule IR, generators, lowering and simulator ('Unit - schedule IR').
Golden strings are the torch-generated IR recorded in SURVEY.md section 3.5 for the reference's largest
interleaved case (P=4, v=2, m=4)."""
import pytest

from ppb200.parallel.schedule import (Action, ActionType, build_schedule, format_schedule, parse_schedule, simulate,
                                      theoretical_bubble, validate_schedule, gpipe_order, one_f_one_b_order,
                                      interleaved_order, add_reduce_grad, add_send_recv, stage_to_rank_loop)
from ppb200.parallel.schedule import ir
from ppb200.parallel.schedule import ActionType as T
from ppb200.parallel.schedule.simulate import DeadlockError

GOLDEN_RANK0 = ("0F0 0SEND_F0 4RECV_F0 0F1 0SEND_F1 4RECV_F1 0F2 0SEND_F2 4RECV_F2 0F3 0SEND_F3 4RECV_F3 "
                "4F0 4SEND_F0 4F1 4SEND_F1 4F2 4SEND_F2 4F3 4SEND_F3 "
                "4RECV_B0 4B0 4SEND_B0 4RECV_B1 4B1 4SEND_B1 4RECV_B2 4B2 4SEND_B2 4RECV_B3 4B3 4SEND_B3 4REDUCE_GRAD "
                "0RECV_B0 0B0 0RECV_B1 0B1 0RECV_B2 0B2 0RECV_B3 0B3 0REDUCE_GRAD")
GOLDEN_RANK3_PREFIX = "3RECV_F0 3F0 3SEND_F0"
GOLDEN_RANK3_TAIL = ("7RECV_F0 7F0 7RECV_F1 7B0 7SEND_B0 7RECV_F2 7F1 7B1 7SEND_B1 "
                     "7RECV_F3 7F2 3RECV_B0 7B2 7SEND_B2 7F3 3RECV_B1 7B3 7SEND_B3 3RECV_B2 7REDUCE_GRAD 3RECV_B3 "
                     "3B0 3SEND_B0 3B1 3SEND_B1 3B2 3SEND_B2 3B3 3SEND_B3 3REDUCE_GRAD")


def test_action_text_roundtrip():
    for s in ["2F0", "3SEND_B1", "4REDUCE_GRAD", "11RECV_F31", "0B0", "5I2", "5W2", "1ALLREDUCE", "0OPTIM"]:
        assert str(Action.from_str(s)) == s
    with pytest.raises(ValueError):
        Action.from_str("F0")
    with pytest.raises(ValueError):
        Action.from_str("2REDUCE_GRAD1")
    with pytest.raises(ValueError):
        Action.from_str("2F")


def test_golden_interleaved_p4_v2_m4():
    order = build_schedule("Interleaved1F1B", 4, 4, n_virtual=2, with_optimizer=False)
    assert " ".join(map(str, order[0])) == GOLDEN_RANK0
    r3 = " ".join(map(str, order[3]))
    assert r3.startswith(GOLDEN_RANK3_PREFIX)
    assert r3.endswith(GOLDEN_RANK3_TAIL)


def test_text_and_csv_roundtrip():
    order = build_schedule("1F1B", 4, 8)
    assert parse_schedule(format_schedule(order)) == {r: list(a) for r, a in order.items()}
    assert ir.from_csv(ir.to_csv(order)) == {r: list(a) for r, a in order.items()}
    g = ir.grid(order, error_at=(1, 3))
    assert "<-- ERROR HERE" in g and g.count("\n") == 3


@pytest.mark.parametrize("name,P,v,m", [
    ("GPipe", 2, 1, 4), ("GPipe", 4, 1, 8), ("GPipe", 8, 1, 32), ("GPipe", 4, 1, 3),
    ("1F1B", 2, 1, 4), ("1F1B", 4, 1, 4), ("1F1B", 4, 1, 8), ("1F1B", 8, 1, 64),
    ("Interleaved1F1B", 2, 2, 4), ("Interleaved1F1B", 4, 2, 4), ("Interleaved1F1B", 4, 2, 8),
    ("Interleaved1F1B", 8, 2, 32), ("Interleaved1F1B", 4, 3, 12), ("Interleaved1F1B", 2, 4, 8),
    ("Interleaved1F1B", 4, 1, 4),
])
def test_properties_and_no_deadlock(name, P, v, m):
    order = build_schedule(name, P, m, n_virtual=v, dp_allreduce=True)
    S = P * v
    validate_schedule(order, S, m, stage_to_rank_loop(P))
    for r, acts in order.items():
        assert sum(a.type == ActionType.F for a in acts) == m * v
        assert sum(a.type == ActionType.B for a in acts) == m * v
        assert sum(a.type == ActionType.REDUCE_GRAD for a in acts) == v
        assert sum(a.type == ActionType.ALLREDUCE for a in acts) == v
        assert sum(a.type == ActionType.OPTIM for a in acts) == v
        # REDUCE_GRAD sits right behind the stage's last backward (plus its send)
        for i, a in enumerate(acts):
            if a.type == ActionType.REDUCE_GRAD:
                prev = [x for x in acts[:i] if x.type == ActionType.B and x.stage == a.stage]
                assert len(prev) == m
    res = simulate(order, S, comm_latency=0.0)
    assert res.makespan > 0
    theory = theoretical_bubble(P, m, v)
    # with F=1, B=2 and free communication the simulated bubble can not beat the bound by more than rounding
    assert res.bubble_max >= theory - 1e-9 or name == "Interleaved1F1B"
    assert res.bubble_max < 1.0


def test_one_f_one_b_matches_theory_exactly():
    for P, m in [(2, 4), (4, 8), (8, 32)]:
        order = build_schedule("1F1B", P, m)
        res = simulate(order, P)
        assert abs(res.bubble_max - theoretical_bubble(P, m)) < 1e-9
        order = build_schedule("GPipe", P, m)
        res = simulate(order, P)
        assert abs(res.bubble_max - theoretical_bubble(P, m)) < 1e-9


def test_interleaved_beats_1f1b_when_m_multiple_of_p():
    P, m = 8, 32
    b1 = simulate(build_schedule("1F1B", P, m), P).bubble_max
    b2 = simulate(build_schedule("Interleaved1F1B", P, m, n_virtual=2), 2 * P, f_cost=0.5, b_cost=1.0).bubble_max
    assert b2 < b1
    assert abs(theoretical_bubble(8, 32, 2) - 0.0986) < 1e-3
    assert abs(theoretical_bubble(4, 8) - 0.2727) < 1e-3
    assert abs(theoretical_bubble(8, 32) - 0.1795) < 1e-3


def test_memory_profile_1f1b_vs_gpipe():
    """In-flight micro-batches on stage 0: GPipe holds m, 1F1B at most P."""
    def peak_inflight(order, rank):
        live = peak = 0
        for a in order[rank]:
            if a.type == ActionType.F:
                live += 1
            elif a.type == ActionType.B:
                live -= 1
            peak = max(peak, live)
        return peak
    assert peak_inflight(build_schedule("GPipe", 4, 16), 0) == 16
    assert peak_inflight(build_schedule("1F1B", 4, 16), 0) == 4
    assert peak_inflight(build_schedule("1F1B", 4, 16), 3) == 1


def test_constraints():
    with pytest.raises(ValueError):
        build_schedule("1F1B", 4, 3)           # m >= stages (torch schedules.py:828)
    with pytest.raises(ValueError):
        build_schedule("Interleaved1F1B", 4, 9, n_virtual=2)  # m % rounds
    with pytest.raises(ValueError):
        build_schedule("Nope", 2, 4)


def test_deadlock_detection():
    order = build_schedule("GPipe", 2, 2, with_optimizer=False)
    bad = {0: list(order[0]), 1: list(order[1])}
    # make rank 1 wait for a micro-batch rank 0 never sends first: swap rank 0's forwards behind its backwards
    f = [a for a in bad[0] if a.type in (ActionType.F, ActionType.SEND_F)]
    rest = [a for a in bad[0] if a.type not in (ActionType.F, ActionType.SEND_F)]
    bad[0] = rest + f
    with pytest.raises(DeadlockError):
        simulate(bad, 2)


def test_matches_torch_generator_when_available():
    """Cross-check the compute order + comm lowering against torch.distributed.pipelining (test oracle only)."""
    sched = pytest.importorskip("torch.distributed.pipelining.schedules")
    for P, v, m in [(2, 2, 4), (4, 2, 4), (4, 2, 8), (8, 2, 32), (4, 3, 12)]:
        mine = interleaved_order(P, v, m)
        try:
            rounds = max(1, m // P)
            per_round = m // rounds
            for r in range(P):
                warm = sched._get_warmup_ops(r, v, per_round, P, m, multiply_factor=2)
                total = m * v

                def fsi(step, r=r):
                    return ((step // per_round) % v) * P + r

                def bsi(step, r=r, warm=warm):
                    return (v - 1 - ((step - warm) // per_round) % v) * P + r

                ops = sched._get_1f1b_rank_ops(v, P, warm, total - warm, warm, r, fsi, bsi)
                theirs = [str(a) for a in ops if a is not None]
                assert [str(a) for a in mine[r]] == theirs, (P, v, m, r)
        except (AttributeError, TypeError):
            pytest.skip("torch internals changed")


# ---- split backward (I / W) ------------------------------------------------------------------------------------
@pytest.mark.parametrize("p,m", [(2, 4), (4, 4), (4, 8), (8, 16), (3, 7)])
def test_zero_bubble_schedule_is_valid_and_beats_1f1b(p, m):
    from ppb200.parallel.schedule import zero_bubble_order, merge_bw
    raw = zero_bubble_order(p, m)
    for r, acts in raw.items():
        assert sum(a.type == T.F for a in acts) == m
        assert sum(a.type == T.I for a in acts) == m
        assert sum(a.type == T.W for a in acts) == m
        live = peak = wait = wpeak = 0
        for a in acts:   # the 1F1B bound on full activation stashes (F until I) is kept; deferred W's are capped at p
            live += (a.type == T.F) - (a.type == T.I)
            wait += (a.type == T.I) - (a.type == T.W)
            peak, wpeak = max(peak, live), max(wpeak, wait)
        assert peak <= p - r and wpeak <= p
    order = build_schedule("ZeroBubble1F1B", p, m, with_optimizer=True)     # validates internally
    zb = simulate(order, p, f_cost=1.0, b_cost=2.0)
    ref = simulate(build_schedule("1F1B", p, m), p, f_cost=1.0, b_cost=2.0)
    assert zb.makespan <= ref.makespan
    if m >= p:
        assert zb.makespan == 3 * m + (p - 1)        # ZB-H1: bubble (p-1)(F + I - W) with unit costs
        assert ref.makespan == 3 * (m + p - 1)
    # adjacent I,W pairs were fused back into B by merge_bw; nothing else changed
    merged = merge_bw(raw)
    for r in raw:
        n_b = sum(a.type == T.B for a in merged[r])
        assert sum(a.type == T.I for a in merged[r]) == m - n_b == sum(a.type == T.W for a in merged[r])


def test_split_and_merge_bw_roundtrip():
    from ppb200.parallel.schedule import split_bw, merge_bw, one_f_one_b_order
    base = one_f_one_b_order(4, 8)
    sp = split_bw(base)
    assert all(a.type != T.B for acts in sp.values() for a in acts)
    assert merge_bw(sp) == base


def test_validator_rejects_w_before_i():
    from ppb200.parallel.schedule import validate_schedule
    bad = {0: [Action(0, T.F, 0), Action(0, T.W, 0), Action(0, T.I, 0)]}
    with pytest.raises(ValueError):
        validate_schedule(bad, 1, 1)
    bad = {0: [Action(0, T.F, 0), Action(0, T.I, 0)]}
    with pytest.raises(ValueError):
        validate_schedule(bad, 1, 1)


def test_concept_figures_and_ascii(tmp_path):
    from ppb200.utils import diagram
    paths = diagram.write_concept_figures(str(tmp_path))
    assert len(paths) == 5
    for p in paths:
        txt = open(p).read()
        assert txt.startswith("<svg") and txt.rstrip().endswith("</svg>")
    tl, span, bub = diagram.timeline_of("1F1B", 4, 8)
    art = diagram.ascii_timeline(tl, span)
    assert art.count("\n") == 3 and art.splitlines()[3].startswith("rank3: ...0aa1")
    assert abs(max(bub.values()) - 3 / 11) < 1e-9


def test_overlap_f_b_text_form_and_passes():
    from ppb200.parallel.schedule import overlap_f_b, expand_overlap, one_f_one_b_order
    a = Action.from_str("(2F1;2B0)OVERLAP_F_B")
    assert a.type == T.OVERLAP_F_B and [str(x) for x in a.sub] == ["2F1", "2B0"] and str(a) == "(2F1;2B0)OVERLAP_F_B"
    assert parse_schedule("rank0: 0F0 (0F1;0B0)OVERLAP_F_B 0B1")[0][1] == a.__class__(0, T.OVERLAP_F_B, 1, (Action(0, T.F, 1), Action(0, T.B, 0)))
    base = one_f_one_b_order(4, 8)
    grouped = overlap_f_b(base)
    assert any(x.type == T.OVERLAP_F_B for x in grouped[0]) and expand_overlap(grouped) == base
    with pytest.raises(ValueError):
        Action.from_str("(2F1)OVERLAP_F_B")


def test_autotuner_predictions_track_the_measured_steps():
    """Simulator + measured per-action costs reproduce the measured B200 steps of the flagship config (profiles/) and rank
    the candidates consistently."""
    from ppb200.models import preset
    from ppb200.utils.autotune import UnitCosts, predict, recommend
    a = preset("ref-gpt2-medium")
    for pp, sched, measured in [(2, "1F1B", 428.8), (4, "1F1B", 260.4), (8, "1F1B", 155.9), (8, "ZeroBubble1F1B", 146.2)]:
        p = predict(a, pp, sched, 32, 128, 1024, costs=UnitCosts())
        assert abs(p["ms"] / measured - 1.0) < 0.08, (pp, sched, p["ms"], measured)
        assert 0.0 <= p["idle_critical"] <= p["idle_max"] < 1.0
    rows = recommend(a, 8, 128, 1024)
    assert rows and rows == sorted(rows, key=lambda r: r["ms"])
    best_1f1b = min(r["ms"] for r in rows if r["schedule"] == "1F1B")
    best_zb = min(r["ms"] for r in rows if r["schedule"] == "ZeroBubble1F1B")
    assert best_zb < best_1f1b                          # as measured: 146 vs 156 ms
    assert all((128 // r["n_microbatches"]) * 1024 >= 4096 for r in rows)
    llama = UnitCosts.from_model(preset("llama-7b"), 2048)
    assert llama.layer_f > 5 * UnitCosts().layer_f      # a 7B layer is far more expensive than a GPT-2-medium one


@pytest.mark.parametrize("p,v,m", [(2, 2, 4), (4, 2, 8), (4, 3, 8), (8, 2, 32), (2, 3, 6)])
def test_zero_bubble_interleaved_is_valid_and_beats_both_parents(p, v, m):
    from ppb200.parallel.schedule import zero_bubble_interleaved_order
    raw = zero_bubble_interleaved_order(p, v, m)
    for r, acts in raw.items():
        assert {a.stage % p for a in acts} == {r}                                   # loop placement
        for typ in (T.F, T.I, T.W):
            assert sum(a.type == typ for a in acts) == m * v
    order = build_schedule("ZeroBubbleInterleaved", p, m, n_virtual=v)              # validates internally
    kw = dict(f_cost=1.0 / v, b_cost=2.0 / v)
    zbi = simulate(order, p * v, **kw).makespan
    inter = simulate(build_schedule("Interleaved1F1B", p, m, n_virtual=v), p * v, **kw).makespan
    zb = simulate(build_schedule("ZeroBubble1F1B", p, m), p).makespan
    assert zbi < inter and zbi <= zb
    assert zbi <= 3 * m + (p - 1) / v + 1e-9 + 1.0 / v          # bubble of ZB-H1 divided by the number of chunks (+ one slot)


def test_eager_1f1b_extra_warmup_is_valid_and_keeps_more_in_flight():
    """`extra_warmup` forwards beyond torch's warm-up count: same actions, still deadlock-free, one more micro-batch in
    flight on every rank but the last; 0 reproduces the torch order exactly."""
    from ppb200.parallel.schedule import build_schedule, simulate
    from ppb200.parallel.schedule.generate import one_f_one_b_order
    from ppb200.parallel.schedule.ir import ActionType as T
    assert one_f_one_b_order(4, 8, 0) == one_f_one_b_order(4, 8)
    base, eager = build_schedule("1F1B", 4, 8), build_schedule("1F1B", 4, 8, extra_warmup=1)
    for r in range(4):
        assert sorted(map(str, base[r])) == sorted(map(str, eager[r]))

        def peak(order):
            live = best = 0
            for a in order:
                if a.type == T.F:
                    live += 1
                    best = max(best, live)
                elif a.type == T.B:
                    live -= 1
            return best
        assert peak(eager[r]) == peak(base[r]) + (1 if r < 3 else 0)
    assert simulate(eager, 4).makespan <= simulate(base, 4).makespan


def test_simulator_predicts_the_round2_sweep_out_of_sample():
    """benchmarks/predict_sweep_r2.py --check: per-token cost factors calibrated on the measured 2-rank cells, every
    measured 4- and 8-rank cell (GPipe / 1F1B / Interleaved x micro-batch counts, B200) predicted within 10 %."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "benchmarks", "predict_sweep_r2.py"), "--check"],
                       capture_output=True, text=True, timeout=600, cwd=root)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    assert "out of sample" in r.stdout
