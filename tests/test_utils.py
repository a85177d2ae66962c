"""CPU unit tests: partitioner, mesh layout, analysis formulas on the reference's published table, checkpoint
shards, config/CLI, trace/bubble helpers, dropout reference, launcher error contract."""
import json
import os
import re

import pytest
import torch

from ppb200.models import ModelArgs, Transformer, preset
from ppb200.ops import reference as R
from ppb200.parallel import partition as Pn
from ppb200.parallel.mesh import coords_of, group_ranks, rank_of
from ppb200.utils import analysis, checkpoint, trace
from ppb200.utils.config import RunConfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ---- partitioner (helper:60-94) -------------------------------------------------------------------------------
def test_reference_ranges_remainder_goes_to_last_stage():
    assert Pn.reference_ranges(12, 4) == [(0, 3), (3, 6), (6, 9), (9, 12)]
    assert Pn.reference_ranges(10, 4) == [(0, 2), (2, 4), (4, 6), (6, 10)]        # helper:74-75
    assert Pn.reference_ranges(4, 8)[-1] == (0, 4) and Pn.reference_ranges(4, 8)[0] == (0, 0)  # quirk 8


def test_balanced_ranges_cover_all_layers_and_relieve_the_head_stage():
    a = preset("ref-gpt2-medium")
    for P in (1, 2, 3, 4, 6, 8):
        rg = Pn.balanced_ranges(24, P, a, 1024)
        assert rg[0][0] == 0 and rg[-1][1] == 24
        assert all(rg[i][1] == rg[i + 1][0] for i in range(P - 1))
        assert all(e > s for s, e in rg)
        assert (rg[-1][1] - rg[-1][0]) <= max(e - s for s, e in rg)   # the head stage never has the most layers


def test_prune_and_direct_build_give_the_same_stage_keys():
    args = ModelArgs(dim=32, n_layers=5, n_heads=4, vocab_size=50, ffn_dim=64)
    full_keys = set(Transformer(args).state_dict())
    union = set()
    for s in range(3):
        pruned = Pn.prune_to_stage(Transformer(args), s, 3)
        built = Pn.build_stage_model(args, s, 3, device=torch.device("cpu"), seed=0)
        assert set(pruned.state_dict()) == set(built.state_dict())
        assert all(p.device.type == "cpu" for p in built.parameters())
        assert not (union & set(pruned.state_dict()))
        union |= set(pruned.state_dict())
    assert union == full_keys
    # global layer indices are preserved
    assert any(k.startswith("layers.4.") for k in Pn.prune_to_stage(Transformer(args), 2, 3).state_dict())


def test_fine_partition_state_dicts_tile_the_full_model():
    import copy
    args = ModelArgs(dim=32, n_layers=5, n_heads=4, vocab_size=50, ffn_dim=64, dropout=0.0)
    full = Transformer(args)
    keys = set(full.state_dict())
    seen, cut_inside = set(), False
    for s in range(4):
        parts = Pn.stage_layer_parts(5, 4, s, "fine", args, 16)
        cut_inside |= any(p != "full" for _, p in parts)
        sd = Pn.prune_to_stage(copy.deepcopy(full), s, 4, rule="fine", seq_len=16).state_dict()
        assert not (seen & set(sd))                       # every parameter lives on exactly one stage
        seen |= set(sd)
        built = Pn.build_stage_model(args, s, 4, rule="fine", seq_len=16, device=torch.device("cpu"), seed=0).state_dict()
        assert set(built) == set(sd)
    assert seen == keys and cut_inside
    units = Pn.fine_unit_ranges(24, 8, ModelArgs(dim=1024, n_layers=24, n_heads=16, vocab_size=50257), 1024)
    assert units[0][0] == 0 and units[-1][1] == 72 and all(a[1] == b[0] for a, b in zip(units, units[1:]))


def test_interleaved_fallback_rule_matches_helper():
    # helper:181-185: two chunks per rank iff n_layers % (2 * world) == 0
    for n_layers, world, want in [(4, 2, 2), (8, 4, 2), (4, 4, 1), (12, 4, 1), (12, 2, 2)]:
        v = 2 if n_layers % (world * 2) == 0 else 1
        assert v == want


# ---- mesh ---------------------------------------------------------------------------------------------------------
def test_mesh_rank_layout_roundtrip():
    dp, pp, tp = 2, 4, 2
    seen = set()
    for d in range(dp):
        for p in range(pp):
            for t in range(tp):
                r = rank_of(d, p, t, pp, tp)
                assert coords_of(r, pp, tp) == (d, p, t)
                seen.add(r)
    assert seen == set(range(16))
    ppg, dpg, tpg = group_ranks(dp, pp, tp)
    assert len(ppg) == dp * tp and all(len(g) == pp for g in ppg)
    assert len(dpg) == pp * tp and all(len(g) == dp for g in dpg)
    assert tpg[0] == [0, 1]
    assert sorted(r for g in ppg for r in g) == list(range(16))


# ---- analysis on the reference's published numbers ---------------------------------------
def _published_df():
    import pandas as pd
    rows = []
    for line in open(os.path.join(ROOT, "BASELINE.md")):
        m = re.match(r"^\|\s*(\d+)\s*\|\s*(\d+)\s*\|\s*(\d+)\s*\|\s*(GPipe|1F1B|Interleaved1F1B)\s*\|\s*(\d)\s*\|\s*([\d.]+)\s*\|", line)
        if m:
            rows.append({"n_layers": int(m.group(1)), "n_heads": int(m.group(2)), "num_processes": int(m.group(3)),
                         "schedule": m.group(4), "throughput": float(m.group(6))})
    return pd.DataFrame(rows[:54])  # section 2 of BASELINE.md (later sections re-quote a few rows)


def test_speedup_and_efficiency_reproduce_the_notebook_table():
    df = _published_df()
    assert len(df) == 54
    sp = analysis.compute_speedup_and_efficiency(df)
    assert len(sp) == 36
    row = sp[(sp.n_layers == 4) & (sp.n_heads == 4) & (sp.num_processes == 2) & (sp.schedule == "1F1B")].iloc[0]
    assert abs(row.speedup - 1.0265) < 1e-3 and abs(row.efficiency - 51.32) < 0.05      # nb:802
    row = sp[(sp.n_layers == 8) & (sp.n_heads == 4) & (sp.num_processes == 4) & (sp.schedule == "Interleaved1F1B")].iloc[0]
    assert abs(row.speedup - 1.1247) < 1e-3 and abs(row.efficiency - 28.12) < 0.05      # nb:817
    assert abs(sp.speedup.max() - 1.171) < 2e-3 and abs(sp.speedup.min() - 0.913) < 2e-3
    ratio = analysis.scaling_ratio(df)
    assert len(ratio) == 27 and abs(ratio.ratio_4_over_2.mean() - 1.114) < 5e-3
    piv = analysis.throughput_pivot(df)
    assert piv.shape == (9, 6)


def test_plots_render_headless(tmp_path):
    df = _published_df()
    sp = analysis.compute_speedup_and_efficiency(df)
    pa = analysis.plot_speedup_and_efficiency(sp, str(tmp_path / "a.svg"))
    pb = analysis.plot_throughput_grid(df, str(tmp_path / "b.svg"))
    assert os.path.getsize(pa) > 1000 and os.path.getsize(pb) > 1000
    env = analysis.environment_report()
    assert "torch" in env and "cuda_available" in env


# ---- checkpoint shards ----------------------------------------------------------------------------------------------
def test_checkpoint_shards_merge_to_full_state_dict_and_reshard(tmp_path):
    from ppb200.api import manual_model_split
    args = ModelArgs(dim=32, n_layers=4, n_heads=4, vocab_size=50, ffn_dim=64, dropout=0.0)
    torch.manual_seed(0)
    full = Transformer(args)
    want = {k: v.clone() for k, v in full.state_dict().items()}
    import copy
    stages = [manual_model_split(copy.deepcopy(full), s, 2, torch.device("cpu")).materialize() for s in range(2)]
    for st in stages:
        st.exp_avg.fill_(0.5)
        st.step_count = 7
        checkpoint.save_stage(st, str(tmp_path))
    merged = checkpoint.merge_shards(str(tmp_path))
    assert set(merged) == set(want)
    assert all(torch.equal(merged[k], want[k]) for k in want)
    # the merged dict loads into the unmodified reference layout and into a different stage count
    Transformer(args).load_state_dict(merged)
    st4 = manual_model_split(Transformer(args), 3, 4, torch.device("cpu")).materialize()
    step = checkpoint.load_stage(st4, str(tmp_path))
    assert step == 7
    for k, v in st4.state_dict().items():
        assert torch.equal(v, want[k])
    assert float(st4.exp_avg[: st4.n_total].max()) == 0.5


def test_checkpoint_roundtrip_across_tensor_parallel_degrees(tmp_path):
    """(pp, tp) = (2, 2) shards -> merged reference-layout checkpoint -> (1, 1) and back into a (2, 2) rank, parameters
    and optimizer moments (Llama-style block: packed QKV and packed gate|up splits)."""
    import copy
    from ppb200.models.tp import shard_model_for_tp
    from ppb200.parallel.partition import prune_to_stage
    from ppb200.parallel.stage import PipelineStage
    args = ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=50, ffn_dim=48, dropout=0.0, causal=True, cross_attn=False,
                     activation="swiglu", norm_first=True, norm="rmsnorm", rope=True)
    torch.manual_seed(1)
    full = Transformer(args)
    want = {k: v.clone() for k, v in full.state_dict().items()}
    group = object()   # stands in for the tp process group: no collective runs in this test

    def make(pp_idx, tp_idx):
        m = copy.deepcopy(full)
        prune_to_stage(m, pp_idx, 2)
        shard_model_for_tp(m, tp_idx, 2)
        return PipelineStage(m, pp_idx, 2, torch.device("cpu"), tp_group=group).materialize()

    for pp_idx in range(2):
        for tp_idx in range(2):
            st = make(pp_idx, tp_idx)
            st.exp_avg.copy_(st.master * 2.0)          # moments with parameter-shaped structure
            st.step_count = 3
            path = checkpoint.save_stage(st, str(tmp_path))
            assert path.endswith(f"stage_{pp_idx:03d}_tp{tp_idx}.pt")
    merged = checkpoint.merge_shards(str(tmp_path))
    assert set(merged) == set(want) and all(torch.equal(merged[k], want[k]) for k in want)
    from ppb200.api import manual_model_split
    one = manual_model_split(Transformer(args), 0, 1, torch.device("cpu")).materialize()
    assert checkpoint.load_stage(one, str(tmp_path)) == 3
    assert all(torch.equal(v, want[k]) for k, v in one.state_dict().items())
    assert torch.allclose(one.exp_avg[: one.n_total], one.master[: one.n_total] * 2.0)
    again = make(1, 1)
    again.master.zero_()
    checkpoint.load_stage(again, str(tmp_path))
    ref = make(1, 1)
    assert all(torch.equal(a, b) for a, b in zip(again.state_dict().values(), ref.state_dict().values()))
    assert torch.allclose(again.exp_avg[: again.n_total], ref.master[: ref.n_total] * 2.0)


# ---- config ---------------------------------------------------------------------------------------------------------
def test_run_config_cli_and_presets():
    cfg = RunConfig.from_args(["--model", "ref-gpt2-medium", "--schedule", "Interleaved1F1B", "--pp", "8",
                               "--virtual-stages", "2", "--n-microbatches", "32", "--optimizer", "false"])
    assert cfg.pp == 8 and cfg.virtual_stages == 2 and cfg.n_microbatches == 32 and cfg.optimizer is False
    a = cfg.model_args()
    assert (a.dim, a.n_layers, a.n_heads, a.vocab_size, a.ffn_dim) == (1024, 24, 16, 50257, 2048)
    assert cfg.optimizer_dict() is None
    assert RunConfig().optimizer_dict()["lr"] == 1e-4
    d = ModelArgs()
    assert (d.dim, d.n_layers, d.n_heads, d.vocab_size) == (768, 8, 8, 10000)      # helper:23-28
    assert d.params_per_layer() == 7_877_888 and d.num_params() == 78_394_640 + 0  # SURVEY.md 2.3 (8 layers: 78.4 M)
    with pytest.raises(ValueError):
        ModelArgs(dim=100, n_heads=8)


# ---- trace helpers ----------------------------------------------------------------------------------------------------
def test_bubble_from_events_and_chrome_trace(tmp_path):
    evs = [{"name": "0F0", "ph": "X", "pid": 0, "tid": 0, "ts": 0.0, "dur": 10.0},
           {"name": "0B0", "ph": "X", "pid": 0, "tid": 0, "ts": 30.0, "dur": 20.0},
           {"name": "1F0", "ph": "X", "pid": 1, "tid": 1, "ts": 10.0, "dur": 10.0},
           {"name": "1B0", "ph": "X", "pid": 1, "tid": 1, "ts": 20.0, "dur": 10.0}]
    b = trace.bubble_from_events(evs)
    assert abs(b["rank0"] - 0.4) < 1e-9 and abs(b["rank1"] - 0.6) < 1e-9 and b["max"] == b["rank1"]
    trace.write_chrome_trace(evs, str(tmp_path / "t.json"))
    assert len(json.load(open(tmp_path / "t.json"))["traceEvents"]) == 4


# ---- dropout reference ------------------------------------------------------------------------------------------------
def test_dropout_reference_rate_and_determinism():
    x = torch.ones(200_000)
    y = R.dropout(x, 0.1, seed=42)
    kept = (y != 0).float().mean().item()
    assert abs(kept - 0.9) < 5e-3
    assert torch.equal(y, R.dropout(x, 0.1, seed=42))
    assert not torch.equal(y, R.dropout(x, 0.1, seed=43))
    assert abs(y.mean().item() - 1.0) < 1e-2
    # masks are addressed by flat index: a shifted window reproduces the same decisions
    m_all = R.dropout_keep_mask(7, 0, 1000, 0.3)
    assert torch.equal(m_all[100:300], R.dropout_keep_mask(7, 100, 200, 0.3))
    assert torch.equal(R.dropout_keep_mask(7, 2**33 + 10, 64, 0.3), R.dropout_keep_mask(7, 2**33 + 10, 64, 0.3))


# ---- launcher error contract (helper:231-235, nb:328-331) ---------------------------------------------------------------
def _boom(rank, world, q, master_port=None):
    if rank == 1:
        q.put({"error": "boom"})
    else:
        import time
        time.sleep(30)


def _silent(rank, world, q, master_port=None):
    return


def test_launcher_reports_errors_and_kills_stragglers():
    from ppb200.parallel.launch import launch_workers
    import time
    t0 = time.time()
    res = launch_workers(_boom, 2, (), timeout=60)
    assert res == {"error": "boom"} and time.time() - t0 < 25
    assert launch_workers(_silent, 2, (), timeout=60) == {"error": "No results returned"}


def test_guarded_allocations_catch_out_of_bounds_writes_and_stale_reads():
    """The in-repo memory checker (utils/debug.py), exercised on CPU tensors: guard bands + NaN poison."""
    import torch
    from ppb200.utils import debug
    with debug.guarded_allocations(device_type="cpu") as rep:
        a = torch.empty(16, 8, dtype=torch.float32, device="cpu")
        assert torch.isnan(a).all()                       # unwritten memory reads as NaN
        z = torch.zeros(5, dtype=torch.float32, device="cpu")
        assert (z == 0).all()
        b = torch.empty_like(a)
        assert b.shape == a.shape and torch.isnan(b).all()
        a.fill_(1.0)
        assert rep.check("clean") == 0
        c = torch.empty(4, dtype=torch.float32, device="cpu")
        # simulate a kernel writing 3 floats past the end of c
        c.as_strided((7,), (1,))[4:].fill_(2.0)
        assert rep.check("oob") == 1
        assert "test_utils.py" in rep.violations[0] and "after" in rep.violations[0]
    assert torch.empty(2).shape == (2,)                   # patches are removed on exit
    assert torch.empty is not None and not rep.entries


def test_determinism_check_reports_mismatches():
    import torch
    from ppb200.utils import debug
    x = torch.arange(64, dtype=torch.float32)
    assert debug.determinism_check(lambda: (x * 2, None), reps=5, device="cpu")["mismatching"] == 0
    state = {"n": 0}

    def flaky():
        state["n"] += 1
        return (x + (1.0 if state["n"] == 3 else 0.0),)
    res = debug.determinism_check(flaky, reps=5, device="cpu")
    assert res["mismatching"] == 1 and res["worst_abs"] == 1.0


def test_token_file_data_gives_next_token_targets(tmp_path):
    import numpy as np
    from ppb200.train import TokenData
    toks = (np.arange(5000) % 97).astype(np.uint16)
    path = tmp_path / "tokens.bin"
    toks.tofile(path)
    cfg = RunConfig(global_batch=8, seq_len=16, data_path=str(path), seed=1, dp=2)
    d0, d1 = TokenData(cfg, 97, 0, 2), TokenData(cfg, 97, 1, 2)
    x, y = d0.batch(5)
    assert x.shape == (4, 16) and y.shape == (4, 16) and x.dtype == torch.int64
    assert torch.equal(x[:, 1:], y[:, :-1])                      # targets are the inputs shifted by one token
    assert torch.equal(d0.batch(5)[0], x)                        # deterministic per (step, dp rank)
    assert not torch.equal(d1.batch(5)[0], x) and not torch.equal(d0.batch(6)[0], x)
    syn = TokenData(RunConfig(global_batch=4, seq_len=8), 50, 0, 1).batch(0)
    assert syn[0].shape == (4, 8) and int(syn[0].max()) < 50
    assert RunConfig(lr=1.0, lr_warmup_steps=4).lr_at(1) == 0.5
    c = RunConfig(lr=1.0, lr_warmup_steps=0, lr_decay_steps=10, lr_min_ratio=0.1)
    assert c.lr_at(0) == 1.0 and abs(c.lr_at(10) - 0.1) < 1e-9 and abs(c.lr_at(5) - 0.55) < 1e-9


def test_partition_units_parts_and_auto_rule():
    """Unit-granular cuts: part names, which parameters a part keeps, measured costs steering the cut positions, and the
    'auto' rule (sub-layer cuts only when they lower the most expensive stage by more than 2 %)."""
    import copy
    args = ModelArgs(dim=32, n_layers=6, n_heads=4, vocab_size=50, ffn_dim=64, dropout=0.0)
    assert Pn.layer_units(args) == ("self", "cross", "ffn")
    assert Pn.layer_units(ModelArgs(dim=32, n_layers=2, n_heads=4, cross_attn=False, norm_first=True, causal=True)) == ("attn", "ffn")
    assert Pn.part_name(0, 3, ("self", "cross", "ffn")) == "full" and Pn.part_name(0, 2, ("self", "cross", "ffn")) == "attn"
    assert Pn.part_name(1, 3, ("self", "cross", "ffn")) == "cross+ffn" and Pn.part_units("cross+ffn", ("self", "cross", "ffn")) == ("cross", "ffn")
    full = Transformer(args)
    layer = copy.deepcopy(full.layers["0"])
    Pn._apply_part(layer, "cross+ffn")
    kept = {n.split(".")[0] for n, _ in layer.named_parameters()}
    assert kept == {"multihead_attn", "norm2", "linear1", "linear2", "norm3"}
    try:
        Pn.set_measured_costs(args, 16, {"self": 1.0, "cross": 1.0, "ffn": 1.0, "head": 2.0, "embed": 0.0})
        assert Pn.unit_costs(args, 16)["source"] == "measured"
        costs = Pn._stage_costs(6, 4, "fine", args, 16)
        assert max(costs) == 5.0 and abs(sum(costs) - 20.0) < 1e-9          # 18 units + head 2 over 4 stages: 5 each
        assert Pn.resolve_rule("auto", 6, 4, args, 16) == "fine"            # whole layers: max 6 (two layers) vs 5 with cuts
        Pn.set_measured_costs(args, 16, {"self": 1.0, "cross": 1.0, "ffn": 1.0, "head": 0.0, "embed": 0.0})
        assert Pn.resolve_rule("auto", 6, 3, args, 16) == "balanced"        # 2 layers per stage is already perfect
        assert Pn.resolve_rule("auto", 6, 4, args, 16, tp=2) == "balanced"  # tensor-parallel stages keep whole layers
    finally:
        Pn.clear_measured_costs()
    assert Pn.unit_costs(args, 16)["source"] == "model"


def test_deadline_lets_a_fast_block_finish_and_fires_on_a_slow_one():
    """bench.py bounds its appended fused-kernel check with this: exactly one side (block / handler) wins."""
    import time
    from ppb200.utils.deadline import Deadline
    fired = []
    with Deadline(5.0, lambda: fired.append("late")) as dl:
        x = 1 + 1
    assert dl.completed and x == 2
    time.sleep(0.05)
    assert fired == []
    with Deadline(0.05, lambda: fired.append("timeout")) as dl:
        time.sleep(0.4)
    assert not dl.completed and fired == ["timeout"]
    try:   # an exception inside the block disarms the watchdog and propagates
        with Deadline(0.05, lambda: fired.append("after-exception")) as dl:
            raise ValueError("boom")
    except ValueError:
        pass
    time.sleep(0.15)
    assert fired == ["timeout"] and dl.completed


def test_min_max_partition_is_optimal_against_brute_force():
    """The stage cutter (parallel/partition.py: _min_max_partition) against exhaustive search on random unit costs,
    with the embedding charged to the first group and the head to the last (hypothesis-driven)."""
    import itertools
    from hypothesis import given, settings, strategies as st
    from ppb200.parallel.partition import _min_max_partition

    @settings(max_examples=60, deadline=None)
    @given(st.lists(st.floats(0.05, 3.0), min_size=2, max_size=9), st.integers(1, 5), st.floats(0.0, 2.5), st.floats(0.0, 0.5))
    def check(costs, S, tail, head):
        S = min(S, len(costs))
        got = _min_max_partition(costs, S, tail, head)
        assert got[0][0] == 0 and got[-1][1] == len(costs) and all(a < b for a, b in got)
        assert all(got[i][1] == got[i + 1][0] for i in range(S - 1))

        def worst(cuts):
            b = [0, *cuts, len(costs)]
            loads = [sum(costs[b[i]:b[i + 1]]) for i in range(S)]
            loads[0] += head
            loads[-1] += tail
            return max(loads)

        best = min(worst(c) for c in itertools.combinations(range(1, len(costs)), S - 1))
        mine = worst([b for _, b in got[:-1]])
        assert mine <= best * (1 + 1e-9) + 1e-12, (costs, S, tail, head, got, mine, best)

    check()


def _dead_peer(rank, world, q, master_port=None):
    """Rank 1 joins the group and then never sends: rank 0's receive must give up after the configured time."""
    import datetime
    import os
    import time
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(master_port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world, timeout=datetime.timedelta(seconds=3))
    if rank == 1:
        time.sleep(60)
        return
    try:
        dist.recv(torch.zeros(4), src=1)
        q.put({"error": "the receive returned although nothing was sent"})
    except Exception as e:  # noqa: BLE001  (same error-dict contract as the workers, helper:231-235)
        q.put({"error": f"peer lost: {type(e).__name__}"})


def test_collective_timeout_turns_a_dead_peer_into_an_error_dict():
    """the reference blocks forever when a rank dies (no timeout anywhere).  With `collective_timeout_s`
    (RunConfig / worker_process / `ppb200.train --collective-timeout-s`) the survivor raises, the worker reports the
    error dict and the launcher tears the rest down."""
    import time
    from ppb200.parallel.launch import launch_workers
    from ppb200.utils.config import RunConfig
    assert RunConfig.from_args(["--collective-timeout-s", "7.5"]).collective_timeout_s == 7.5
    t0 = time.time()
    res = launch_workers(_dead_peer, 2, (), timeout=120)
    assert "error" in res and res["error"].startswith("peer lost"), res
    assert time.time() - t0 < 45
