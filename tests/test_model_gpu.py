"""GPU integration: the bf16 CUDA stage program (all ops in libppb200.so) against the fp32 CPU program on the same
weights -- loss and per-parameter gradients (SURVEY.md section 4 'Distributed GPU integration', 1-GPU part)."""
import copy

import pytest
import torch

from ppb200.api import make_schedule, manual_model_split
from ppb200.models import ModelArgs, Transformer
from ppb200.ops import _lib

pytestmark = pytest.mark.gpu


def _run(dev, full, x, y, m, dropout_seed=3, steps=1, opt=None):
    model = copy.deepcopy(full)
    st = manual_model_split(model, 0, 1, torch.device(dev))
    st.base_seed = dropout_seed
    sched = make_schedule("1F1B", st, m, None, optimizer=opt)
    hist = []
    for _ in range(steps):
        losses = []
        sched.step(x.to(dev), target=y.to(dev), losses=losses)
        hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
    return st, hist


@pytest.mark.parametrize("heads,dropout", [(4, 0.1), (6, 0.0), (8, 0.1), (12, 0.1)])
def test_cuda_matches_cpu_reference(heads, dropout):
    torch.manual_seed(0)
    args = ModelArgs(dim=768, n_layers=2, n_heads=heads, vocab_size=1000, dropout=dropout)
    full = Transformer(args)
    x = torch.randint(0, 1000, (8, 128))
    y = torch.randint(0, 1000, (8, 128))
    st_c, loss_c = _run("cpu", full, x, y, 4)
    st_g, loss_g = _run("cuda", full, x, y, 4)
    assert abs(loss_c[0] - loss_g[0]) < 3e-2 * abs(loss_c[0]), (loss_c, loss_g)
    gc, gg = st_c.grad_views(), st_g.grad_views()
    worst = 0.0
    for n in gc:
        a, b = gc[n].float(), gg[n].float().cpu()
        rel = ((a - b).norm() / (a.norm() + 1e-6)).item()
        worst = max(worst, rel)
        assert rel < 0.12, (n, rel)   # bf16 activations end-to-end through 2 layers
    assert worst > 0


def test_training_loss_goes_down_and_library_is_used():
    torch.manual_seed(1)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
    full = Transformer(args)
    x = torch.randint(0, 512, (8, 128))
    y = torch.randint(0, 512, (8, 128))
    _lib.reset_launches()
    st, hist = _run("cuda", full, x, y, 4, steps=8, opt={"lr": 2e-3, "weight_decay": 0.0})
    assert hist[-1] < hist[0] - 0.1, hist
    assert _lib.launches() > 100
    sd = st.state_dict()
    assert set(sd) == set(full.state_dict())
    assert all(v.dtype == torch.float32 for v in sd.values())  # fp32 master weights are what gets checkpointed


def test_smoke_entry():
    import __graft_entry__ as g
    g.smoke()


def test_cuda_graph_path_matches_eager():
    """Graph-captured forward/backward (per in-flight slot) must train exactly like the eager launches (dropout 0)."""
    torch.manual_seed(2)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.0)
    full = Transformer(args)
    x = torch.randint(0, 512, (8, 128))
    y = torch.randint(0, 512, (8, 128))
    res = {}
    for use in (False, True):
        model = copy.deepcopy(full)
        st = manual_model_split(model, 0, 1, torch.device("cuda"))
        sched = make_schedule("1F1B", st, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0}, cuda_graphs=use,
                              graph_warmup_steps=1)
        hist = []
        for _ in range(5):
            losses = []
            sched.step(x.cuda(), target=y.cuda(), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        res[use] = (hist, st.master[:st.n_total].clone())
        assert sched.use_graphs == use
    assert max(abs(a - b) for a, b in zip(res[False][0], res[True][0])) < 1e-4, (res[False][0], res[True][0])
    rel = ((res[False][1] - res[True][1]).norm() / res[False][1].norm()).item()
    assert rel < 1e-5, rel


@pytest.mark.parametrize("graphs", [False, True])
def test_split_backward_schedule_matches_fused_backward(graphs):
    """ZeroBubble1F1B (backward as I + deferred W actions, eager and graph-captured) == 1F1B (fused B), dropout on:
    the same counter-based masks are regenerated whichever action consumes them.

    Tolerance: the fp32 atomics of the bias / embedding gradient reductions land in arrival order, a 1e-7-level
    difference that now and then flips the bf16 rounding of one parameter and grows to ~5e-5 over the next steps
    (measured; both orders show it against themselves).  A wrong mask or a clobbered operand is >= 1e-3."""
    torch.manual_seed(5)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
    full = Transformer(args)
    x = torch.stack([torch.randperm(512)[:128] for _ in range(8)])   # few repeated tokens per micro-batch
    y = torch.randint(0, 512, (8, 128))
    res = {}
    for name in ("1F1B", "ZeroBubble1F1B"):
        st = manual_model_split(copy.deepcopy(full), 0, 1, torch.device("cuda"))
        # one rank cannot show a bubble, so hand the runner an explicit order with two deferred W's in flight
        custom = None if name == "1F1B" else {0: "0F0 0I0 0F1 0I1 0W0 0F2 0I2 0W1 0F3 0I3 0W2 0W3".split()}
        sched = make_schedule("1F1B", st, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0}, cuda_graphs=graphs,
                              graph_warmup_steps=1, actions=custom)
        if custom is not None:
            assert [str(a) for a in sched.order][-2:] == ["0REDUCE_GRAD", "0OPTIM"] and sched.name == "custom"
        hist = []
        for _ in range(4):
            losses = []
            sched.step(x.cuda(), target=y.cuda(), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        res[name] = (hist, st.master[:st.n_total].clone(), list(st.param_names))
    assert max(abs(a - b) for a, b in zip(res["1F1B"][0], res["ZeroBubble1F1B"][0])) < 5e-4, res
    a, b = res["1F1B"][1], res["ZeroBubble1F1B"][1]
    rel = ((a - b).norm() / a.norm()).item()
    worst = sorted(((((a[o:o + torch.Size(s).numel()] - b[o:o + torch.Size(s).numel()]).norm() /
                      a[o:o + torch.Size(s).numel()].norm()).item(), n) for n, o, s in res["1F1B"][2]), reverse=True)[:6]
    assert rel < 3e-4, (rel, worst, res["1F1B"][0], res["ZeroBubble1F1B"][0])


def test_graph_slots_that_see_both_B_and_I_graphs():
    """A graph slot of a non-last stage can serve a fused backward (B) for one micro-batch and a split backward (I + W)
    for another (ZeroBubble orders merge only some micro-batches back into B).  Both captured graphs must read the
    slot's ONE static output-gradient buffer; with graphs on, losses and parameters must track the eager run."""
    torch.manual_seed(9)
    args = ModelArgs(dim=256, n_layers=4, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
    full = Transformer(args)
    x = torch.stack([torch.randperm(512)[:128] for _ in range(8)])
    y = torch.randint(0, 512, (8, 128))
    order = {0: ("0F0 1F0 1B0 0I0 0F1 1F1 1B1 0B1 0W0 0F2 1F2 1B2 0B2 0F3 1F3 1B3 0I3 0W3").split()}
    res = {}
    for graphs in (False, True):
        stages = [manual_model_split(copy.deepcopy(full), s, 2, torch.device("cuda"), rule="balanced") for s in range(2)]
        for st in stages:
            st.base_seed = 11
        sched = make_schedule("Interleaved1F1B", stages, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0},
                              cuda_graphs=graphs, graph_warmup_steps=1, actions=order)
        if graphs:
            assert sched.use_graphs and stages[0].nslots == 2
        hist = []
        for _ in range(5):
            losses = []
            sched.step(x.cuda(), target=y.cuda(), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        res[graphs] = (hist, torch.cat([st.master[:st.n_total] for st in stages]).clone())
    assert max(abs(a - b) for a, b in zip(res[False][0], res[True][0])) < 5e-4, res
    rel = ((res[False][1] - res[True][1]).norm() / res[False][1].norm()).item()
    assert rel < 3e-4, (rel, res[False][0], res[True][0])


def test_fine_partition_with_packed_boundary_cuda():
    """Sub-layer cuts on the GPU (a boundary behind a self-attention block carries the packed [x1 ; x0] tensor): three
    local virtual stages with forced cut positions against the unsplit model, graphs on, dropout on."""
    from ppb200.parallel import partition as Pn
    torch.manual_seed(10)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
    full = Transformer(args)
    x = torch.stack([torch.randperm(512)[:128] for _ in range(6)])
    y = torch.randint(0, 512, (6, 128))
    res = {}
    for split in (False, True):
        if split:
            Pn.set_measured_costs(args, 1024, {"self": 1.0, "cross": 1.0, "ffn": 1.0, "head": 0.0, "embed": 0.0})
        try:
            n = 3 if split else 1
            stages = [manual_model_split(copy.deepcopy(full), s, n, torch.device("cuda"), rule="fine" if split else "reference")
                      for s in range(n)]
        finally:
            Pn.clear_measured_costs()
        if split:   # L0.self+cross | L0.ffn + L1.self | L1.cross+ffn
            assert [getattr(l, "part", "full") for st in stages for l in st.submod.layers.values()] == ["attn", "ffn", "self", "cross+ffn"]
            assert stages[1].out_rows == 2 and stages[2].in_rows == 2
        for st in stages:
            st.base_seed = 5
        sched = make_schedule("Interleaved1F1B" if split else "1F1B", stages if split else stages[0], 3, None,
                              optimizer={"lr": 1e-3, "weight_decay": 0.0}, graph_warmup_steps=1)
        hist = []
        for _ in range(4):
            losses = []
            sched.step(x.cuda(), target=y.cuda(), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        views = {}
        for st in stages:
            for name, o, shp in st.param_names:
                views[name] = st.master[o:o + torch.Size(shp).numel()].clone()
        res[split] = (hist, views)
    assert max(abs(a - b) for a, b in zip(res[False][0], res[True][0])) < 5e-4, res
    assert set(res[False][1]) == set(res[True][1])
    # Adam turns rounding-level gradient differences of near-zero parameters (biases start at 0) into full-size steps
    # of either sign, so the parameters are compared as one vector (a wrong boundary tensor is a >= 1e-2 difference)
    ks = sorted(res[False][1])
    a, b = torch.cat([res[False][1][k].flatten() for k in ks]), torch.cat([res[True][1][k].flatten() for k in ks])
    assert ((a - b).norm() / a.norm()).item() < 1e-3


def test_measured_unit_costs_drive_the_partition():
    from ppb200.parallel import partition as Pn
    args = ModelArgs(dim=256, n_layers=6, n_heads=4, vocab_size=2048, ffn_dim=512, dropout=0.1, dtype=torch.bfloat16)
    try:
        c = Pn.calibrate(args, 2, 128, torch.device("cuda"))
        assert c["source"] == "measured" and all(c[k] > 0 for k in ("self", "cross", "ffn", "head", "embed")), c
        assert Pn.unit_costs(args, 128)["source"] == "measured"
        parts = [Pn.stage_layer_parts(6, 4, s, "fine", args, 128) for s in range(4)]
        assert sum(len(p) for p in parts) >= 6
    finally:
        Pn.clear_measured_costs()


@pytest.mark.parametrize("act,dropout", [("gelu", 0.0), ("relu", 0.1)])
def test_prenorm_gpt_block_cuda_matches_cpu(act, dropout):
    torch.manual_seed(4)
    args = ModelArgs(dim=512, n_layers=2, n_heads=8, vocab_size=1000, ffn_dim=1024, dropout=dropout, causal=True,
                     cross_attn=False, activation=act, norm_first=True)
    full = Transformer(args)
    x = torch.randint(0, 1000, (8, 128))
    y = torch.randint(0, 1000, (8, 128))
    st_c, loss_c = _run("cpu", full, x, y, 4)
    st_g, loss_g = _run("cuda", full, x, y, 4)
    assert abs(loss_c[0] - loss_g[0]) < 3e-2 * abs(loss_c[0]), (loss_c, loss_g)
    gc, gg = st_c.grad_views(), st_g.grad_views()
    for n in gc:
        a, b = gc[n].float(), gg[n].float().cpu()
        rel = ((a - b).norm() / (a.norm() + 1e-6)).item()
        assert rel < 0.12, (n, rel)


def test_llama_style_block_cuda_matches_cpu():
    """SwiGLU + RoPE + RMSNorm pre-LN block: bf16 CUDA program (ppb_swiglu_*, ppb_rope_inplace + the GEMM/attention/norm
    kernels) against the fp32 CPU program on the same weights."""
    torch.manual_seed(6)
    args = ModelArgs(dim=512, n_layers=2, n_heads=8, vocab_size=1000, ffn_dim=1024, dropout=0.0, causal=True,
                     cross_attn=False, activation="swiglu", norm_first=True, norm="rmsnorm", rope=True)
    full = Transformer(args)
    x = torch.randint(0, 1000, (8, 128))
    y = torch.randint(0, 1000, (8, 128))
    st_c, loss_c = _run("cpu", full, x, y, 4)
    st_g, loss_g = _run("cuda", full, x, y, 4)
    assert abs(loss_c[0] - loss_g[0]) < 3e-2 * abs(loss_c[0]), (loss_c, loss_g)
    gc, gg = st_c.grad_views(), st_g.grad_views()
    for n in gc:
        a, b = gc[n].float(), gg[n].float().cpu()
        rel = ((a - b).norm() / (a.norm() + 1e-6)).item()
        assert rel < 0.12, (n, rel)


def test_gradient_clipping_cuda_matches_cpu():
    """Global-norm clipping: device-side sum of squares + coefficient (ppb_sumsq_f32, ppb_clip_coef) feeding the AdamW
    kernel, against the CPU program on the same weights."""
    torch.manual_seed(8)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.0)
    full = Transformer(args)
    x = torch.randint(0, 512, (8, 128))
    y = torch.randint(0, 512, (8, 128))
    res = {}
    for dev in ("cpu", "cuda"):
        st = manual_model_split(copy.deepcopy(full), 0, 1, torch.device(dev))
        sched = make_schedule("1F1B", st, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0, "max_grad_norm": 0.5})
        hist, norms = [], []
        for _ in range(3):
            losses = []
            sched.step(x.to(dev), target=y.to(dev), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
            norms.append(float(sched.last_grad_norm_sq.sqrt()) / 4)
        res[dev] = (hist, norms)
    assert res["cpu"][1][0] > 0.5                                            # clipping is active
    for a, b in zip(res["cpu"][1], res["cuda"][1]):
        assert abs(a - b) < 0.05 * a, res
    assert res["cuda"][0][-1] < res["cuda"][0][0]
    assert abs(res["cuda"][0][-1] - res["cpu"][0][-1]) < 3e-2 * res["cpu"][0][-1], res


def test_activation_recomputation_cuda():
    """recompute=True on CUDA (eager launches): same losses and the same training trajectory as the stashing run."""
    torch.manual_seed(10)
    args = ModelArgs(dim=256, n_layers=2, n_heads=4, vocab_size=512, ffn_dim=512, dropout=0.1)
    full = Transformer(args)
    x = torch.stack([torch.randperm(512)[:128] for _ in range(8)])
    y = torch.randint(0, 512, (8, 128))
    res = {}
    for rc in (False, True):
        st = manual_model_split(copy.deepcopy(full), 0, 1, torch.device("cuda"))
        sched = make_schedule("1F1B", st, 4, None, optimizer={"lr": 1e-3, "weight_decay": 0.0}, cuda_graphs=False, recompute=rc)
        hist = []
        for _ in range(3):
            losses = []
            sched.step(x.cuda(), target=y.cuda(), losses=losses)
            hist.append(torch.stack([l.reshape(()) for l in losses]).mean().item())
        res[rc] = (hist, st.master[:st.n_total].clone())
        assert sched.use_graphs is False
    assert max(abs(a - b) for a, b in zip(res[False][0], res[True][0])) < 5e-4, res
    rel = ((res[False][1] - res[True][1]).norm() / res[False][1].norm()).item()
    assert rel < 3e-4, rel
