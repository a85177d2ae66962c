"""Model parity with the UNMODIFIED reference helper (SURVEY.md section 4 'Oracle vs. reference'): identical
state_dict key layout, identical logits and parameter gradients (fp32, dropout off), on CPU."""
import os
import sys

import pytest
import torch

from ppb200.models import ModelArgs, Transformer

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIRS = [os.path.join(ROOT, "baseline", "_ref"), "/root/reference"]


def _load_ref():
    for d in REF_DIRS:
        if os.path.exists(os.path.join(d, "LLMsDistributedTrainingHelper.py")):
            if d not in sys.path:
                sys.path.insert(0, d)
            import LLMsDistributedTrainingHelper as ref  # noqa: WPS433
            return ref
    pytest.skip("reference helper not available")


def test_state_dict_layout_matches_reference():
    ref = _load_ref()
    rm = ref.Transformer(ref.ModelArgs(n_layers=2, n_heads=4))
    mm = Transformer(ModelArgs(n_layers=2, n_heads=4))
    rsd, msd = rm.state_dict(), mm.state_dict()
    assert list(rsd.keys()) == list(msd.keys())
    for k in rsd:
        assert tuple(rsd[k].shape) == tuple(msd[k].shape), k
    mm.load_state_dict(rsd)  # the checkpoint format is shared
    assert sum(p.numel() for p in mm.parameters()) == ModelArgs(n_layers=2, n_heads=4).num_params()


@pytest.mark.parametrize("heads", [4, 8, 12])
def test_logits_and_grads_match_reference(heads):
    ref = _load_ref()
    torch.manual_seed(0)
    rm = ref.Transformer(ref.ModelArgs(dim=96 * 2, n_layers=2, n_heads=heads, vocab_size=211)).eval()  # eval: dropout off
    mm = Transformer(ModelArgs(dim=96 * 2, n_layers=2, n_heads=heads, vocab_size=211, dropout=0.0))
    mm.load_state_dict(rm.state_dict())
    x = torch.randint(0, 211, (3, 16))
    y = torch.randint(0, 211, (3, 16))
    lr = rm(x)
    lm = mm(x)
    assert torch.allclose(lr, lm, atol=2e-5, rtol=1e-4), (lr - lm).abs().max()
    ce = torch.nn.CrossEntropyLoss()
    ce(lr.reshape(-1, 211), y.reshape(-1)).backward()
    ce(lm.reshape(-1, 211), y.reshape(-1)).backward()
    rg = dict(rm.named_parameters())
    for n, p in mm.named_parameters():
        assert p.grad is not None, n
        assert torch.allclose(p.grad, rg[n].grad, atol=3e-5, rtol=1e-3), (n, (p.grad - rg[n].grad).abs().max())


def test_pruned_model_is_a_stage():
    torch.manual_seed(1)
    args = ModelArgs(dim=64, n_layers=4, n_heads=4, vocab_size=97, dropout=0.0, ffn_dim=128)
    full = Transformer(args)
    x = torch.randint(0, 97, (2, 8))
    want = full(x)
    import copy
    s0, s1 = copy.deepcopy(full), copy.deepcopy(full)
    for i in ("2", "3"):
        del s0.layers[i]
    s0.norm = None
    s0.output = None
    for i in ("0", "1"):
        del s1.layers[i]
    s1.tok_embeddings = None
    mid = s0(x)
    assert mid.shape == (2, 8, 64)
    got = s1(mid)
    assert torch.allclose(want, got, atol=1e-5)
    assert set(s0.state_dict()) | set(s1.state_dict()) == set(full.state_dict())
    assert not (set(s0.state_dict()) & set(s1.state_dict()))


def test_dropout_is_deterministic_and_active():
    args = ModelArgs(dim=64, n_layers=1, n_heads=4, vocab_size=50, ffn_dim=128)
    m = Transformer(args)
    x = torch.randint(0, 50, (2, 8))
    m.set_seed(5)
    a = m(x)
    m.set_seed(5)
    b = m(x)
    m.set_seed(6)
    c = m(x)
    assert torch.equal(a, b)
    assert not torch.equal(a, c)
    m.eval()
    d1, d2 = m(x), m(x)
    assert torch.equal(d1, d2)


@pytest.mark.parametrize("act", ["gelu", "relu"])
def test_prenorm_gpt_block_matches_torch_encoder_layer(act):
    """The pre-LN (GPT-style, causal, no cross-attention) block against nn.TransformerEncoderLayer(norm_first=True)."""
    torch.manual_seed(3)
    d, H, f, V, S = 64, 4, 128, 97, 16
    args = ModelArgs(dim=d, n_layers=2, n_heads=H, vocab_size=V, ffn_dim=f, dropout=0.0, causal=True, cross_attn=False,
                     activation=act, norm_first=True)
    m = Transformer(args)
    layers = [torch.nn.TransformerEncoderLayer(d, H, dim_feedforward=f, dropout=0.0, activation=act, batch_first=True,
                                               norm_first=True) for _ in range(2)]
    for i, tl in enumerate(layers):
        ml = m.layers[str(i)]
        tl.self_attn.in_proj_weight.data.copy_(ml.self_attn.in_proj_weight.data)
        tl.self_attn.in_proj_bias.data.copy_(ml.self_attn.in_proj_bias.data)
        tl.self_attn.out_proj.weight.data.copy_(ml.self_attn.out_proj.weight.data)
        tl.self_attn.out_proj.bias.data.copy_(ml.self_attn.out_proj.bias.data)
        tl.linear1.weight.data.copy_(ml.linear1.weight.data)
        tl.linear1.bias.data.copy_(ml.linear1.bias.data)
        tl.linear2.weight.data.copy_(ml.linear2.weight.data)
        tl.linear2.bias.data.copy_(ml.linear2.bias.data)
        ml.norm1.weight.data.uniform_(0.5, 1.5)
        ml.norm3.weight.data.uniform_(0.5, 1.5)
        tl.norm1.weight.data.copy_(ml.norm1.weight.data)
        tl.norm2.weight.data.copy_(ml.norm3.weight.data)
    x = torch.randint(0, V, (2, S))
    y = torch.randint(0, V, (2, S))
    mask = torch.nn.Transformer.generate_square_subsequent_mask(S)
    h = torch.nn.functional.embedding(x, m.tok_embeddings.weight)
    for tl in layers:
        h = tl(h, src_mask=mask, is_causal=True)
    h = torch.nn.functional.layer_norm(h, (d,), m.norm.weight, m.norm.bias)
    want = torch.nn.functional.linear(h, m.output.weight, m.output.bias)
    got = m(x)
    assert torch.allclose(want, got, atol=3e-5, rtol=1e-4), (want - got).abs().max()
    ce = torch.nn.CrossEntropyLoss()
    ce(got.reshape(-1, V), y.reshape(-1)).backward()
    gm = {n: p.grad.clone() for n, p in m.named_parameters()}
    m.zero_grad()
    ce(want.reshape(-1, V), y.reshape(-1)).backward()   # same parameters feed the torch graph for embedding/norm/head
    assert torch.allclose(gm["output.weight"], m.output.weight.grad, atol=3e-5, rtol=1e-3)
    assert torch.allclose(gm["tok_embeddings.weight"], m.tok_embeddings.weight.grad, atol=3e-5, rtol=1e-3)
    for i, tl in enumerate(layers):
        for mine, theirs in [(f"layers.{i}.self_attn.in_proj_weight", tl.self_attn.in_proj_weight),
                             (f"layers.{i}.linear1.weight", tl.linear1.weight), (f"layers.{i}.linear2.bias", tl.linear2.bias),
                             (f"layers.{i}.norm1.weight", tl.norm1.weight), (f"layers.{i}.norm3.bias", tl.norm2.bias)]:
            assert torch.allclose(gm[mine], theirs.grad, atol=3e-5, rtol=1e-3), mine


def _llama_reference(sd, args, tokens):
    """Independent plain-PyTorch (autograd) Llama-style model on the framework's state dict: RMSNorm, rotate-half RoPE,
    causal attention, SwiGLU with linear1 packed as gate | up."""
    import torch.nn.functional as Fn
    d, H = args.dim, args.n_heads
    dh = d // H
    B, S = tokens.shape

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + args.layer_norm_eps) * w

    inv = args.rope_theta ** (-torch.arange(0, dh, 2, dtype=torch.float64) / dh)
    ang = (torch.arange(S, dtype=torch.float64)[:, None] * inv[None, :]).float()
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)   # [S, dh]

    def rope(t):   # [B, H, S, dh]
        rot = torch.cat([-t[..., dh // 2:], t[..., :dh // 2]], -1)
        return t * cos + rot * sin

    x = sd["tok_embeddings.weight"][tokens]
    for i in range(args.n_layers):
        p = f"layers.{i}."
        h = rms(x, sd[p + "norm1.weight"])
        qkv = h @ sd[p + "self_attn.in_proj_weight"].t() + sd[p + "self_attn.in_proj_bias"]
        q, k, v = (t.reshape(B, S, H, dh).transpose(1, 2) for t in qkv.chunk(3, -1))
        if args.rope:
            q, k = rope(q), rope(k)
        att = Fn.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, S, d)
        x = x + att @ sd[p + "self_attn.out_proj.weight"].t() + sd[p + "self_attn.out_proj.bias"]
        h = rms(x, sd[p + "norm3.weight"])
        g, u = (h @ sd[p + "linear1.weight"].t() + sd[p + "linear1.bias"]).chunk(2, -1)
        x = x + (Fn.silu(g) * u) @ sd[p + "linear2.weight"].t() + sd[p + "linear2.bias"]
    x = rms(x, sd["norm.weight"])
    return x @ sd["output.weight"].t() + sd["output.bias"]


def test_llama_style_block_matches_autograd_reference():
    torch.manual_seed(11)
    args = ModelArgs(dim=64, n_layers=2, n_heads=4, vocab_size=53, ffn_dim=48, dropout=0.0, causal=True, cross_attn=False,
                     activation="swiglu", norm_first=True, norm="rmsnorm", rope=True)
    model = Transformer(args)
    assert model.layers["0"].linear1.weight.shape == (96, 64)       # gate | up
    tokens = torch.randint(0, 53, (3, 16))
    target = torch.randint(0, 53, (3, 16))
    loss = torch.nn.functional.cross_entropy(model(tokens).reshape(-1, 53), target.reshape(-1))
    loss.backward()
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in model.state_dict().items()}
    loss_ref = torch.nn.functional.cross_entropy(_llama_reference(sd, args, tokens).reshape(-1, 53), target.reshape(-1))
    loss_ref.backward()
    assert abs(loss.item() - loss_ref.item()) < 1e-5
    for n, p in model.named_parameters():
        err = (p.grad - sd[n].grad).abs().max().item() / (sd[n].grad.abs().max().item() + 1e-9)
        assert err < 2e-4, (n, err)


def test_rope_and_swiglu_reference_ops():
    from ppb200.ops import functional as F, reference as ref
    x = torch.randn(32, 3 * 64)
    y = x.clone()
    F.rope_(y, 0, 16, 2 * 4, 16, 10000.0)
    assert torch.equal(y[:, 128:], x[:, 128:]) and not torch.allclose(y[:, :128], x[:, :128])   # v untouched
    assert torch.allclose(y[::16, :128], x[::16, :128])          # position 0 is the identity rotation
    F.rope_(y, 0, 16, 2 * 4, 16, 10000.0, inverse=True)
    assert torch.allclose(y, x, atol=1e-5)
    gu = torch.randn(8, 48, requires_grad=True)
    h = torch.nn.functional.silu(gu[:, :24]) * gu[:, 24:]
    dh = torch.randn(8, 24)
    h.backward(dh)
    assert torch.allclose(F.swiglu_fwd(gu.detach()), h.detach(), atol=1e-6)
    assert torch.allclose(F.swiglu_bwd(dh, gu.detach()), gu.grad, atol=1e-5)
