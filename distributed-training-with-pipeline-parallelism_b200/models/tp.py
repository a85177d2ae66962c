"""Intra-stage tensor parallelism (Megatron style, with sequence-parallel activations) for the pre-LN block.

Not in the reference ("TP (intra-stage): no"); required by BASELINE.json config 4 ("4-stage
1F1B x 2-way intra-stage TP, fused GEMM->reduce-scatter / all-gather->GEMM").  Layout per decoder layer, tp ranks:

    residual stream, norms, residual adds ......... [T/tp, d] row shards (sequence parallel, every rank its own rows)
    in_proj  (column-parallel, heads h0..h1) ....... qkv_local [T, 3d/tp] = all-gather(norm1(x)) @ Wqkv_local^T   (K12)
    attention ..................................... local heads only, all T rows
    out_proj (row-parallel) ....................... a1 [T/tp, d] = reduce-scatter(o_local @ Wo_local^T)          (K11)
    linear1  (column-parallel) .................... h_local [T, f/tp] = act(all-gather(norm3(z)) @ W1_local^T)   (K12)
    linear2  (row-parallel) ....................... a3 [T/tp, d] = reduce-scatter(h_local @ W2_local^T)          (K11)

Backward mirrors it: the input gradient of a row-parallel layer is an all-gather -> GEMM, the input gradient of a
column-parallel layer is a GEMM -> reduce-scatter, weight gradients are local GEMMs against the gathered rows.

Sharded parameters keep the reference's key names; split dims (for checkpoint merge, ``merge_tp_shards``):
    self_attn.in_proj_weight / in_proj_bias : dim 0, per head WITHIN each of q, k, v  (packed [3, H, dh, ...])
    self_attn.out_proj.weight               : dim 1
    linear1.weight / linear1.bias           : dim 0
    linear2.weight                          : dim 1
    everything else (norms, out_proj.bias, linear2.bias, embedding, head) is replicated; its gradient is computed
    on the local token rows and summed over the tp group once per step (``PipelineStage.reduce_tp_grads``).

Dropout (attention probabilities, the two residual branches) works with per-rank mask streams: the tensor-parallel
rank is mixed into every seed, so ranks draw independent masks for their rows / heads.  Only the ReLU block's extra
dropout on the hidden activations is not available here (its fused ReLU+dropout epilogue lives in the plain GEMM).
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.nn as nn

from ..ops import functional as F

SHARD_DIM = {"self_attn.in_proj_weight": "qkv0", "self_attn.in_proj_bias": "qkv0", "self_attn.out_proj.weight": 1,
             "multihead_attn.in_proj_weight": "qkv0", "multihead_attn.in_proj_bias": "qkv0", "multihead_attn.out_proj.weight": 1,
             "linear1.weight": 0, "linear1.bias": 0, "linear2.weight": 1}
# with activation == 'swiglu' linear1 packs gate | up on dim 0: each half is split separately ("pair0")


def _shard_key(name: str):
    for suffix, dim in SHARD_DIM.items():
        if name.endswith(suffix):
            return dim
    return None


def shard_tensor(name: str, t: torch.Tensor, n_heads: int, tp_rank: int, tp_size: int, swiglu: bool = False) -> torch.Tensor:
    dim = _shard_key(name)
    if dim is None:
        return t
    if swiglu and name.endswith(("linear1.weight", "linear1.bias")):
        v = t.reshape(2, t.shape[0] // 2, *t.shape[1:])
        n = v.shape[1]
        return v[:, tp_rank * n // tp_size:(tp_rank + 1) * n // tp_size].reshape(-1, *t.shape[1:]).contiguous()
    if dim == "qkv0":
        H = n_heads
        v = t.reshape(3, H, t.shape[0] // (3 * H), *t.shape[1:])
        h0, h1 = tp_rank * H // tp_size, (tp_rank + 1) * H // tp_size
        return v[:, h0:h1].reshape(-1, *t.shape[1:]).contiguous()
    n = t.shape[dim]
    return t.narrow(dim, tp_rank * n // tp_size, n // tp_size).contiguous()


def shard_model_for_tp(model, tp_rank: int, tp_size: int):
    """Replace the shardable parameters of every layer of a (pruned) ``Transformer`` by this rank's slices, in place."""
    a = model.args
    if tp_size == 1:
        return model
    if a.norm_first and a.dropout != 0.0 and a.activation == "relu":
        raise ValueError("tensor parallelism of the pre-LN block with dropout needs activation 'gelu' or 'swiglu'")
    if not a.norm_first and not a.cross_attn:
        raise ValueError("the post-LN tensor-parallel program is the reference block (self + cross attention)")
    if a.n_heads % tp_size or a.ffn_dim % tp_size or (a.ffn_dim // tp_size) % 8 or (a.dim // tp_size) % 8:
        raise ValueError(f"heads ({a.n_heads}) and ffn_dim ({a.ffn_dim}) must divide by tp={tp_size} into multiples of 8 columns")
    for name, p in list(model.named_parameters()):
        if _shard_key(name) is None:
            p._tp_replicated = True
            continue
        mod = model
        parts = name.split(".")
        for part in parts[:-1]:
            mod = getattr(mod, part) if not isinstance(mod, nn.ModuleDict) else mod[part]
        new = nn.Parameter(shard_tensor(name, p.data, a.n_heads, tp_rank, tp_size, a.activation == "swiglu"),
                           requires_grad=p.requires_grad)
        new._tp_replicated = False
        setattr(mod, parts[-1], new)
    model.tp_rank, model.tp_size = tp_rank, tp_size
    return model


def merge_tp_shards(shards: Sequence[Dict[str, torch.Tensor]], n_heads: int, swiglu: bool = False) -> Dict[str, torch.Tensor]:
    """Inverse of ``shard_tensor`` over whole state dicts: tp shards (same stage) -> the reference-layout state dict."""
    tp = len(shards)
    out: Dict[str, torch.Tensor] = {}
    for name in shards[0]:
        dim = _shard_key(name)
        ts = [sd[name] for sd in shards]
        if dim is None:
            out[name] = ts[0].clone()
        elif swiglu and name.endswith(("linear1.weight", "linear1.bias")):
            vs = [t.reshape(2, t.shape[0] // 2, *t.shape[1:]) for t in ts]
            out[name] = torch.cat(vs, dim=1).reshape(-1, *ts[0].shape[1:]).contiguous()
        elif dim == "qkv0":
            hl = n_heads // tp
            vs = [t.reshape(3, hl, t.shape[0] // (3 * hl), *t.shape[1:]) for t in ts]
            out[name] = torch.cat(vs, dim=1).reshape(-1, *ts[0].shape[1:]).contiguous()
        else:
            out[name] = torch.cat(ts, dim=dim).contiguous()
    return out


def _rs_finish(tp, y32: torch.Tensor, bias, dtype) -> torch.Tensor:
    """fp32 reduce-scatter result (+ replicated bias) -> activation dtype (one fused pass on CUDA)."""
    return tp.rs_finish(y32, bias, dtype)


def _tp_seed(ctx, layer_id: int, site: int) -> int:
    """Dropout stream of one site on THIS tensor-parallel rank: the ranks hold different token rows / heads under the same
    local indices, so the rank is mixed into the seed (independent masks; forward and backward use the same value)."""
    return (ctx.seed(layer_id, site) + 0x5851F42D4C957F2D * (ctx.tp.rank + 1)) & ((1 << 64) - 1)


def tp_prenorm_forward(L, layer_id: int, x_in: torch.Tensor, ctx, st: dict) -> torch.Tensor:
    """x_in: this rank's [T/tp, d] rows of the residual stream.  Returns the same rows of the layer output."""
    a, tp = ctx.args, ctx.tp
    if getattr(L, "part", "full") != "full":
        raise NotImplementedError("tensor parallelism with stage boundaries inside a layer (partition 'fine') is not implemented")
    dl, Hl = a.dim // tp.tp, a.n_heads // tp.tp
    rms, eps, p = a.norm == "rmsnorm", a.layer_norm_eps, ctx.p
    from .transformer import SITE_ATTN1, SITE_RES1, SITE_RES3
    _, y1, mean1, rstd1 = F.add_dropout_norm_fwd(x_in, None, L.norm1.weight, L.norm1.bias, eps, 0.0, 0, rms, need_z=False)
    qkv, y1_full = tp.ag_linear(y1, L.self_attn.in_proj_weight, L.self_attn.in_proj_bias)
    if a.rope:
        F.rope_(qkv, 0, ctx.S, 2 * Hl, a.dim // a.n_heads, a.rope_theta)
    o1, lse1 = F.attention_fwd(qkv[:, :dl], qkv[:, dl:2 * dl], qkv[:, 2 * dl:], ctx.B, Hl, a.causal, None, p,
                               _tp_seed(ctx, layer_id, SITE_ATTN1))
    a1 = _rs_finish(tp, tp.linear_rs(o1, L.self_attn.out_proj.weight), L.self_attn.out_proj.bias, x_in.dtype)
    z3, y3, mean3, rstd3 = F.add_dropout_norm_fwd(x_in, a1, L.norm3.weight, L.norm3.bias, eps, p,
                                                  _tp_seed(ctx, layer_id, SITE_RES1), rms)
    if a.activation == "relu":
        h, y3_full = tp.ag_linear(y3, L.linear1.weight, L.linear1.bias, "relu")
    elif a.activation == "swiglu":
        hpre, y3_full = tp.ag_linear(y3, L.linear1.weight, L.linear1.bias)    # [T, 2f/tp]: local gate | local up
        h = F.swiglu_fwd(hpre)
        st["hpre"] = hpre
    else:
        hpre, y3_full = tp.ag_linear(y3, L.linear1.weight, L.linear1.bias)
        h = F.gelu_fwd(hpre)
        st["hpre"] = hpre
    a3 = _rs_finish(tp, tp.linear_rs(h, L.linear2.weight), L.linear2.bias, x_in.dtype)
    x_out = F.add_dropout_fwd(z3, a3, p, _tp_seed(ctx, layer_id, SITE_RES3))
    st.update(x_in=x_in, y1_full=y1_full, mean1=mean1, rstd1=rstd1, qkv=qkv, o1=o1, lse1=lse1, z3=z3, y3_full=y3_full,
              mean3=mean3, rstd3=rstd3, h=h)
    return x_out


def tp_prenorm_backward(L, layer_id: int, dys: Sequence[torch.Tensor], ctx, st: dict, G) -> List[torch.Tensor]:
    a, tp = ctx.args, ctx.tp
    dl, Hl = a.dim // tp.tp, a.n_heads // tp.tp
    rms, p = a.norm == "rmsnorm", ctx.p
    from .transformer import SITE_ATTN1, SITE_RES1, SITE_RES3
    g, da3 = F.sum_dropout_bwd(dys, p, _tp_seed(ctx, layer_id, SITE_RES3))    # d x_out rows; d a3 = dropout'(d x_out)
    # linear2 (row-parallel):  d h_local = gather(d a3) @ W2_local ;  dW2_local += gather(d a3)^T h_local
    # (the activation derivative rides in the epilogue of the gather GEMM: ReLU mask / GELU' from the stashed activations)
    if a.activation == "relu":
        dh, g_full = tp.ag_linear(da3, L.linear2.weight, w_mn_major=True, mask_pos_aux=st["h"])
    elif a.activation == "swiglu":
        dh, g_full = tp.ag_linear(da3, L.linear2.weight, w_mn_major=True)
        dh = F.swiglu_bwd(dh, st["hpre"])
    else:
        dh, g_full = tp.ag_linear(da3, L.linear2.weight, w_mn_major=True, gelu_aux=st["hpre"])
    G.wgrad(g_full, st["h"], G(L.linear2.weight), None)
    G.bias_grad(da3, G(L.linear2.bias))
    # linear1 (column-parallel):  dW1_local += dh^T gather(y3) ;  d y3 rows = reduce-scatter(dh @ W1_local)
    G.wgrad(dh, st["y3_full"], G(L.linear1.weight), G(L.linear1.bias))
    dy3 = tp.rs_finish(tp.linear_rs(dh, L.linear1.weight, w_mn_major=True), None, g.dtype)
    dz3_ln, _ = F.norm_bwd([dy3], st["z3"], st["mean3"], st["rstd3"], L.norm3.weight, G(L.norm3.weight), G(L.norm3.bias),
                           0.0, 0, rms, need_da=False)
    total, da1 = F.sum_dropout_bwd([g, dz3_ln], p, _tp_seed(ctx, layer_id, SITE_RES1))   # d z3 rows; d a1 = dropout'(d z3)
    do1, t_full = tp.ag_linear(da1, L.self_attn.out_proj.weight, w_mn_major=True)
    G.wgrad(t_full, st["o1"], G(L.self_attn.out_proj.weight), None)
    G.bias_grad(da1, G(L.self_attn.out_proj.bias))
    qkv = st["qkv"]
    dqkv = torch.empty_like(qkv)
    F.attention_bwd(qkv[:, :dl], qkv[:, dl:2 * dl], qkv[:, 2 * dl:], st["o1"], do1, st["lse1"], ctx.B, Hl, a.causal, None,
                    p, _tp_seed(ctx, layer_id, SITE_ATTN1), out=(dqkv[:, :dl], dqkv[:, dl:2 * dl], dqkv[:, 2 * dl:]))
    if a.rope:
        F.rope_(dqkv, 0, ctx.S, 2 * Hl, a.dim // a.n_heads, a.rope_theta, inverse=True)
    G.wgrad(dqkv, st["y1_full"], G(L.self_attn.in_proj_weight), G(L.self_attn.in_proj_bias))
    dy1 = tp.rs_finish(tp.linear_rs(dqkv, L.self_attn.in_proj_weight, w_mn_major=True), None, g.dtype)
    dx_ln, _ = F.norm_bwd([dy1], st["x_in"], st["mean1"], st["rstd1"], L.norm1.weight, G(L.norm1.weight), G(L.norm1.bias),
                          0.0, 0, rms, need_da=False)
    return [total, dx_ln]


# ------------------------------------------------------------------------------------------------------------
# the reference's post-LN block (self-attention, cross-attention over the layer input, ReLU/GELU FFN) under TP
# ------------------------------------------------------------------------------------------------------------
def tp_postln_forward(L, layer_id: int, x0: torch.Tensor, ctx, st: dict) -> torch.Tensor:
    """x0: this rank's [T/tp, d] rows.  Same layout as the pre-LN program; the cross-attention block adds a
    column-parallel q projection of x1 and a column-parallel kv projection of the (already gathered) layer input."""
    from .transformer import SITE_ATTN1, SITE_ATTN2, SITE_FFN, SITE_RES1, SITE_RES2, SITE_RES3
    a, tp = ctx.args, ctx.tp
    if getattr(L, "part", "full") != "full":
        raise NotImplementedError("tensor parallelism with stage boundaries inside a layer (partition 'fine') is not implemented")
    dl, Hl, p = a.dim // tp.tp, a.n_heads // tp.tp, ctx.p
    rms, eps = a.norm == "rmsnorm", a.layer_norm_eps
    sd = lambda site: _tp_seed(ctx, layer_id, site)   # noqa: E731
    # ---- self-attention ----
    qkv, x0_full = tp.ag_linear(x0, L.self_attn.in_proj_weight, L.self_attn.in_proj_bias)
    o1, lse1 = F.attention_fwd(qkv[:, :dl], qkv[:, dl:2 * dl], qkv[:, 2 * dl:], ctx.B, Hl, a.causal, None, p, sd(SITE_ATTN1))
    a1 = _rs_finish(tp, tp.linear_rs(o1, L.self_attn.out_proj.weight), L.self_attn.out_proj.bias, x0.dtype)
    z1, x1, mean1, rstd1 = F.add_dropout_norm_fwd(x0, a1, L.norm1.weight, L.norm1.bias, eps, p, sd(SITE_RES1), rms)
    # ---- cross-attention: q from x1, k/v from the layer input ----
    W, bvec = L.multihead_attn.in_proj_weight, L.multihead_attn.in_proj_bias
    q2, x1_full = tp.ag_linear(x1, W[:dl], bvec[:dl])
    kv2 = F.linear_fwd(x0_full, W[dl:], bvec[dl:])
    o2, lse2 = F.attention_fwd(q2, kv2[:, :dl], kv2[:, dl:], ctx.B, Hl, False, None, p, sd(SITE_ATTN2))
    a2 = _rs_finish(tp, tp.linear_rs(o2, L.multihead_attn.out_proj.weight), L.multihead_attn.out_proj.bias, x0.dtype)
    z2, x2, mean2, rstd2 = F.add_dropout_norm_fwd(x1, a2, L.norm2.weight, L.norm2.bias, eps, p, sd(SITE_RES2), rms)
    # ---- feed-forward ----
    if a.activation == "relu":
        h, x2_full = tp.ag_linear(x2, L.linear1.weight, L.linear1.bias, "relu")
        if p > 0:   # the reference block also drops hidden activations; dropped entries stay 0, which is all backward needs
            h = F.add_dropout_fwd(torch.zeros_like(h), h, p, sd(SITE_FFN))
    else:
        hpre, x2_full = tp.ag_linear(x2, L.linear1.weight, L.linear1.bias)
        h = F.gelu_fwd(hpre)
        st["hpre"] = hpre
    a3 = _rs_finish(tp, tp.linear_rs(h, L.linear2.weight), L.linear2.bias, x0.dtype)
    z3, x3, mean3, rstd3 = F.add_dropout_norm_fwd(x2, a3, L.norm3.weight, L.norm3.bias, eps, p, sd(SITE_RES3), rms)
    st.update(x0_full=x0_full, qkv=qkv, o1=o1, lse1=lse1, z1=z1, mean1=mean1, rstd1=rstd1, x1_full=x1_full, q2=q2, kv2=kv2,
              o2=o2, lse2=lse2, z2=z2, mean2=mean2, rstd2=rstd2, x2_full=x2_full, h=h, z3=z3, mean3=mean3, rstd3=rstd3)
    return x3


def tp_postln_backward(L, layer_id: int, dys: Sequence[torch.Tensor], ctx, st: dict, G) -> List[torch.Tensor]:
    from .transformer import SITE_ATTN1, SITE_ATTN2, SITE_RES1, SITE_RES2, SITE_RES3
    a, tp = ctx.args, ctx.tp
    dl, Hl, p = a.dim // tp.tp, a.n_heads // tp.tp, ctx.p
    rms = a.norm == "rmsnorm"
    sd = lambda site: _tp_seed(ctx, layer_id, site)   # noqa: E731
    # ---- FFN ----
    dz3, da3 = F.norm_bwd(dys, st["z3"], st["mean3"], st["rstd3"], L.norm3.weight, G(L.norm3.weight), G(L.norm3.bias), p,
                          sd(SITE_RES3), rms)
    dtype = dz3.dtype
    if a.activation == "relu":   # ReLU (+ hidden dropout) mask and 1/(1-p) in the epilogue of the gather GEMM
        dh, da3_full = tp.ag_linear(da3, L.linear2.weight, w_mn_major=True, mask_pos_aux=st["h"],
                                    alpha=(1.0 / (1.0 - p)) if p > 0 else 1.0)
    else:
        dh, da3_full = tp.ag_linear(da3, L.linear2.weight, w_mn_major=True, gelu_aux=st["hpre"])
    G.wgrad(da3_full, st["h"], G(L.linear2.weight), None)
    G.bias_grad(da3, G(L.linear2.bias))
    G.wgrad(dh, st["x2_full"], G(L.linear1.weight), G(L.linear1.bias))
    dx2 = tp.rs_finish(tp.linear_rs(dh, L.linear1.weight, w_mn_major=True), None, dtype)
    # ---- cross-attention ----
    W = L.multihead_attn.in_proj_weight
    gW, gb = G(W), G(L.multihead_attn.in_proj_bias)
    dz2, da2 = F.norm_bwd([dz3, dx2], st["z2"], st["mean2"], st["rstd2"], L.norm2.weight, G(L.norm2.weight), G(L.norm2.bias),
                          p, sd(SITE_RES2), rms)
    do2, da2_full = tp.ag_linear(da2, L.multihead_attn.out_proj.weight, w_mn_major=True)
    G.wgrad(da2_full, st["o2"], G(L.multihead_attn.out_proj.weight), None)
    G.bias_grad(da2, G(L.multihead_attn.out_proj.bias))
    kv2 = st["kv2"]
    dq2 = torch.empty_like(st["q2"])
    dkv2 = torch.empty_like(kv2)
    F.attention_bwd(st["q2"], kv2[:, :dl], kv2[:, dl:], st["o2"], do2, st["lse2"], ctx.B, Hl, False, None, p, sd(SITE_ATTN2),
                    out=(dq2, dkv2[:, :dl], dkv2[:, dl:]))
    G.wgrad(dq2, st["x1_full"], gW[:dl], gb[:dl])
    G.wgrad(dkv2, st["x0_full"], gW[dl:], gb[dl:])
    dx1 = tp.rs_finish(tp.linear_rs(dq2, W[:dl], w_mn_major=True), None, dtype)   # (a fresh tensor: the symmetric buffer is reused two calls on)
    dx0_extra = tp.rs_finish(tp.linear_rs(dkv2, W[dl:], w_mn_major=True), None, dtype)
    # ---- self-attention ----
    dz1, da1 = F.norm_bwd([dz2, dx1], st["z1"], st["mean1"], st["rstd1"], L.norm1.weight, G(L.norm1.weight), G(L.norm1.bias),
                          p, sd(SITE_RES1), rms)
    do1, da1_full = tp.ag_linear(da1, L.self_attn.out_proj.weight, w_mn_major=True)
    G.wgrad(da1_full, st["o1"], G(L.self_attn.out_proj.weight), None)
    G.bias_grad(da1, G(L.self_attn.out_proj.bias))
    qkv = st["qkv"]
    dqkv = torch.empty_like(qkv)
    F.attention_bwd(qkv[:, :dl], qkv[:, dl:2 * dl], qkv[:, 2 * dl:], st["o1"], do1, st["lse1"], ctx.B, Hl, a.causal, None, p,
                    sd(SITE_ATTN1), out=(dqkv[:, :dl], dqkv[:, dl:2 * dl], dqkv[:, 2 * dl:]))
    G.wgrad(dqkv, st["x0_full"], G(L.self_attn.in_proj_weight), G(L.self_attn.in_proj_bias))
    dx0 = tp.rs_finish(tp.linear_rs(dqkv, L.self_attn.in_proj_weight, w_mn_major=True), None, dtype)
    return [dz1, dx0, dx0_extra]
