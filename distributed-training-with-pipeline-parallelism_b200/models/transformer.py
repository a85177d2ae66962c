"""The helper-compatible model (``Transformer``) and its explicit forward/backward stage program.

API parity with /root/reference/LLMsDistributedTrainingHelper.py:31-55:
  * ``Transformer(model_args)`` has ``.tok_embeddings``, ``.layers`` (``ModuleDict`` keyed by the GLOBAL layer
    index as a string), ``.norm``, ``.output``; any of them may be ``None``/deleted, so a pruned copy IS a
    pipeline stage (helper:48-55) and a stage's ``state_dict()`` is a strict subset of the full model's
    (SURVEY.md section 2.3 lists the exact keys; they are reproduced here name for name, shape for shape).
  * the decoder layer computes what ``nn.TransformerDecoderLayer(dim, heads, batch_first=True)`` computes when
    called as ``layer(h, h)``: post-LN self-attention, cross-attention over the layer's own input, ReLU FFN,
    six dropouts (torch transformer.py:1143-1199).

What is different: nothing here goes through autograd op by op.  ``layer_forward`` / ``layer_backward`` call the
fused ops of ``ppb200.ops.functional`` (hand-written sm_100a kernels on CUDA, fp32 torch math on CPU) and keep an
explicit activation stash, so one (stage, micro-batch) forward or backward is a fixed launch sequence that the
runner can capture in a CUDA graph.  ``Transformer.forward`` wraps the same program in a single
``autograd.Function`` so that user code (``loss.backward()``) keeps working.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn as nn

from ..ops import functional as F
from .args import ModelArgs

# dropout "sites" inside one decoder layer: each gets its own counter-based stream
SITE_ATTN1, SITE_RES1, SITE_ATTN2, SITE_RES2, SITE_FFN, SITE_RES3 = range(6)
_MASK64 = (1 << 64) - 1


def site_seed(base_seed: int, layer_id: int, mb: int, site: int) -> int:
    """Static part of a dropout seed (the step-varying part is added on the device, see StageProgram)."""
    x = (base_seed * 0x9E3779B97F4A7C15 + (layer_id + 1) * 0xD1B54A32D192ED03 + (mb + 1) * 0x94D049BB133111EB +
         (site + 1) * 0xBF58476D1CE4E5B9) & _MASK64
    x ^= x >> 31
    return (x * 0xD6E8FEB86659FD93) & _MASK64


# ------------------------------------------------------------------------------------------------------------
# parameter containers (names == the reference checkpoint keys)
# ------------------------------------------------------------------------------------------------------------
class _Proj(nn.Module):
    def __init__(self, n_in: int, n_out: int, dtype):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(n_out, n_in, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(n_out, dtype=dtype))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(n_in)
        nn.init.uniform_(self.bias, -bound, bound)


class _Attn(nn.Module):
    """Holds ``in_proj_weight`` (packed q,k,v on dim 0), ``in_proj_bias``, ``out_proj.{weight,bias}``."""

    def __init__(self, dim: int, dtype):
        super().__init__()
        self.in_proj_weight = nn.Parameter(torch.empty(3 * dim, dim, dtype=dtype))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * dim, dtype=dtype))
        self.out_proj = _Proj(dim, dim, dtype)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.zeros_(self.out_proj.bias)


class _Norm(nn.Module):
    def __init__(self, dim: int, dtype, rms: bool):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim, dtype=dtype))
        if not rms:
            self.bias = nn.Parameter(torch.zeros(dim, dtype=dtype))
        else:
            self.register_parameter("bias", None)


class DecoderLayer(nn.Module):
    def __init__(self, args: ModelArgs):
        super().__init__()
        d, dt, rms = args.dim, args.dtype, args.norm == "rmsnorm"
        self.self_attn = _Attn(d, dt)
        if args.cross_attn:
            self.multihead_attn = _Attn(d, dt)
        self.linear1 = _Proj(d, args.ffn_dim * (2 if args.activation == "swiglu" else 1), dt)   # swiglu: gate | up
        self.linear2 = _Proj(args.ffn_dim, d, dt)
        self.norm1 = _Norm(d, dt, rms)
        if args.cross_attn:
            self.norm2 = _Norm(d, dt, rms)
        self.norm3 = _Norm(d, dt, rms)


class _Embedding(nn.Module):
    def __init__(self, vocab: int, dim: int, dtype):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(vocab, dim, dtype=dtype))
        nn.init.normal_(self.weight)


# ------------------------------------------------------------------------------------------------------------
# explicit per-layer program
# ------------------------------------------------------------------------------------------------------------
class GradSink:
    """Maps parameter tensors to fp32 gradient accumulators (views of the stage's grad arena, or p.grad)."""

    def __init__(self):
        self._g: Dict[int, torch.Tensor] = {}
        self.fused_ar = None   # ops.collective.FusedGradAllReduce while the stage's LAST backward runs
        self.deferred: Optional[list] = None   # a list while an input-gradient-only pass (IR action I) runs
        # Weight-gradient work on a SIDE stream: nothing downstream of a backward needs dW / db before the optimizer, so the
        # wgrad GEMMs and bias column sums of a fused backward are forked off the input-gradient chain.  The persistent
        # kernels of the two streams cannot share an SM, but whenever the chain's kernel leaves SMs idle (the last partial
        # wave of an attention / GEMM launch) the side stream's CTAs fill them.  Works inside CUDA-graph capture (the fork
        # and the join become graph edges).  Operands are kept alive until the join.
        self.side_stream = None   # set by the stage (ops.functional.side_stream(device)); None = everything on one stream
        self._side_live: list = []
        self._side_prev = None

    def _on_side(self) -> bool:
        return self.side_stream is not None and self.fused_ar is None and self.deferred is None

    def _fork(self):
        ev = torch.cuda.Event()
        ev.record()
        self.side_stream.wait_event(ev)
        return torch.cuda.stream(self.side_stream)

    def side_layer_done(self) -> None:
        """End of a layer's backward: the main stream waits for the side work of the PREVIOUS layer (one layer of lag keeps
        the fork busy across the layer boundary) and drops the operands that work was reading."""
        if self.side_stream is None or not self._side_live:
            return
        ev = torch.cuda.Event()
        ev.record(self.side_stream)
        if self._side_prev is not None:
            torch.cuda.current_stream().wait_event(self._side_prev[0])
        self._side_prev = (ev, self._side_live)
        self._side_live = []

    def side_join(self) -> None:
        """End of a stage backward: everything forked so far is complete before the caller goes on (and before a graph
        capture ends)."""
        if self.side_stream is None:
            return
        if self._side_live or self._side_prev is not None:
            torch.cuda.current_stream().wait_stream(self.side_stream)
        self._side_live, self._side_prev = [], None

    def wgrad(self, dy: torch.Tensor, x: torch.Tensor, gw: torch.Tensor, gb: Optional[torch.Tensor]) -> None:
        """gw += dy^T x, gb += colsum(dy): now (action B) or recorded for the matching W action (torch's
        stage_backward_input / stage_backward_weight split, _backward.py:165/:283)."""
        if self.deferred is not None:
            self.deferred.append(("linear", dy, x, gw, gb))
        elif self._on_side() and dy.is_cuda:
            self._side_live.append((dy, x))
            with self._fork():
                F.linear_wgrad(dy, x, gw, gb)
        else:
            F.linear_wgrad(dy, x, gw, gb, ar=self.fused_ar)

    def bias_grad(self, dy: torch.Tensor, gb: torch.Tensor) -> None:
        """gb += colsum(dy) for a bias whose weight gradient is computed elsewhere (tensor-parallel row-parallel layers)."""
        if self.deferred is not None:
            self.deferred.append(("bias", dy, None, None, gb))
        elif self._on_side() and dy.is_cuda:
            self._side_live.append((dy,))
            with self._fork():
                F.colsum_accum(dy, gb)
        else:
            F.colsum_accum(dy, gb)

    def embedding_wgrad(self, tokens: torch.Tensor, dys: Sequence[torch.Tensor], gw: torch.Tensor) -> None:
        if self.deferred is not None:
            self.deferred.append(("embedding", tokens, list(dys), gw, None))
        else:
            F.embedding_bwd(tokens, dys, gw)

    def run_deferred(self, items: list) -> None:
        for kind, a, b, gw, gb in items:
            if kind == "linear":
                F.linear_wgrad(a, b, gw, gb, ar=self.fused_ar)
            elif kind == "bias":
                F.colsum_accum(a, gb)
            else:
                F.embedding_bwd(a, b, gw)

    def bind(self, param: torch.Tensor, grad: torch.Tensor) -> None:
        self._g[id(param)] = grad

    def __call__(self, param: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if param is None:
            return None
        return self._g[id(param)]


class Ctx:
    """Per-call context of the explicit program: micro-batch geometry, model args, dropout seeding."""

    def __init__(self, args: ModelArgs, batch: int, seq: int, base_seed: int, mb: int, training: bool = True,
                 step_offset: int = 0):
        self.args = args
        self.B, self.S = batch, seq
        self.base_seed, self.mb = base_seed, mb
        self.p = args.dropout if training else 0.0
        self.step_offset = step_offset  # CPU path only: on CUDA the offset lives in device memory
        self.tp = None                  # ops.collective.TPComm when the stage is tensor-parallel (models/tp.py)

    def seed(self, layer_id: int, site: int) -> int:
        return (site_seed(self.base_seed, layer_id, self.mb, site) + self.step_offset) & _MASK64


def _prenorm_forward(L: DecoderLayer, layer_id: int, x_in: torch.Tensor, ctx: Ctx, st: dict) -> torch.Tensor:
    """Pre-LN (GPT-style) block:  x += drop(attn(LN1(x)));  x += drop(ffn(LN3(x)))   -- causal self-attention only."""
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms, eps = a.norm == "rmsnorm", a.layer_norm_eps
    part = getattr(L, "part", "full")      # 'attn' / 'ffn': the stage boundary falls inside this layer (partition 'fine')
    if part == "ffn":                      # input = the residual stream after the attention block
        z3 = x_in
        _, y3, mean3, rstd3 = F.add_dropout_norm_fwd(z3, None, L.norm3.weight, L.norm3.bias, eps, 0.0, 0, rms, need_z=False)
        return _prenorm_ffn_forward(L, layer_id, z3, y3, mean3, rstd3, ctx, st)
    _, y1, mean1, rstd1 = F.add_dropout_norm_fwd(x_in, None, L.norm1.weight, L.norm1.bias, eps, 0.0, 0, rms, need_z=False)
    qkv = F.linear_fwd(y1, L.self_attn.in_proj_weight, L.self_attn.in_proj_bias)
    if a.rope:   # q and k are adjacent column blocks of the packed activation: one in-place pass over 2H heads
        F.rope_(qkv, 0, ctx.S, 2 * H, d // H, a.rope_theta)
    o1, lse1 = F.attention_fwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], ctx.B, H, a.causal, None, p,
                               ctx.seed(layer_id, SITE_ATTN1))
    a1 = F.linear_fwd(o1, L.self_attn.out_proj.weight, L.self_attn.out_proj.bias)
    st.update(x_in=x_in, y1=y1, mean1=mean1, rstd1=rstd1, qkv=qkv, o1=o1, lse1=lse1)
    if part == "attn":                     # the feed-forward block (and norm3) live on the next stage
        return F.add_dropout_fwd(x_in, a1, p, ctx.seed(layer_id, SITE_RES1))
    # residual add + dropout fused with the NEXT norm: z3 is the updated residual stream, y3 = LN3(z3)
    z3, y3, mean3, rstd3 = F.add_dropout_norm_fwd(x_in, a1, L.norm3.weight, L.norm3.bias, eps, p,
                                                  ctx.seed(layer_id, SITE_RES1), rms)
    return _prenorm_ffn_forward(L, layer_id, z3, y3, mean3, rstd3, ctx, st)


def _prenorm_ffn_forward(L: DecoderLayer, layer_id: int, z3, y3, mean3, rstd3, ctx: Ctx, st: dict) -> torch.Tensor:
    a, p = ctx.args, ctx.p
    if a.activation == "relu":
        h = F.linear_fwd(y3, L.linear1.weight, L.linear1.bias, "relu", p, ctx.seed(layer_id, SITE_FFN))
        st["h"] = h
    elif a.activation == "swiglu":
        hpre = F.linear_fwd(y3, L.linear1.weight, L.linear1.bias)     # [T, 2f]: gate | up
        h = F.swiglu_fwd(hpre)
        st.update(hpre=hpre, h=h)
    else:
        hpre = F.linear_fwd(y3, L.linear1.weight, L.linear1.bias)
        h = F.gelu_fwd(hpre)
        st.update(hpre=hpre, h=h)
    a3 = F.linear_fwd(h, L.linear2.weight, L.linear2.bias)
    x_out = F.add_dropout_fwd(z3, a3, p, ctx.seed(layer_id, SITE_RES3))
    st.update(z3=z3, y3=y3, mean3=mean3, rstd3=rstd3)
    return x_out


def _prenorm_backward(L: DecoderLayer, layer_id: int, dys: Sequence[torch.Tensor], ctx: Ctx, st: dict,
                      G: "GradSink") -> List[torch.Tensor]:
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms = a.norm == "rmsnorm"
    part = getattr(L, "part", "full")
    if part == "attn":      # dys are gradients w.r.t. z3 (the residual stream after the attention block)
        total, da1 = F.sum_dropout_bwd(list(dys), p, ctx.seed(layer_id, SITE_RES1))
        return _prenorm_attention_backward(L, layer_id, total, da1, ctx, st, G)
    g, da3 = F.sum_dropout_bwd(dys, p, ctx.seed(layer_id, SITE_RES3))          # d x_out, and through the FFN dropout
    G.wgrad(da3, st["h"], G(L.linear2.weight), G(L.linear2.bias))
    if a.activation == "relu":
        dh = F.linear_dgrad(da3, L.linear2.weight, mask_pos_aux=st["h"], alpha=(1.0 / (1.0 - p)) if p > 0 else 1.0)
    elif a.activation == "swiglu":
        dh = F.swiglu_bwd(F.linear_dgrad(da3, L.linear2.weight), st["hpre"])
    else:
        dh = F.linear_dgrad(da3, L.linear2.weight, gelu_aux=st["hpre"])
    G.wgrad(dh, st["y3"], G(L.linear1.weight), G(L.linear1.bias))
    dy3 = F.linear_dgrad(dh, L.linear1.weight)
    dz3_ln, _ = F.norm_bwd([dy3], st["z3"], st["mean3"], st["rstd3"], L.norm3.weight, G(L.norm3.weight), G(L.norm3.bias),
                           0.0, 0, rms, need_da=False)
    if part == "ffn":
        return [g, dz3_ln]
    total, da1 = F.sum_dropout_bwd([g, dz3_ln], p, ctx.seed(layer_id, SITE_RES1))  # d z3 = residual + through LN3
    return _prenorm_attention_backward(L, layer_id, total, da1, ctx, st, G)


def _prenorm_attention_backward(L: DecoderLayer, layer_id: int, total: torch.Tensor, da1: torch.Tensor, ctx: Ctx, st: dict,
                                G: "GradSink") -> List[torch.Tensor]:
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms = a.norm == "rmsnorm"
    G.wgrad(da1, st["o1"], G(L.self_attn.out_proj.weight), G(L.self_attn.out_proj.bias))
    do1 = F.linear_dgrad(da1, L.self_attn.out_proj.weight)
    qkv = st["qkv"]
    dqkv = torch.empty_like(qkv)
    F.attention_bwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], st["o1"], do1, st["lse1"], ctx.B, H, a.causal, None, p,
                    ctx.seed(layer_id, SITE_ATTN1), out=(dqkv[:, :d], dqkv[:, d:2 * d], dqkv[:, 2 * d:]))
    if a.rope:   # gradient of a rotation = the transpose rotation
        F.rope_(dqkv, 0, ctx.S, 2 * H, d // H, a.rope_theta, inverse=True)
    G.wgrad(dqkv, st["y1"], G(L.self_attn.in_proj_weight), G(L.self_attn.in_proj_bias))
    dy1 = F.linear_dgrad(dqkv, L.self_attn.in_proj_weight)
    dx_ln, _ = F.norm_bwd([dy1], st["x_in"], st["mean1"], st["rstd1"], L.norm1.weight, G(L.norm1.weight), G(L.norm1.bias),
                          0.0, 0, rms, need_da=False)
    return [total, dx_ln]


# Units of the post-LN reference layer (partition rule 'fine' may put a stage boundary behind any of them):
#   'self'  : x0 -> x1 = LN1(x0 + drop(SelfAttn(x0)))
#   'cross' : (x1, x0) -> x2 = LN2(x1 + drop(CrossAttn(q = x1, memory = x0)))
#   'ffn'   : x2 -> x3 = LN3(x2 + drop(W2 drop(relu(W1 x2))))
# A cut behind 'self' has to carry BOTH x1 and the layer input x0 (the cross-attention memory): the boundary tensor is
# the row-wise packing [x1 ; x0] of shape [2T, d], and its gradient is [d x1 ; d x0 (cross-attention share)].
_PARTS = {"full": ("self", "cross", "ffn"), "attn": ("self", "cross"), "self": ("self",), "cross": ("cross",),
          "cross+ffn": ("cross", "ffn"), "ffn": ("ffn",)}


def part_io_rows(part_first: str, part_last: str, cross_attn: bool) -> tuple:
    """(input rows multiplier, output rows multiplier) of a stage whose first / last layer own these parts."""
    if not cross_attn:
        return 1, 1
    return (2 if _PARTS[part_first][0] == "cross" else 1), (2 if _PARTS[part_last][-1] == "self" else 1)


def _postln_self_forward(L: DecoderLayer, layer_id: int, x0: torch.Tensor, ctx: Ctx, st: dict, packed_out: bool) -> torch.Tensor:
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms, eps = a.norm == "rmsnorm", a.layer_norm_eps
    st["x0"] = x0
    qkv = F.linear_fwd(x0, L.self_attn.in_proj_weight, L.self_attn.in_proj_bias)
    o1, lse1 = F.attention_fwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], ctx.B, H, a.causal, None, p,
                               ctx.seed(layer_id, SITE_ATTN1))
    a1 = F.linear_fwd(o1, L.self_attn.out_proj.weight, L.self_attn.out_proj.bias)
    out = None
    if packed_out:   # the cross-attention block lives on the next stage: ship [x1 ; x0]
        T = x0.shape[0]
        packed = torch.empty((2 * T, d), dtype=x0.dtype, device=x0.device)
        out = (torch.empty_like(x0), packed[:T], torch.empty(T, dtype=torch.float32, device=x0.device),
               torch.empty(T, dtype=torch.float32, device=x0.device))
    z1, x1, mean1, rstd1 = F.add_dropout_norm_fwd(x0, a1, L.norm1.weight, L.norm1.bias, eps, p,
                                                  ctx.seed(layer_id, SITE_RES1), rms, out=out)
    st.update(qkv=qkv, o1=o1, lse1=lse1, z1=z1, x1=x1, mean1=mean1, rstd1=rstd1)
    if packed_out:
        F.add_n([x0], out=packed[x0.shape[0]:])
        return packed
    return x1


def _postln_cross_forward(L: DecoderLayer, layer_id: int, x1: torch.Tensor, x0: torch.Tensor, ctx: Ctx, st: dict,
                          kv2: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Cross-attention over the layer input (memory = x0).  ``kv2``: the K/V projection when the caller already forked it
    onto the side stream (joined here)."""
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms, eps = a.norm == "rmsnorm", a.layer_norm_eps
    W, bvec = L.multihead_attn.in_proj_weight, L.multihead_attn.in_proj_bias
    q2 = F.linear_fwd(x1, W[:d], bvec[:d])
    if kv2 is None:
        kv2 = F.linear_fwd(x0, W[d:], bvec[d:])
    else:
        F.join(x0.device)
    o2, lse2 = F.attention_fwd(q2, kv2[:, :d], kv2[:, d:], ctx.B, H, False, None, p, ctx.seed(layer_id, SITE_ATTN2))
    a2 = F.linear_fwd(o2, L.multihead_attn.out_proj.weight, L.multihead_attn.out_proj.bias)
    z2, x2, mean2, rstd2 = F.add_dropout_norm_fwd(x1, a2, L.norm2.weight, L.norm2.bias, eps, p,
                                                  ctx.seed(layer_id, SITE_RES2), rms)
    st.update(x0=x0, x1=x1, q2=q2, kv2=kv2, o2=o2, lse2=lse2, z2=z2, x2=x2, mean2=mean2, rstd2=rstd2)
    return x2


def layer_forward(L: DecoderLayer, layer_id: int, x_in: torch.Tensor, ctx: Ctx, st: dict) -> torch.Tensor:
    """Decoder layer forward on [T, d] tokens (post-LN reference block, or the pre-LN GPT block when
    ``args.norm_first``).  Fills the stash ``st`` and returns the layer output.  A layer may hold only some of its
    units (``L.part``, partition rule 'fine'); see ``_PARTS`` for the boundary tensors."""
    if ctx.tp is not None:
        from .tp import tp_postln_forward, tp_prenorm_forward
        return (tp_prenorm_forward if ctx.args.norm_first else tp_postln_forward)(L, layer_id, x_in, ctx, st)
    if ctx.args.norm_first:
        return _prenorm_forward(L, layer_id, x_in, ctx, st)
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms, eps = a.norm == "rmsnorm", a.layer_norm_eps
    part = getattr(L, "part", "full")
    units = _PARTS[part] if a.cross_attn else tuple(u for u in _PARTS[part] if u != "cross")
    T = ctx.B * ctx.S if ctx.tp is None else x_in.shape[0]
    h = x_in
    # (forking the cross-attention K/V projection onto the side stream in FORWARD was measured and dropped: 8.08 -> 8.17 ms
    #  per micro-batch -- the forward chain has no long partial waves for it to fill, it only delays the chain's CTAs)
    if "self" in units:
        h = _postln_self_forward(L, layer_id, x_in, ctx, st, packed_out=(a.cross_attn and units[-1] == "self"))
        if units[-1] == "self":
            return h
    if "cross" in units:
        if units[0] == "cross":     # boundary tensor [x1 ; x0] from the stage that owns the self-attention block
            if x_in.shape[0] != 2 * T:
                raise ValueError(f"layer {layer_id} part {part!r} expects the packed [x1 ; x0] input of {2 * T} rows, got {x_in.shape[0]}")
            h = _postln_cross_forward(L, layer_id, x_in[:T], x_in[T:], ctx, st)
        else:
            h = _postln_cross_forward(L, layer_id, h, x_in, ctx, st)
        if units[-1] == "cross":
            return h
    # ---- feed-forward ----
    x2 = h
    st["x2" if a.cross_attn else "x1"] = x2
    if a.activation == "relu":
        hid = F.linear_fwd(x2, L.linear1.weight, L.linear1.bias, "relu", p, ctx.seed(layer_id, SITE_FFN))
        st["h"] = hid
    else:  # gelu: keep the pre-activation for backward (dropout on the hidden is not applied in this variant)
        hpre = F.linear_fwd(x2, L.linear1.weight, L.linear1.bias)
        hid = F.gelu_fwd(hpre)
        st.update(hpre=hpre, h=hid)
    a3 = F.linear_fwd(hid, L.linear2.weight, L.linear2.bias)
    z3, x3, mean3, rstd3 = F.add_dropout_norm_fwd(x2, a3, L.norm3.weight, L.norm3.bias, eps, p,
                                                  ctx.seed(layer_id, SITE_RES3), rms)
    st.update(z3=z3, mean3=mean3, rstd3=rstd3)
    return x3


def layer_backward(L: DecoderLayer, layer_id: int, dys: Sequence[torch.Tensor], ctx: Ctx, st: dict,
                   G: GradSink) -> List[torch.Tensor]:
    """Backward of ``layer_forward``.  ``dys`` = up to three gradients w.r.t. the layer output (summed on load by the
    norm kernel).  Accumulates parameter gradients into ``G`` and returns the (up to three) gradients w.r.t. the
    layer input (one packed [2T, d] tensor when the input was the packed [x1 ; x0] boundary tensor)."""
    if ctx.tp is not None:
        from .tp import tp_postln_backward, tp_prenorm_backward
        return (tp_prenorm_backward if ctx.args.norm_first else tp_postln_backward)(L, layer_id, dys, ctx, st, G)
    if ctx.args.norm_first:
        return _prenorm_backward(L, layer_id, dys, ctx, st, G)
    a, d, H, p = ctx.args, ctx.args.dim, ctx.args.n_heads, ctx.p
    rms = a.norm == "rmsnorm"
    part = getattr(L, "part", "full")
    units = _PARTS[part] if a.cross_attn else tuple(u for u in _PARTS[part] if u != "cross")
    dys = list(dys)
    dx0_extras: List[torch.Tensor] = []
    if "ffn" in units:
        dz3, da3 = F.norm_bwd(dys, st["z3"], st["mean3"], st["rstd3"], L.norm3.weight, G(L.norm3.weight), G(L.norm3.bias), p,
                              ctx.seed(layer_id, SITE_RES3), rms)
        x2 = st["x2"] if a.cross_attn else st["x1"]
        G.wgrad(da3, st["h"], G(L.linear2.weight), G(L.linear2.bias))
        if a.activation == "relu":
            dh = F.linear_dgrad(da3, L.linear2.weight, mask_pos_aux=st["h"], alpha=(1.0 / (1.0 - p)) if p > 0 else 1.0)
        else:
            dh = F.linear_dgrad(da3, L.linear2.weight, gelu_aux=st["hpre"])
        G.wgrad(dh, x2, G(L.linear1.weight), G(L.linear1.bias))
        dx2 = F.linear_dgrad(dh, L.linear1.weight)
        dys = [dz3, dx2]
        if units[0] == "ffn":
            return dys
    if "cross" in units:
        x0, T = st["x0"], st["x0"].shape[0]
        W = L.multihead_attn.in_proj_weight
        gW, gb = G(W), G(L.multihead_attn.in_proj_bias)
        dz2, da2 = F.norm_bwd(dys, st["z2"], st["mean2"], st["rstd2"], L.norm2.weight, G(L.norm2.weight),
                              G(L.norm2.bias), p, ctx.seed(layer_id, SITE_RES2), rms)
        G.wgrad(da2, st["o2"], G(L.multihead_attn.out_proj.weight), G(L.multihead_attn.out_proj.bias))
        do2 = F.linear_dgrad(da2, L.multihead_attn.out_proj.weight)
        kv2 = st["kv2"]
        dq2 = torch.empty_like(st["q2"])
        dkv2 = torch.empty_like(kv2)
        F.attention_bwd(st["q2"], kv2[:, :d], kv2[:, d:], st["o2"], do2, st["lse2"], ctx.B, H, False, None, p,
                        ctx.seed(layer_id, SITE_ATTN2), out=(dq2, dkv2[:, :d], dkv2[:, d:]))
        G.wgrad(dq2, st["x1"], gW[:d], gb[:d])
        G.wgrad(dkv2, x0, gW[d:], gb[d:])
        dx1 = F.linear_dgrad(dq2, W[:d])
        if units[0] == "cross":     # the self-attention block lives on the previous stage: ship [d x1 ; d x0]
            packed = torch.empty((2 * T, d), dtype=x0.dtype, device=x0.device)
            F.add_n([dz2, dx1], out=packed[:T])
            F.linear_dgrad(dkv2, W[d:], out=packed[T:])
            return [packed]
        dx0_extras = [F.linear_dgrad(dkv2, W[d:])]
        dys = [dz2, dx1]
    # ---- self-attention ----
    x0 = st["x0"]
    if a.cross_attn and units[-1] == "self":   # gradients of the packed [x1 ; x0] boundary tensor
        T = x0.shape[0]
        dx0_extras = [g[T:] for g in dys]
        dys = [g[:T] for g in dys]
    dz1, da1 = F.norm_bwd(dys, st["z1"], st["mean1"], st["rstd1"], L.norm1.weight, G(L.norm1.weight), G(L.norm1.bias), p,
                          ctx.seed(layer_id, SITE_RES1), rms)
    G.wgrad(da1, st["o1"], G(L.self_attn.out_proj.weight), G(L.self_attn.out_proj.bias))
    do1 = F.linear_dgrad(da1, L.self_attn.out_proj.weight)
    qkv = st["qkv"]
    dqkv = torch.empty_like(qkv)
    F.attention_bwd(qkv[:, :d], qkv[:, d:2 * d], qkv[:, 2 * d:], st["o1"], do1, st["lse1"], ctx.B, H, a.causal, None, p,
                    ctx.seed(layer_id, SITE_ATTN1), out=(dqkv[:, :d], dqkv[:, d:2 * d], dqkv[:, 2 * d:]))
    G.wgrad(dqkv, x0, G(L.self_attn.in_proj_weight), G(L.self_attn.in_proj_bias))
    dx0 = F.linear_dgrad(dqkv, L.self_attn.in_proj_weight)
    out = [dz1, dx0] + dx0_extras
    if len(out) > 3:    # the fan-in kernels sum at most three gradients
        out = [dz1, dx0, F.add_n(dx0_extras)]
    return out


# ------------------------------------------------------------------------------------------------------------
# stage-level program (embedding -> layers -> final norm -> LM head -> loss)
# ------------------------------------------------------------------------------------------------------------
def stage_forward(model: "Transformer", inp: torch.Tensor, ctx: Ctx, st: dict) -> torch.Tensor:
    """Forward of whatever pieces this (possibly pruned) model holds.  ``inp`` is int64 tokens [B, S] when the
    model has the embedding, else activations [B*S, d].  Returns [T, d] activations, or logits [T, V] (a view of a
    row-padded buffer) when the model holds the LM head."""
    a = ctx.args
    if model.tok_embeddings is not None:
        if ctx.tp is not None:   # sequence-parallel: this rank embeds its own rows of the flattened token stream
            inp = inp.reshape(-1)[ctx.tp.rank * ctx.tp.rows:(ctx.tp.rank + 1) * ctx.tp.rows]
        st["tokens"] = inp
        h = F.embedding_fwd(inp.reshape(-1), model.tok_embeddings.weight)
    else:
        h = inp.reshape(-1, a.dim)
    st["layers"] = {}
    for name, layer in model.layers.items():
        ls: dict = {}
        h = layer_forward(layer, int(name), h, ctx, ls)
        st["layers"][name] = ls
    if model.norm is not None:
        st["norm_in"] = h
        _, h, mean, rstd = F.add_dropout_norm_fwd(h, None, model.norm.weight, model.norm.bias, a.layer_norm_eps, 0.0, 0,
                                                  a.norm == "rmsnorm", need_z=False)
        st.update(norm_mean=mean, norm_rstd=rstd)
    if model.output is not None:
        st["head_in"] = h
        V = model.output.weight.shape[0]
        buf = torch.empty((h.shape[0], F.pad8(V)), dtype=h.dtype, device=h.device)
        h = F.linear_fwd(h, model.output.weight, model.output.bias, out=buf[:, :V])
    return h


def stage_loss(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float, st: dict, tp=None) -> torch.Tensor:
    """Mean token cross-entropy (helper:197-201) of one micro-batch; overwrites ``logits`` with its gradient
    (scaled by ``grad_scale``, e.g. 1 for a plain mean).  Returns a 1-element fp32 tensor.  With a tensor-parallel
    communicator ``logits`` are this rank's rows: the mean is over ALL rows and the result is this rank's share."""
    T = logits.shape[0]
    targets = targets.reshape(-1)
    if tp is not None:
        targets = targets[tp.rank * tp.rows:(tp.rank + 1) * tp.rows]
        T = T * tp.tp
    rows = F.xent_fwd_bwd(logits, targets, grad_scale / T, True)
    loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
    F.reduce_sum(rows, loss, 1.0 / T)
    st["dlogits"] = logits
    return loss


def stage_backward(model: "Transformer", douts: Sequence[torch.Tensor], ctx: Ctx, st: dict, G: GradSink,
                   need_input_grad: bool = True) -> Optional[torch.Tensor]:
    """Backward of ``stage_forward``.  ``douts`` = gradient(s) w.r.t. the stage output (for the last stage: the
    in-place dlogits).  Returns the gradient w.r.t. the stage input activations (None for the first stage)."""
    a = ctx.args
    dys: List[torch.Tensor] = [d.reshape(-1, d.shape[-1]) for d in douts]
    if model.output is not None:
        dlog = dys[0]
        G.wgrad(dlog, st["head_in"], G(model.output.weight), G(model.output.bias))
        dys = [F.linear_dgrad(dlog, model.output.weight)]
    if model.norm is not None:
        dz, _ = F.norm_bwd(dys, st["norm_in"], st["norm_mean"], st["norm_rstd"], model.norm.weight, G(model.norm.weight),
                           G(model.norm.bias), 0.0, 0, a.norm == "rmsnorm", need_da=False)
        dys = [dz]
    for name in reversed(list(model.layers.keys())):
        dys = layer_backward(model.layers[name], int(name), dys, ctx, st["layers"][name], G)
        G.side_layer_done()
    if model.tok_embeddings is not None:
        G.embedding_wgrad(st["tokens"].reshape(-1), dys, G(model.tok_embeddings.weight))
        G.side_join()
        return None
    out = F.add_n(dys) if need_input_grad else None
    G.side_join()
    return out


# ------------------------------------------------------------------------------------------------------------
# nn.Module front-end
# ------------------------------------------------------------------------------------------------------------
class _StageFn(torch.autograd.Function):
    """Bridges the explicit program into autograd: ONE node for the whole (pruned) model."""

    @staticmethod
    def forward(fctx, model, inp, dummy, *params):
        B, S = inp.shape[0], inp.shape[1]
        ctx = Ctx(model.args, B, S, model._seed, model._calls, model.training, 0)
        model._calls += 1
        st: dict = {}
        with torch.no_grad():
            out = stage_forward(model, inp, ctx, st)
        fctx.model, fctx.ctx, fctx.st = model, ctx, st
        fctx.inp_is_tokens = model.tok_embeddings is not None
        fctx.in_shape = inp.shape
        return out.reshape(B, S, -1)

    @staticmethod
    def backward(fctx, dout):
        model, ctx, st = fctx.model, fctx.ctx, fctx.st
        G = GradSink()
        params = list(model.parameters())
        acc = []
        for p in params:  # fp32 accumulators (p.grad must have the parameter's dtype, so they are folded in below)
            g = torch.zeros(p.shape, dtype=torch.float32, device=p.device)
            acc.append(g)
            G.bind(p, g)
        with torch.no_grad():
            dout = dout.contiguous()
            if model.output is not None and dout.is_cuda:
                V = dout.shape[-1]
                buf = torch.empty((dout.shape[0] * dout.shape[1], F.pad8(V)), dtype=dout.dtype, device=dout.device)
                buf[:, :V].copy_(dout.reshape(-1, V))
                dout = buf[:, :V]
            dx = stage_backward(model, [dout], ctx, st, G, need_input_grad=not fctx.inp_is_tokens)
            for p, g in zip(params, acc):
                if p.requires_grad:
                    p.grad = g.to(p.dtype) if p.grad is None else p.grad + g.to(p.dtype)
        if dx is not None:
            dx = dx.reshape(fctx.in_shape)
        return (None, dx, None) + (None,) * len(params)


class Transformer(nn.Module):
    """transformer model for pipeline parallelism (same constructor / attributes as helper:31-46)."""

    def __init__(self, model_args: ModelArgs):
        super().__init__()
        self.args = model_args
        dt = model_args.dtype
        self.tok_embeddings = _Embedding(model_args.vocab_size, model_args.dim, dt)
        self.layers = nn.ModuleDict()
        for layer_id in range(model_args.n_layers):
            self.layers[str(layer_id)] = DecoderLayer(model_args)
        self.norm = _Norm(model_args.dim, dt, model_args.norm == "rmsnorm")
        self.output = _Proj(model_args.dim, model_args.vocab_size, dt)
        self._seed = 0
        self._calls = 0
        self._dummy = None

    def set_seed(self, seed: int) -> None:
        self._seed, self._calls = int(seed), 0

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        """tokens: int64 [B, S] if this model holds the embedding, else activations [B, S, dim].
        Returns [B, S, dim] activations or [B, S, vocab] logits (helper:48-55)."""
        if tokens.dim() == 2 and self.tok_embeddings is None:
            raise ValueError("this stage has no embedding: pass activations [B, S, dim]")
        needs_grad = torch.is_grad_enabled() and (any(p.requires_grad for p in self.parameters())
                                                  or (tokens.is_floating_point() and tokens.requires_grad))
        if not needs_grad:
            B, S = tokens.shape[0], tokens.shape[1]
            ctx = Ctx(self.args, B, S, self._seed, self._calls, self.training, 0)
            self._calls += 1
            with torch.no_grad():
                return stage_forward(self, tokens, ctx, {}).reshape(B, S, -1)
        if self._dummy is None or self._dummy.device != tokens.device:
            self._dummy = torch.zeros(1, device=tokens.device, requires_grad=True)
        return _StageFn.apply(self, tokens, self._dummy, *list(self.parameters()))
