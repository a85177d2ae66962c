"""Functional op layer: every hot op of a pipeline stage, with explicit forward AND backward entry points.

CUDA bf16 tensors go to the hand-written sm_100a kernels of libppb200.so (ctypes, C ABI, launched on the
current torch stream, CUDA-graph capturable).  CPU tensors go to the fp32 references in ``ops.reference``.
There is no PyTorch fallback for CUDA tensors: a missing library raises.

The op set mirrors SURVEY.md section 8 (K1-K9) and replaces the ATen op stream of the reference
(/root/reference/LLMsDistributedTrainingHelper.py:31-55).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch

from . import _lib, reference as ref

OUT_BF16, OUT_F32, OUT_F32_ACCUM = 0, 1, 2
EPI_NONE, EPI_BIAS, EPI_BIAS_RELU, EPI_BIAS_GELU, EPI_BIAS_RELU_DROPOUT, EPI_MASK_POS, EPI_GELU_BWD = range(7)


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _lib_required():
    return _lib.load(required=True)


def _is_cuda(t: torch.Tensor) -> bool:
    return t.is_cuda


def _check_2d(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be 2-D with a contiguous last dim, got shape {tuple(t.shape)} stride {t.stride()}")


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


# ------------------------------------------------------------------------------------------------------------
# side stream: independent branches of the stage program (weight gradients, the cross-attention K/V projection) are
# forked off the critical chain so that their CTAs fill the SMs the chain's kernels leave idle in their last partial wave
# (persistent tcgen05 kernels hold a whole SM, so two launches overlap exactly where one of them runs out of tiles).
# Fork and join are event edges, which CUDA-graph capture records as graph dependencies.  PPB_SIDE_STREAM=0 turns it off.
# ------------------------------------------------------------------------------------------------------------
_side_streams: dict = {}
SIDE_STREAM = __import__("os").environ.get("PPB_SIDE_STREAM", "1") != "0"


def side_stream(device: torch.device):
    """The per-device side stream (None on CPU or when disabled)."""
    if not SIDE_STREAM or device.type != "cuda":
        return None
    key = device.index if device.index is not None else torch.cuda.current_device()
    st = _side_streams.get(key)
    if st is None:
        st = _side_streams[key] = torch.cuda.Stream(device)
    return st


class fork:
    """``with fork(device) as on_side:`` -- work issued inside runs on the side stream, ordered after everything issued so
    far on the current stream (no-op context when there is no side stream).  ``join(device)`` orders the current stream
    after everything forked so far."""

    def __init__(self, device: torch.device):
        self.side = side_stream(device)
        self.ctx = None

    def __enter__(self):
        if self.side is not None:
            ev = torch.cuda.Event()
            ev.record()
            self.side.wait_event(ev)
            self.ctx = torch.cuda.stream(self.side)
            self.ctx.__enter__()
        return self.side is not None

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def join(device: torch.device) -> None:
    st = side_stream(device)
    if st is not None:
        torch.cuda.current_stream().wait_stream(st)


# ------------------------------------------------------------------------------------------------------------
# GEMM family
# ------------------------------------------------------------------------------------------------------------
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, a_mn: bool, b_mn: bool,
         out_mode: int = OUT_BF16, epi: int = EPI_NONE, alpha: float = 1.0, bias: Optional[torch.Tensor] = None,
         aux: Optional[torch.Tensor] = None, seed: int = 0, dropout_p: float = 0.0, force_bn: int = 0,
         force_splits: int = 0) -> torch.Tensor:
    """Raw tcgen05 GEMM: out[M,N] (=|+=) epi(alpha * op(a)[M,K] @ op(b)[N,K]^T).  CUDA only."""
    lib = _lib_required()
    rc = lib.ppb_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a.stride(0), b.stride(0),
                           out.stride(0), int(a_mn), int(b_mn), out_mode, epi, float(alpha), _lib.ptr(bias),
                           _lib.ptr(aux), 0 if aux is None else aux.stride(0), int(seed) & (2**64 - 1),
                           float(dropout_p), force_bn, force_splits, _stream())
    _lib.check(rc, "ppb_gemm_bf16")
    _lib.count_launch()
    return out


def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
               dropout_p: float = 0.0, seed: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """y = dropout(act(x @ w^T + bias)).  x [T, in], w [out, in] -> y [T, out] (``out`` may be a padded-ld view)."""
    _check_2d(x, "x")
    _check_2d(w, "w")
    if not _is_cuda(x):
        y = ref.linear_fwd(x, w, bias, act, dropout_p, seed)
        if out is not None:
            out.copy_(y)
            return out
        return y
    T, K = x.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty((T, N), dtype=torch.bfloat16, device=x.device)
    if bias is None:
        if act is not None or dropout_p > 0:
            raise ValueError("activation/dropout epilogues need a bias")
        epi = EPI_NONE
    elif act is None:
        epi = EPI_BIAS
    elif act == "relu":
        epi = EPI_BIAS_RELU_DROPOUT if dropout_p > 0 else EPI_BIAS_RELU
    elif act == "gelu":
        if dropout_p > 0:
            raise ValueError("gelu+dropout epilogue not implemented")
        epi = EPI_BIAS_GELU
    else:
        raise ValueError(act)
    return gemm(x, w, out, T, N, K, False, False, OUT_BF16, epi, 1.0, bias, None, seed, dropout_p)


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, mask_pos_aux: Optional[torch.Tensor] = None,
                 gelu_aux: Optional[torch.Tensor] = None, alpha: float = 1.0,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """dx = alpha * (dy @ w) [* (aux > 0)] [* gelu'(aux)].  dy [T, out], w [out, in] -> dx [T, in]."""
    _check_2d(dy, "dy")
    _check_2d(w, "w")
    if not _is_cuda(dy):
        dx = ref.linear_dgrad(dy, w, mask_pos_aux, gelu_aux, alpha)
        if out is not None:
            out.copy_(dx)
            return out
        return dx
    T, K = dy.shape  # K = out features
    N = w.shape[1]
    if out is None:
        out = torch.empty((T, N), dtype=torch.bfloat16, device=dy.device)
    epi, aux = EPI_NONE, None
    if mask_pos_aux is not None:
        epi, aux = EPI_MASK_POS, mask_pos_aux
    elif gelu_aux is not None:
        epi, aux = EPI_GELU_BWD, gelu_aux
    return gemm(dy, w, out, T, N, K, False, True, OUT_BF16, epi, alpha, None, aux)


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, dw_accum: torch.Tensor,
                 db_accum: Optional[torch.Tensor] = None, ar=None) -> None:
    """dw_accum[out, in] += dy^T @ x (fp32, split-K + TMA reduce-add);  db_accum[out] += colsum(dy).

    ``ar``: an armed ``ops.collective.FusedGradAllReduce`` -- the GEMM then pushes (previous accumulation + this
    micro-batch) into every data-parallel replica's reduced-gradient buffer instead (K10)."""
    _check_2d(dy, "dy")
    _check_2d(x, "x")
    if not _is_cuda(dy):
        ref.linear_wgrad(dy, x, dw_accum, db_accum)
        return
    T, Nout = dy.shape
    Nin = x.shape[1]
    if ar is not None and ar.armed:
        ar.wgrad(dy, x, dw_accum)
    else:
        # split-K partials land with TMA reduce-add in arrival order: fp32-rounding-level run-to-run differences
        # (which Adam can amplify for near-zero gradients); the deterministic mode keeps each output tile in one CTA
        gemm(dy, x, dw_accum, Nout, Nin, T, True, True, OUT_F32_ACCUM, force_splits=1 if DETERMINISTIC else 0)
    if db_accum is not None:
        lib = _lib_required()
        _lib.check(lib.ppb_colsum_accum(dy.data_ptr(), dy.stride(0), db_accum.data_ptr(), T, Nout, _stream()),
                   "ppb_colsum_accum")
        _lib.count_launch()


def colsum_accum(dy: torch.Tensor, db_accum: torch.Tensor) -> None:
    """db_accum[out] += colsum(dy)  (bias gradient without its weight-gradient GEMM)."""
    _check_2d(dy, "dy")
    if not _is_cuda(dy):
        db_accum.add_(dy.float().sum(0).to(db_accum.dtype))
        return
    lib = _lib_required()
    _lib.check(lib.ppb_colsum_accum(dy.data_ptr(), dy.stride(0), db_accum.data_ptr(), dy.shape[0], dy.shape[1], _stream()),
               "ppb_colsum_accum")
    _lib.count_launch()


# ------------------------------------------------------------------------------------------------------------
# fused residual + dropout + norm
# ------------------------------------------------------------------------------------------------------------
_scratch: dict = {}
import os as _os
DETERMINISTIC = _os.environ.get("PPB_DETERMINISTIC", "0") == "1"   # bit-reproducible weight gradients (no split-K)


def _norm_scratch(device: torch.device, d: int) -> torch.Tensor:
    lib = _lib_required()
    key = (device.index, torch.cuda.current_stream().cuda_stream, "norm")
    need = int(lib.ppb_norm_bwd_scratch_bytes(0, d))
    if torch.cuda.is_current_stream_capturing():
        # a capture gets its own scratch from the graph's private pool (static address, reusable by the rest of the
        # capture once this call is over); a process-wide cache entry would tie graphs of unrelated pools together
        return torch.empty(need, dtype=torch.uint8, device=device)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.zeros(need, dtype=torch.uint8, device=device)  # zero-filled once; the kernel self-resets
        _scratch[key] = buf
    return buf


def add_dropout_norm_fwd(resid: Optional[torch.Tensor], a: Optional[torch.Tensor], gamma: torch.Tensor,
                         beta: Optional[torch.Tensor], eps: float = 1e-5, dropout_p: float = 0.0, seed: int = 0,
                         rms: bool = False, need_z: bool = True, out=None):
    """z = resid + dropout(a); y = norm(z).  Returns (z, y, mean, rstd); ``out`` = optional (z, y, mean, rstd) buffers."""
    t0 = resid if resid is not None else a
    if not _is_cuda(t0):
        z, y, mean, rstd = ref.add_dropout_norm_fwd(resid, a, gamma, beta, eps, dropout_p, seed, rms)
        if out is not None:
            for dst, src in zip(out, (z, y, mean, rstd)):
                if dst is not None:
                    dst.copy_(src)
            return out
        return z, y, mean, rstd
    lib = _lib_required()
    T, d = t0.shape
    if out is None:
        z = torch.empty_like(t0) if need_z else None
        y = torch.empty_like(t0)
        mean = torch.empty(T, dtype=torch.float32, device=t0.device)
        rstd = torch.empty(T, dtype=torch.float32, device=t0.device)
    else:
        z, y, mean, rstd = out
    rc = lib.ppb_add_dropout_norm_fwd(_lib.ptr(resid), _lib.ptr(a), gamma.data_ptr(), _lib.ptr(beta), _lib.ptr(z),
                                      y.data_ptr(), _lib.ptr(mean), rstd.data_ptr(), T, d, float(eps),
                                      float(dropout_p), int(seed) & (2**64 - 1), int(rms), _stream())
    _lib.check(rc, "ppb_add_dropout_norm_fwd")
    _lib.count_launch()
    return z, y, mean, rstd


def norm_bwd(dys: Sequence[Optional[torch.Tensor]], z: torch.Tensor, mean: Optional[torch.Tensor], rstd: torch.Tensor,
             gamma: torch.Tensor, dgamma: torch.Tensor, dbeta: Optional[torch.Tensor], dropout_p: float = 0.0,
             seed: int = 0, rms: bool = False, need_da: bool = True, out=None):
    """Backward of add_dropout_norm_fwd.  dys: up to three gradients w.r.t. y (summed).  Returns (dz, da);
    da is dz itself when there is no dropout."""
    dys = [d for d in dys if d is not None]
    if not _is_cuda(z):
        dz, da = ref.norm_bwd(dys, z, mean, rstd, gamma, dgamma, dbeta, dropout_p, seed, rms, need_da)
        if out is not None:
            out[0].copy_(dz)
            if da is not None and out[1] is not None:
                out[1].copy_(da)
                return out[0], out[1]
            return out[0], out[0]
        return dz, da
    lib = _lib_required()
    T, d = z.shape
    if len(dys) > 3:
        raise ValueError("norm_bwd sums at most three incoming gradients")
    dys = dys + [None] * (3 - len(dys))
    if out is None:
        dz = torch.empty_like(z)
        da = torch.empty_like(z) if (need_da and dropout_p > 0) else None
    else:
        dz, da = out
        if not (need_da and dropout_p > 0):
            da = None
    scratch = _norm_scratch(z.device, d)
    rc = lib.ppb_norm_bwd(dys[0].data_ptr(), _lib.ptr(dys[1]), _lib.ptr(dys[2]), z.data_ptr(), _lib.ptr(mean),
                          rstd.data_ptr(), gamma.data_ptr(), dz.data_ptr(), _lib.ptr(da), dgamma.data_ptr(),
                          _lib.ptr(dbeta), scratch.data_ptr(), scratch.numel(), T, d, float(dropout_p),
                          int(seed) & (2**64 - 1), int(rms), _stream())
    _lib.check(rc, "ppb_norm_bwd")
    _lib.count_launch()
    return dz, (da if da is not None else dz)


# ------------------------------------------------------------------------------------------------------------
# loss
# ------------------------------------------------------------------------------------------------------------
def xent_fwd_bwd(logits: torch.Tensor, targets: torch.Tensor, grad_scale: float = 1.0, write_grad: bool = True,
                 row_loss: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-row cross-entropy; overwrites ``logits`` ([T, V] view, row stride % 8 == 0) with grad_scale*(softmax-onehot)."""
    T, V = logits.shape
    if not _is_cuda(logits):
        loss = ref.xent_fwd_bwd(logits, targets, grad_scale, write_grad)
        if row_loss is not None:
            row_loss.copy_(loss)
            return row_loss
        return loss
    lib = _lib_required()
    if row_loss is None:
        row_loss = torch.empty(T, dtype=torch.float32, device=logits.device)
    rc = lib.ppb_xent_fwd_bwd(logits.data_ptr(), logits.stride(0), targets.data_ptr(), row_loss.data_ptr(), T, V,
                              float(grad_scale), int(write_grad), _stream())
    _lib.check(rc, "ppb_xent_fwd_bwd")
    _lib.count_launch()
    return row_loss


def reduce_sum(x: torch.Tensor, out: torch.Tensor, scale: float = 1.0, accumulate: bool = False) -> torch.Tensor:
    if not _is_cuda(x):
        s = x.float().sum() * scale
        out.copy_(out + s if accumulate else s)
        return out
    lib = _lib_required()
    _lib.check(lib.ppb_reduce_sum_f32(x.data_ptr(), x.numel(), out.data_ptr(), float(scale), int(accumulate), _stream()),
               "ppb_reduce_sum_f32")
    _lib.count_launch()
    return out


# ------------------------------------------------------------------------------------------------------------
# optimizer
# ------------------------------------------------------------------------------------------------------------
def adamw_step(param: Optional[torch.Tensor], master: torch.Tensor, m: torch.Tensor, v: torch.Tensor,
               grad: torch.Tensor, lr: float, beta1: float, beta2: float, eps: float, weight_decay: float, step: int,
               grad_scale: float = 1.0, clip_coef: Optional[torch.Tensor] = None, zero_grad: bool = True) -> None:
    """One fused AdamW step over flat arenas (param = bf16 working copy or None; the rest fp32)."""
    if not _is_cuda(master):
        ref.adamw_step(param, master, m, v, grad, lr, beta1, beta2, eps, weight_decay, step, grad_scale, clip_coef,
                       zero_grad)
        return
    lib = _lib_required()
    rc = lib.ppb_adamw_step(_lib.ptr(param), master.data_ptr(), m.data_ptr(), v.data_ptr(), grad.data_ptr(),
                            master.numel(), lr, beta1, beta2, eps, weight_decay, step, grad_scale,
                            _lib.ptr(clip_coef), int(zero_grad), _stream())
    _lib.check(rc, "ppb_adamw_step")
    _lib.count_launch()


def sumsq_accum(x: torch.Tensor, out: torch.Tensor) -> None:
    if not _is_cuda(x):
        out.add_((x.float() ** 2).sum())
        return
    lib = _lib_required()
    _lib.check(lib.ppb_sumsq_f32(x.data_ptr(), x.numel(), out.data_ptr(), _stream()), "ppb_sumsq_f32")
    _lib.count_launch()


def clip_coef(sumsq: torch.Tensor, grad_scale: float, max_norm: float, coef: torch.Tensor) -> None:
    if not _is_cuda(sumsq):
        nrm = float(sumsq.sqrt()) * grad_scale
        coef.fill_(min(1.0, max_norm / (nrm + 1e-6)))
        return
    lib = _lib_required()
    _lib.check(lib.ppb_clip_coef(sumsq.data_ptr(), grad_scale, max_norm, coef.data_ptr(), _stream()), "ppb_clip_coef")
    _lib.count_launch()


# ------------------------------------------------------------------------------------------------------------
# embedding / small elementwise
# ------------------------------------------------------------------------------------------------------------
def embedding_fwd(idx: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not _is_cuda(w):
        y = ref.embedding_fwd(idx, w)
        if out is not None:
            out.copy_(y)
            return out
        return y
    lib = _lib_required()
    T = idx.numel()
    V, d = w.shape
    if out is None:
        out = torch.empty((T, d), dtype=w.dtype, device=w.device)
    _lib.check(lib.ppb_embedding_fwd(idx.data_ptr(), w.data_ptr(), out.data_ptr(), T, d, V, _stream()),
               "ppb_embedding_fwd")
    _lib.count_launch()
    return out


def embedding_bwd(idx: torch.Tensor, dys: Sequence[Optional[torch.Tensor]], dw_accum: torch.Tensor) -> None:
    dys = [d for d in dys if d is not None]
    if not _is_cuda(dw_accum):
        ref.embedding_bwd(idx, dys, dw_accum)
        return
    lib = _lib_required()
    dys = dys + [None] * (3 - len(dys))
    V, d = dw_accum.shape
    _lib.check(lib.ppb_embedding_bwd(idx.data_ptr(), dys[0].data_ptr(), _lib.ptr(dys[1]), _lib.ptr(dys[2]),
                                     dw_accum.data_ptr(), idx.numel(), d, V, _stream()), "ppb_embedding_bwd")
    _lib.count_launch()


def add_n(ts: Sequence[Optional[torch.Tensor]], out: Optional[torch.Tensor] = None) -> torch.Tensor:
    ts = [t for t in ts if t is not None]
    if len(ts) == 1 and out is None:
        return ts[0]
    if not _is_cuda(ts[0]):
        y = ref.add_n(ts)
        if out is not None:
            out.copy_(y)
            return out
        return y
    lib = _lib_required()
    if out is None:
        out = torch.empty_like(ts[0])
    ts = ts + [None] * (3 - len(ts))
    _lib.check(lib.ppb_add_n_bf16(ts[0].data_ptr(), _lib.ptr(ts[1]), _lib.ptr(ts[2]), out.data_ptr(), out.numel(),
                                  _stream()), "ppb_add_n_bf16")
    _lib.count_launch()
    return out


def add_dropout_fwd(resid: torch.Tensor, a: torch.Tensor, dropout_p: float = 0.0, seed: int = 0) -> torch.Tensor:
    """z = resid + dropout(a) (no norm; the residual branch of pre-LN blocks)."""
    if not _is_cuda(resid):
        return (resid.float() + (ref.dropout(a.float(), dropout_p, seed) if dropout_p > 0 else a.float())).to(resid.dtype)
    lib = _lib_required()
    z = torch.empty_like(resid)
    _lib.check(lib.ppb_add_dropout_fwd(resid.data_ptr(), a.data_ptr(), z.data_ptr(), z.numel(), float(dropout_p),
                                       int(seed) & (2**64 - 1), _stream()), "ppb_add_dropout_fwd")
    _lib.count_launch()
    return z


def sum_dropout_bwd(dys: Sequence[Optional[torch.Tensor]], dropout_p: float = 0.0, seed: int = 0, need_da: bool = True):
    """dz = sum(dys) (up to three); da = dropout'(dz).  Returns (dz, da) -- da is dz itself without dropout."""
    dys = [d for d in dys if d is not None]
    if not _is_cuda(dys[0]):
        dz = ref.add_n(dys)
        da = ref.dropout(dz.float(), dropout_p, seed).to(dz.dtype) if (need_da and dropout_p > 0) else dz
        return dz, da
    lib = _lib_required()
    if len(dys) == 1 and not (need_da and dropout_p > 0):
        return dys[0], dys[0]
    dz = torch.empty_like(dys[0])
    da = torch.empty_like(dys[0]) if (need_da and dropout_p > 0) else None
    dys = dys + [None] * (3 - len(dys))
    _lib.check(lib.ppb_sum_dropout_bwd(dys[0].data_ptr(), _lib.ptr(dys[1]), _lib.ptr(dys[2]), dz.data_ptr(), _lib.ptr(da),
                                       dz.numel(), float(dropout_p), int(seed) & (2**64 - 1), _stream()), "ppb_sum_dropout_bwd")
    _lib.count_launch()
    return dz, (da if da is not None else dz)


def gelu_fwd(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if not _is_cuda(x):
        y = torch.nn.functional.gelu(x.float()).to(x.dtype)
        if out is not None:
            out.copy_(y)
            return out
        return y
    lib = _lib_required()
    if out is None:
        out = torch.empty_like(x)
    _lib.check(lib.ppb_gelu_fwd(x.data_ptr(), out.data_ptr(), x.numel(), _stream()), "ppb_gelu_fwd")
    _lib.count_launch()
    return out


# ------------------------------------------------------------------------------------------------------------
# attention
# ------------------------------------------------------------------------------------------------------------
def attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, B: int, H: int, causal: bool = False,
                  scale: Optional[float] = None, dropout_p: float = 0.0, seed: int = 0,
                  out: Optional[torch.Tensor] = None, lse: Optional[torch.Tensor] = None):
    """Flash attention forward.  q [B*Sq, H*dh], k/v [B*Skv, H*dh] (column-sliced views of packed QKV are fine).
    Returns (o [B*Sq, H*dh], lse [B, H, Sq])."""
    if not _is_cuda(q):
        o, l = ref.attention_fwd(q, k, v, B, H, causal, scale, dropout_p, seed)
        if out is not None:
            out.copy_(o)
            o = out
        if lse is not None:
            lse.copy_(l)
            l = lse
        return o, l
    lib = _lib_required()
    Tq, D = q.shape
    dh = D // H
    Sq, Skv = Tq // B, k.shape[0] // B
    scale = scale if scale is not None else 1.0 / math.sqrt(dh)
    if out is None:
        out = torch.empty((Tq, D), dtype=q.dtype, device=q.device)
    if lse is None:
        lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device)
    rc = lib.ppb_attn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(), B, H, Sq, Skv, dh,
                          q.stride(0), k.stride(0), v.stride(0), out.stride(0), int(causal), 0, 0, float(scale),
                          float(dropout_p), int(seed) & (2**64 - 1), _stream())
    _lib.check(rc, "ppb_attn_fwd")
    _lib.count_launch()
    return out, lse


def _attn_scratch(device: torch.device, B: int, H: int, Sq: int, dh: int) -> torch.Tensor:
    lib = _lib_required()
    need = int(lib.ppb_attn_bwd_scratch_bytes(B, H, Sq, dh))
    key = (device.index, torch.cuda.current_stream().cuda_stream, "attn")
    if torch.cuda.is_current_stream_capturing():
        return torch.empty(need, dtype=torch.uint8, device=device)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < need:
        buf = torch.empty(need, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def attention_bwd(q, k, v, o, do, lse, B: int, H: int, causal: bool = False, scale: Optional[float] = None,
                  dropout_p: float = 0.0, seed: int = 0, out=None):
    """Flash attention backward -> (dq, dk, dv) with the layouts of q, k, v (``out`` = optional views to fill)."""
    if not _is_cuda(q):
        dq, dk, dv = ref.attention_bwd(q, k, v, o, do, lse, B, H, causal, scale, dropout_p, seed)
        if out is not None:
            out[0].copy_(dq)
            out[1].copy_(dk)
            out[2].copy_(dv)
            return out
        return dq, dk, dv
    lib = _lib_required()
    Tq, D = q.shape
    dh = D // H
    Sq, Skv = Tq // B, k.shape[0] // B
    scale = scale if scale is not None else 1.0 / math.sqrt(dh)
    if out is None:
        # the kernel addresses dq/dk/dv with the row strides of q/k/v
        dq = torch.empty((Tq, q.stride(0)), dtype=q.dtype, device=q.device)[:, :D]
        dk = torch.empty((k.shape[0], k.stride(0)), dtype=q.dtype, device=q.device)[:, :D]
        dv = torch.empty((k.shape[0], v.stride(0)), dtype=q.dtype, device=q.device)[:, :D]
    else:
        dq, dk, dv = out
    if dq.stride(0) != q.stride(0) or dk.stride(0) != k.stride(0) or dv.stride(0) != v.stride(0) or do.stride(0) != o.stride(0):
        raise ValueError("attention_bwd: gradient buffers must share the row stride of their forward tensors")
    scratch = _attn_scratch(q.device, B, H, max(Sq, Skv), dh)
    rc = lib.ppb_attn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(),
                          dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), scratch.data_ptr(), 0, B, H, Sq, Skv, dh,
                          q.stride(0), k.stride(0), v.stride(0), o.stride(0), int(causal), 0, 0, float(scale),
                          float(dropout_p), int(seed) & (2**64 - 1), _stream())
    _lib.check(rc, "ppb_attn_bwd")
    _lib.count_launch(3)
    return dq, dk, dv


# ------------------------------------------------------------------------------------------------------------
# Llama-style pieces: SwiGLU, rotary position embedding (csrc/llama_ops.cu)
# ------------------------------------------------------------------------------------------------------------
def swiglu_fwd(gu: torch.Tensor) -> torch.Tensor:
    """h [T, f] = silu(gu[:, :f]) * gu[:, f:]  (gu = packed gate/up output of linear1)."""
    _check_2d(gu, "gu")
    if not _is_cuda(gu):
        return ref.swiglu_fwd(gu)
    T, f = gu.shape[0], gu.shape[1] // 2
    h = torch.empty((T, f), dtype=gu.dtype, device=gu.device)
    lib = _lib_required()
    _lib.check(lib.ppb_swiglu_fwd(gu.data_ptr(), gu.stride(0), h.data_ptr(), h.stride(0), T, f, _stream()), "ppb_swiglu_fwd")
    _lib.count_launch()
    return h


def swiglu_bwd(dh: torch.Tensor, gu: torch.Tensor) -> torch.Tensor:
    """d gu [T, 2f] from d h [T, f] and the stashed pre-activation gu."""
    _check_2d(dh, "dh")
    _check_2d(gu, "gu")
    if not _is_cuda(gu):
        return ref.swiglu_bwd(dh, gu)
    T, f = dh.shape
    dgu = torch.empty((T, 2 * f), dtype=gu.dtype, device=gu.device)
    lib = _lib_required()
    _lib.check(lib.ppb_swiglu_bwd(dh.data_ptr(), dh.stride(0), gu.data_ptr(), gu.stride(0), dgu.data_ptr(), dgu.stride(0),
                                  T, f, _stream()), "ppb_swiglu_bwd")
    _lib.count_launch()
    return dgu


_rope_tables: dict = {}


def rope_table(S: int, dh: int, theta: float, device: torch.device) -> torch.Tensor:
    key = (S, dh, float(theta), str(device))
    t = _rope_tables.get(key)
    if t is None:
        t = ref.rope_table(S, dh, theta, device)
        _rope_tables[key] = t
    return t


def rope_(x: torch.Tensor, col0: int, S: int, n_heads: int, dh: int, theta: float = 10000.0, inverse: bool = False) -> torch.Tensor:
    """Rotary position embedding, in place, on ``n_heads`` heads of ``dh`` columns starting at column ``col0`` of the
    packed activation x [T, *]; position = row % S.  ``inverse`` applies the transpose rotation (the backward)."""
    _check_2d(x, "x")
    table = rope_table(S, dh, theta, x.device)
    if not _is_cuda(x):
        return ref.rope_(x, col0, S, n_heads, dh, table, inverse)
    lib = _lib_required()
    _lib.check(lib.ppb_rope_inplace(x.data_ptr(), x.stride(0), col0, x.shape[0], S, n_heads, dh, table.data_ptr(),
                                    int(inverse), _stream()), "ppb_rope_inplace")
    _lib.count_launch()
    return x
