"""Plain PyTorch (fp32-math) implementations of every op in ``ops.functional``.

Two jobs: (1) the CPU execution path (gloo plumbing runs, BASELINE.json config 1) and (2) the numerics oracle
for the CUDA kernels (tests/ compare each kernel against these).  The dropout mask generator reproduces
``ppb::dropout_keep`` (csrc/common.cuh) bit-for-bit with int64 arithmetic, so CPU and GPU runs with the same
seed drop the same elements.
"""
from __future__ import annotations

import math

import torch

_M32 = 0xFFFFFFFF


def _mix32(x: torch.Tensor) -> torch.Tensor:
    x = x & _M32
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    x = x ^ (x >> 16)
    return x


def _mix32_int(x: int) -> int:
    x &= _M32
    x ^= x >> 16
    x = (x * 0x7FEB352D) & _M32
    x ^= x >> 15
    x = (x * 0x846CA68B) & _M32
    x ^= x >> 16
    return x


def dropout_thresh13(p: float) -> int:
    t = p * 8192.0 + 0.5
    return 0 if t <= 0 else (8191 if t >= 8191 else int(t))


def dropout_thr_pattern(p: float) -> int:
    """Threshold as the 16-bit pattern the kernels compare against (csrc/common.cuh: bit 14 set, bits 15 and 10 clear)."""
    t = dropout_thresh13(p)
    return 0x4000 | ((t >> 10) << 11) | (t & 0x3FF)


def dropout_keep_mask(seed: int, start: int, numel: int, p: float, device=None) -> torch.Tensor:
    """keep[i] for flat element indices start .. start+numel-1 (bool); bit-identical to ppb::dropout_keep:
    one hash per pair of elements, x = (lo + (k0 ^ mix32(hi ^ k1 ^ golden))) * C1, x ^= x >> 15, y = x * C2; the even
    element takes the low 16 bits of y, the odd one the high 16 bits; 13 bits of each half are compared with the
    threshold (effective dropout probability round(p * 8192) / 8192)."""
    thr = dropout_thr_pattern(p)
    if numel <= 0:
        return torch.zeros(0, dtype=torch.bool, device=device)
    p0, p1 = start >> 1, (start + numel - 1) >> 1
    pair = torch.arange(p0, p1 + 1, dtype=torch.int64, device=device)      # one hash per PAIR, like the kernels
    k0, k1 = seed & _M32, (seed >> 32) & _M32
    if (p0 >> 32) == (p1 >> 32):      # the usual case: the high word of the pair index is one constant -> scalar key
        key = k0 ^ _mix32_int(((p0 >> 32) & _M32) ^ ((k1 ^ 0x9E3779B9) & _M32))
    else:
        key = k0 ^ _mix32(((pair >> 32) & _M32) ^ ((k1 ^ 0x9E3779B9) & _M32))
    x = ((((pair & _M32) + key) & _M32) * 0x9E3779B1) & _M32
    x = x ^ (x >> 15)
    y = (x * 0x85EBCA77) & _M32
    halves = torch.stack((y & 0xFFFF, y >> 16), dim=1).reshape(-1)          # element order: even, odd, even, odd, ...
    off = start & 1
    half = halves[off:off + numel]
    return ((half & 0x3BFF) | 0x4000) >= thr


def dropout(x: torch.Tensor, p: float, seed: int, index_offset: int = 0) -> torch.Tensor:
    if p <= 0.0:
        return x
    keep = dropout_keep_mask(seed, index_offset, x.numel(), p, x.device).view(x.shape)
    return torch.where(keep, x * (1.0 / (1.0 - p)), torch.zeros_like(x))


def _act(x, act):
    if act is None:
        return x
    if act == "relu":
        return torch.relu(x)
    if act == "gelu":
        return torch.nn.functional.gelu(x)
    raise ValueError(act)


def linear_fwd(x, w, bias=None, act=None, dropout_p=0.0, seed=0):
    y = x.float() @ w.float().t()
    if bias is not None:
        y = y + bias.float()
    y = _act(y, act)
    if dropout_p > 0.0:
        y = dropout(y, dropout_p, seed)
    return y.to(x.dtype)


def linear_dgrad(dy, w, mask_pos_aux=None, gelu_aux=None, alpha=1.0):
    dx = (dy.float() @ w.float()) * alpha
    if mask_pos_aux is not None:
        dx = torch.where(mask_pos_aux.float() > 0, dx, torch.zeros_like(dx))
    if gelu_aux is not None:
        a = gelu_aux.float()
        cdf = 0.5 * (1.0 + torch.erf(a * 0.70710678118))
        pdf = 0.3989422804 * torch.exp(-0.5 * a * a)
        dx = dx * (cdf + a * pdf)
    return dx.to(dy.dtype)


def linear_wgrad(dy, x, dw_accum, db_accum=None):
    dw_accum.add_((dy.float().t() @ x.float()).to(dw_accum.dtype))
    if db_accum is not None:
        db_accum.add_(dy.float().sum(0).to(db_accum.dtype))


def add_dropout_norm_fwd(resid, a, gamma, beta, eps=1e-5, dropout_p=0.0, seed=0, rms=False):
    ref = resid if resid is not None else a
    z = torch.zeros_like(ref, dtype=torch.float32)
    if resid is not None:
        z = z + resid.float()
    if a is not None:
        z = z + (dropout(a.float(), dropout_p, seed) if dropout_p > 0 else a.float())
    z = z.to(ref.dtype)  # statistics on the storage-rounded z, like the kernel
    zf = z.float()
    if rms:
        mean = torch.zeros(zf.shape[0], device=zf.device)
        var = (zf * zf).mean(-1)
    else:
        mean = zf.mean(-1)
        var = ((zf - mean[:, None]) ** 2).mean(-1)
    rstd = torch.rsqrt(var + eps)
    y = (zf - mean[:, None]) * rstd[:, None] * gamma.float()
    if beta is not None:
        y = y + beta.float()
    return z, y.to(ref.dtype), mean, rstd


def norm_bwd(dys, z, mean, rstd, gamma, dgamma, dbeta, dropout_p=0.0, seed=0, rms=False, need_da=True):
    dy = sum(d.float() for d in dys if d is not None)
    zf = z.float()
    m = torch.zeros_like(rstd) if rms else mean
    xh = (zf - m[:, None]) * rstd[:, None]
    g = dy * gamma.float()
    c1 = torch.zeros_like(rstd) if rms else g.mean(-1)
    c2 = (g * xh).mean(-1)
    dz = rstd[:, None] * (g - c1[:, None] - xh * c2[:, None])
    dgamma.add_((dy * xh).sum(0).to(dgamma.dtype))
    if dbeta is not None:
        dbeta.add_(dy.sum(0).to(dbeta.dtype))
    dz = dz.to(z.dtype)
    da = None
    if need_da:
        da = dropout(dz.float(), dropout_p, seed).to(z.dtype) if dropout_p > 0 else dz
    return dz, da


def xent_fwd_bwd(logits, targets, grad_scale=1.0, write_grad=True):
    """logits [T, V] (possibly a strided view); returns per-row loss, overwrites logits with the gradient."""
    lf = logits.float()
    lse = torch.logsumexp(lf, dim=-1)
    valid = (targets >= 0) & (targets < lf.shape[1])
    tg = targets.clamp(0, lf.shape[1] - 1)
    lt = lf.gather(1, tg[:, None]).squeeze(1)
    loss = torch.where(valid, lse - lt, torch.zeros_like(lse))
    if write_grad:
        g = torch.softmax(lf, dim=-1)
        g[torch.arange(lf.shape[0], device=lf.device), tg] -= 1.0
        g = g * grad_scale * valid[:, None].float()
        logits.copy_(g.to(logits.dtype))
    return loss


def adamw_step(param, master, m, v, grad, lr, beta1, beta2, eps, weight_decay, step, grad_scale=1.0,
               clip_coef=None, zero_grad=True):
    g = grad * (grad_scale * (float(clip_coef) if clip_coef is not None else 1.0))
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1.0 / (1.0 - beta1 ** step)
    bc2 = 1.0 / (1.0 - beta2 ** step)
    upd = (m * bc1) / ((v * bc2).sqrt() + eps) + weight_decay * master
    master.add_(upd, alpha=-lr)
    if param is not None and param.data_ptr() != master.data_ptr():
        param.copy_(master.to(param.dtype))
    if zero_grad:
        grad.zero_()


def embedding_fwd(idx, w):
    return w[idx.reshape(-1).clamp(0, w.shape[0] - 1)]


def embedding_bwd(idx, dys, dw_accum):
    dy = sum(d.float() for d in dys if d is not None)
    dw_accum.index_add_(0, idx.reshape(-1), dy.to(dw_accum.dtype))


def add_n(ts):
    ts = [t for t in ts if t is not None]
    out = ts[0].float()
    for t in ts[1:]:
        out = out + t.float()
    return out.to(ts[0].dtype)


def attention_fwd(q, k, v, B, H, causal=False, scale=None, dropout_p=0.0, seed=0):
    """q [B*Sq, H*dh] (strided views allowed), k/v [B*Skv, H*dh] -> o [B*Sq, H*dh], lse [B, H, Sq] (natural log).

    Dropout acts on the normalised probabilities; element index = ((b*H + h)*Sq + i)*Skv + j."""
    Tq, D = q.shape
    dh = D // H
    Sq, Skv = Tq // B, k.shape[0] // B
    scale = scale if scale is not None else 1.0 / math.sqrt(dh)
    qf = q.float().view(B, Sq, H, dh).permute(0, 2, 1, 3)
    kf = k.float().view(B, Skv, H, dh).permute(0, 2, 1, 3)
    vf = v.float().view(B, Skv, H, dh).permute(0, 2, 1, 3)
    s = (qf @ kf.transpose(-1, -2)) * scale
    if causal:
        mask = torch.ones(Sq, Skv, dtype=torch.bool, device=q.device).tril(Skv - Sq)
        s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse[..., None])
    if dropout_p > 0:
        p = dropout(p.contiguous(), dropout_p, seed)
    o = (p @ vf).permute(0, 2, 1, 3).reshape(Tq, D)
    return o.to(q.dtype), lse


def attention_bwd(q, k, v, o, do, lse, B, H, causal=False, scale=None, dropout_p=0.0, seed=0):
    Tq, D = q.shape
    dh = D // H
    Sq, Skv = Tq // B, k.shape[0] // B
    scale = scale if scale is not None else 1.0 / math.sqrt(dh)
    qf = q.float().view(B, Sq, H, dh).permute(0, 2, 1, 3)
    kf = k.float().view(B, Skv, H, dh).permute(0, 2, 1, 3)
    vf = v.float().view(B, Skv, H, dh).permute(0, 2, 1, 3)
    dof = do.float().view(B, Sq, H, dh).permute(0, 2, 1, 3)
    of = o.float().view(B, Sq, H, dh).permute(0, 2, 1, 3)
    s = (qf @ kf.transpose(-1, -2)) * scale
    if causal:
        mask = torch.ones(Sq, Skv, dtype=torch.bool, device=q.device).tril(Skv - Sq)
        s = s.masked_fill(~mask, float("-inf"))
    p = torch.exp(s - lse[..., None])
    if dropout_p > 0:
        keep = dropout_keep_mask(seed, 0, p.numel(), dropout_p, p.device).view(p.shape).float() / (1.0 - dropout_p)
    else:
        keep = torch.ones_like(p)
    pd = p * keep
    dv = pd.transpose(-1, -2) @ dof
    dpd = dof @ vf.transpose(-1, -2)
    dp = dpd * keep
    delta = (dof * of).sum(-1, keepdim=True)  # == sum_j pd_ij * dpd_ij
    ds = p * (dp - delta) * scale
    dq = ds @ kf
    dk = ds.transpose(-1, -2) @ qf
    back = lambda t, S: t.permute(0, 2, 1, 3).reshape(B * S, D).to(q.dtype)
    return back(dq, Sq), back(dk, Skv), back(dv, Skv)


# ---- Llama-style pieces -------------------------------------------------------------------------------------
def swiglu_fwd(gu):
    f = gu.shape[1] // 2
    g, u = gu[:, :f].float(), gu[:, f:].float()
    return (torch.nn.functional.silu(g) * u).to(gu.dtype)


def swiglu_bwd(dh, gu):
    f = gu.shape[1] // 2
    g, u, d = gu[:, :f].float(), gu[:, f:].float(), dh.float()
    s = torch.sigmoid(g)
    return torch.cat([d * u * (s + g * s * (1 - s)), d * g * s], dim=1).to(gu.dtype)


def rope_table(S, dh, theta=10000.0, device="cpu"):
    """[S, dh/2, 2] fp32 (cos, sin) of pos * theta^(-2i/dh)."""
    inv = theta ** (-torch.arange(0, dh, 2, dtype=torch.float64) / dh)
    ang = torch.arange(S, dtype=torch.float64)[:, None] * inv[None, :]
    return torch.stack([ang.cos(), ang.sin()], dim=-1).float().to(device).contiguous()


def rope_(x, col0, S, n_heads, dh, table, inverse=False):
    """In-place rotary embedding on columns [col0, col0 + n_heads*dh) of x [T, *] (rotate-half pairs, position = row % S)."""
    T = x.shape[0]
    v = x[:, col0:col0 + n_heads * dh].float().reshape(T, n_heads, dh)
    pos = torch.arange(T, device=x.device) % S
    cos, sin = table[pos, :, 0][:, None, :], table[pos, :, 1][:, None, :]
    if inverse:
        sin = -sin
    a, b = v[..., :dh // 2], v[..., dh // 2:]
    out = torch.cat([a * cos - b * sin, b * cos + a * sin], dim=-1)
    x[:, col0:col0 + n_heads * dh] = out.reshape(T, n_heads * dh).to(x.dtype)
    return x
