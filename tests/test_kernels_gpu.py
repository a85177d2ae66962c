"""Numerics of every sm_100a kernel against the plain PyTorch fp32 reference of the same op (SURVEY.md section 4,
"Kernel unit tests").  Shapes are the real shapes of SURVEY.md section 8 plus ragged edge cases."""
import math

import pytest
import torch

from ppb200.ops import functional as F
from ppb200.ops import reference as R
from ppb200.ops import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def test_library_loaded():
    assert _lib.load(required=True) is not None


@pytest.mark.parametrize("M,N,K", [(1024, 2304, 768), (4096, 1024, 1024), (300, 200, 136), (128, 64, 64), (1024, 10000, 768)])
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_gemm_tn(M, N, K, bn):
    torch.manual_seed(0)
    a = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
    b = torch.randn(N, K, device=DEV, dtype=torch.bfloat16)
    ldc = F.pad8(N)
    cbuf = torch.zeros(M, ldc, device=DEV, dtype=torch.bfloat16)
    c = cbuf[:, :N]
    F.gemm(a, b, c, M, N, K, False, False, force_bn=bn)
    ref = a.float() @ b.float().t()
    assert rel_err(c, ref) < 5e-3
    if ldc > N:
        assert cbuf[:, N:].abs().max().item() == 0  # TMA store clipped at N


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(512, 768, 1024), (200, 136, 264)])
def test_gemm_majors(a_mn, b_mn, M, N, K):
    torch.manual_seed(1)
    a = torch.randn((K, M) if a_mn else (M, K), device=DEV, dtype=torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device=DEV, dtype=torch.bfloat16)
    c = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    F.gemm(a, b, c, M, N, K, a_mn, b_mn)
    af = a.float().t() if a_mn else a.float()
    bf = b.float().t() if b_mn else b.float()
    assert rel_err(c, af @ bf.t()) < 5e-3


@pytest.mark.parametrize("act,p", [(None, 0.0), ("relu", 0.0), ("gelu", 0.0), ("relu", 0.1)])
def test_linear_fwd_epilogues(act, p):
    torch.manual_seed(2)
    T, K, N = 1024, 768, 2048
    x = torch.randn(T, K, device=DEV, dtype=torch.bfloat16)
    w = torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.05
    b = torch.randn(N, device=DEV, dtype=torch.bfloat16)
    y = F.linear_fwd(x, w, b, act, p, seed=1234)
    ref = R.linear_fwd(x, w, b, act, p, seed=1234)
    assert rel_err(y, ref) < 1e-2
    if p > 0:
        frac = (y == 0).float().mean().item()
        assert 0.5 < frac < 0.6  # relu zeros ~50% + dropout 10% of the rest


def test_linear_dgrad_wgrad():
    torch.manual_seed(3)
    T, Nin, Nout = 1024, 768, 2048
    x = torch.randn(T, Nin, device=DEV, dtype=torch.bfloat16)
    w = torch.randn(Nout, Nin, device=DEV, dtype=torch.bfloat16) * 0.05
    dy = torch.randn(T, Nout, device=DEV, dtype=torch.bfloat16)
    h = torch.relu(torch.randn(T, Nin, device=DEV)).to(torch.bfloat16)
    dx = F.linear_dgrad(dy, w)
    assert rel_err(dx, dy.float() @ w.float()) < 5e-3
    dxm = F.linear_dgrad(dy, w, mask_pos_aux=h, alpha=1.0 / 0.9)
    assert rel_err(dxm, R.linear_dgrad(dy, w, mask_pos_aux=h, alpha=1.0 / 0.9)) < 5e-3
    dxg = F.linear_dgrad(dy, w, gelu_aux=h)
    assert rel_err(dxg, R.linear_dgrad(dy, w, gelu_aux=h)) < 5e-3
    dw = torch.zeros(Nout, Nin, device=DEV)
    db = torch.zeros(Nout, device=DEV)
    F.linear_wgrad(dy, x, dw, db)
    F.linear_wgrad(dy, x, dw, db)  # accumulates
    assert rel_err(dw, 2 * (dy.float().t() @ x.float())) < 5e-3
    assert rel_err(db, 2 * dy.float().sum(0)) < 5e-3


@pytest.mark.parametrize("M,N,K,splits", [(1024, 1024, 4096, 0), (768, 768, 1024, 3), (2048, 768, 8192, 0)])
def test_wgrad_splitk(M, N, K, splits):
    torch.manual_seed(4)
    dy = torch.randn(K, M, device=DEV, dtype=torch.bfloat16)
    x = torch.randn(K, N, device=DEV, dtype=torch.bfloat16)
    dw = torch.ones(M, N, device=DEV)
    F.gemm(dy, x, dw, M, N, K, True, True, F.OUT_F32_ACCUM, force_splits=splits)
    assert rel_err(dw, 1 + dy.float().t() @ x.float()) < 5e-3


@pytest.mark.parametrize("T,d", [(1024, 768), (4096, 1024), (333, 1600), (64, 256), (512, 4096)])
@pytest.mark.parametrize("p", [0.0, 0.1])
@pytest.mark.parametrize("rms", [False, True])
def test_add_dropout_norm(T, d, p, rms):
    torch.manual_seed(5)
    resid = torch.randn(T, d, device=DEV, dtype=torch.bfloat16)
    a = torch.randn(T, d, device=DEV, dtype=torch.bfloat16)
    g = (1 + 0.1 * torch.randn(d, device=DEV)).to(torch.bfloat16)
    b = None if rms else (0.1 * torch.randn(d, device=DEV)).to(torch.bfloat16)
    z, y, mean, rstd = F.add_dropout_norm_fwd(resid, a, g, b, 1e-5, p, seed=77, rms=rms)
    zr, yr, meanr, rstdr = R.add_dropout_norm_fwd(resid, a, g, b, 1e-5, p, seed=77, rms=rms)
    assert rel_err(z, zr) < 1e-3
    assert rel_err(y, yr) < 1e-2
    assert rel_err(rstd, rstdr) < 1e-3
    dys = [torch.randn(T, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    dg, dbt = torch.zeros(d, device=DEV), (None if rms else torch.zeros(d, device=DEV))
    dgr, dbr = torch.zeros(d, device=DEV), (None if rms else torch.zeros(d, device=DEV))
    for _ in range(2):  # twice: accumulators add up and the scratch self-resets
        dz, da = F.norm_bwd(dys, z, mean, rstd, g, dg, dbt, p, seed=77, rms=rms)
        dzr, dar = R.norm_bwd(dys, zr, meanr, rstdr, g, dgr, dbr, p, seed=77, rms=rms)
    assert rel_err(dz, dzr) < 1e-2
    assert rel_err(da, dar) < 1e-2
    assert rel_err(dg, dgr) < 1e-2
    if not rms:
        assert rel_err(dbt, dbr) < 1e-2


@pytest.mark.parametrize("T,V", [(1024, 10000), (512, 50257), (64, 32000), (8, 120000)])
def test_xent(T, V):
    torch.manual_seed(6)
    ld = F.pad8(V)
    buf = torch.zeros(T, ld, device=DEV, dtype=torch.bfloat16)
    logits = buf[:, :V]
    logits.copy_(torch.randn(T, V, device=DEV) * 3)
    tgt = torch.randint(0, V, (T,), device=DEV)
    lref = logits.clone()
    loss_ref = R.xent_fwd_bwd(lref, tgt, 0.5)
    loss = F.xent_fwd_bwd(logits, tgt, 0.5)
    assert rel_err(loss, loss_ref) < 1e-3
    assert rel_err(logits, lref) < 2e-2
    out = torch.zeros(1, device=DEV)
    F.reduce_sum(loss, out, 1.0 / T)
    assert abs(out.item() - loss_ref.mean().item()) < 1e-3 * abs(loss_ref.mean().item())


def test_adamw():
    torch.manual_seed(7)
    n = 1_000_003
    master = torch.randn(n, device=DEV)
    grad = torch.randn(n, device=DEV)
    m, v = torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
    param = master.to(torch.bfloat16)
    mr, gr, mmr, vr = master.clone(), grad.clone(), m.clone(), v.clone()
    pr = param.clone()
    coef = torch.tensor([0.5], device=DEV)
    for step in (1, 2):
        F.adamw_step(param, master, m, v, grad, 1e-3, 0.9, 0.95, 1e-8, 0.1, step, 0.25, coef, zero_grad=False)
        R.adamw_step(pr, mr, mmr, vr, gr, 1e-3, 0.9, 0.95, 1e-8, 0.1, step, 0.25, coef, zero_grad=False)
    assert rel_err(master, mr) < 1e-5
    assert rel_err(param, pr) < 1e-3
    assert rel_err(v, vr) < 1e-5
    F.adamw_step(param, master, m, v, grad, 1e-3, 0.9, 0.95, 1e-8, 0.1, 3, 1.0, None, zero_grad=True)
    assert grad.abs().max().item() == 0
    ss = torch.zeros(1, device=DEV)
    F.sumsq_accum(gr, ss)
    assert abs(ss.item() - (gr ** 2).sum().item()) < 1e-3 * ss.item()
    c = torch.zeros(1, device=DEV)
    F.clip_coef(ss, 1.0, 1.0, c)
    assert abs(c.item() - min(1.0, 1.0 / math.sqrt(ss.item()))) < 1e-5


def test_embedding_and_small_ops():
    torch.manual_seed(8)
    V, d, T = 10000, 768, 1024
    w = torch.randn(V, d, device=DEV, dtype=torch.bfloat16)
    idx = torch.randint(0, V, (8, 128), device=DEV)
    out = F.embedding_fwd(idx, w)
    assert torch.equal(out, w[idx.reshape(-1)])
    dys = [torch.randn(T, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    dw, dwr = torch.zeros(V, d, device=DEV), torch.zeros(V, d, device=DEV)
    F.embedding_bwd(idx, dys, dw)
    R.embedding_bwd(idx, dys, dwr)
    assert rel_err(dw, dwr) < 1e-5
    s = F.add_n(dys)
    assert rel_err(s, R.add_n(dys)) < 1e-2
    g = F.gelu_fwd(dys[0])
    assert rel_err(g, torch.nn.functional.gelu(dys[0].float())) < 1e-2


ATTN_CASES = [
    # B, H, Sq, Skv, dh, causal, p
    (2, 4, 256, 256, 64, False, 0.0),
    (2, 4, 256, 256, 64, True, 0.0),
    (2, 4, 256, 256, 64, False, 0.1),
    (8, 8, 128, 128, 96, False, 0.1),
    (1, 2, 1024, 1024, 64, True, 0.1),
    (2, 2, 384, 384, 128, True, 0.0),
    (2, 3, 200, 200, 64, False, 0.0),
    (2, 3, 136, 328, 64, False, 0.0),
    (1, 4, 128, 128, 32, False, 0.0),
    (2, 4, 128, 128, 192, False, 0.1),   # the reference sweep's 4-head case: wide-head tcgen05 kernels (two-part backward)
    (1, 2, 96, 160, 192, False, 0.0),
    (1, 2, 384, 384, 192, True, 0.1),    # several query / kv tiles per head, causal, dropout
    (1, 3, 200, 328, 160, False, 0.0),   # 129..192 zero-padded to three 64-column chunks, ragged tiles
    (1, 2, 128, 128, 256, True, 0.0),    # CUDA-core path (head dims above 192)
]


def _attn_inputs(B, H, Sq, Skv, dh):
    torch.manual_seed(9)
    D = H * dh
    qbuf = torch.randn(B * Sq, 3 * D, device=DEV, dtype=torch.bfloat16)
    kvbuf = torch.randn(B * Skv, 3 * D, device=DEV, dtype=torch.bfloat16)
    return qbuf[:, :D], kvbuf[:, D:2 * D], kvbuf[:, 2 * D:]


@pytest.mark.parametrize("B,H,Sq,Skv,dh,causal,p", ATTN_CASES)
def test_attention_fwd(B, H, Sq, Skv, dh, causal, p):
    q, k, v = _attn_inputs(B, H, Sq, Skv, dh)
    o, lse = F.attention_fwd(q, k, v, B, H, causal, None, p, seed=42)
    oref, lref = R.attention_fwd(q, k, v, B, H, causal, None, p, seed=42)
    assert rel_err(lse, lref) < 1e-3
    assert rel_err(o, oref) < 2e-2


@pytest.mark.parametrize("B,H,Sq,Skv,dh,causal,p", ATTN_CASES)
def test_attention_bwd(B, H, Sq, Skv, dh, causal, p):
    q, k, v = _attn_inputs(B, H, Sq, Skv, dh)
    o, lse = F.attention_fwd(q, k, v, B, H, causal, None, p, seed=42)
    torch.manual_seed(10)
    do = torch.randn_like(o)
    dq, dk, dv = F.attention_bwd(q, k, v, o, do, lse, B, H, causal, None, p, seed=42)
    dqr, dkr, dvr = R.attention_bwd(q, k, v, o, do, lse, B, H, causal, None, p, seed=42)
    assert rel_err(dv, dvr) < 2e-2
    assert rel_err(dk, dkr) < 2e-2
    assert rel_err(dq, dqr) < 2e-2
    # second call re-uses (and re-zeroes) the scratch accumulator
    dq2, _, _ = F.attention_bwd(q, k, v, o, do, lse, B, H, causal, None, p, seed=42)
    assert rel_err(dq2, dqr) < 2e-2


@pytest.mark.parametrize("M,N,K,n_dst,splits", [(1024, 1024, 4096, 1, 0), (768, 2304, 1024, 2, 0), (2048, 1024, 4096, 3, 2), (200, 136, 520, 2, 0)])
def test_gemm_allreduce_push_single_gpu(M, N, K, n_dst, splits):
    """K10 numerics on one GPU: the 'replicas' are n_dst separate buffers of the same device."""
    from ppb200.ops.collective import gemm_allreduce, push_add
    torch.manual_seed(12)
    dy = torch.randn(K, M, device=DEV, dtype=torch.bfloat16)
    x = torch.randn(K, N, device=DEV, dtype=torch.bfloat16)
    prev = torch.randn(M, N, device=DEV)
    accum = prev.clone()
    dsts = [torch.full((M, N), float(i + 1), device=DEV) for i in range(n_dst)]
    for rep in range(2):  # twice: the arrival counters must have reset themselves
        gemm_allreduce(dy, x, accum, [d.data_ptr() for d in dsts], force_splits=splits)
    prod = dy.float().t() @ x.float()
    assert rel_err(accum, prev + 2 * prod) < 5e-3
    for i, d in enumerate(dsts):   # push 1 saw prev + prod, push 2 saw prev + 2 prod
        assert rel_err(d - (i + 1), 2 * prev + 3 * prod) < 5e-3
    src = torch.randn(4096, device=DEV)
    outs = [torch.ones(4096, device=DEV) for _ in range(n_dst)]
    push_add(src, [o.data_ptr() for o in outs], 0.5)
    for o in outs:
        assert rel_err(o, 1 + 0.5 * src) < 1e-6


@pytest.mark.parametrize("tp,rows,N,K,act", [(2, 256, 512, 256, None), (4, 128, 2048, 1024, "relu"), (2, 512, 200, 136, "gelu")])
def test_allgather_gemm_single_gpu(tp, rows, N, K, act):
    """K12 numerics with the tp shards living in separate buffers of one GPU."""
    from ppb200.ops.collective import allgather_linear
    torch.manual_seed(13)
    shards = [torch.randn(rows, K, device=DEV, dtype=torch.bfloat16) for _ in range(tp)]
    w = torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.05
    b = torch.randn(N, device=DEV, dtype=torch.bfloat16)
    for rank in range(tp):
        y = allgather_linear([s.data_ptr() for s in shards], rows, K, K, w, rank, b, act)
        ref = R.linear_fwd(torch.cat(shards), w, b, act)
        assert rel_err(y, ref) < 1e-2


@pytest.mark.parametrize("tp,rows,N,K,act", [(2, 256, 512, 256, None), (4, 128, 2048, 1024, "relu"), (2, 1024, 1000, 512, "gelu"),
                                             (8, 512, 3072, 1024, None)])
def test_fused_allgather_gemm_single_gpu(tp, rows, N, K, act):
    """K12, fused form (copier warps + panel counters inside the GEMM kernel), with the tp shards in separate
    buffers of one GPU: every 'rank' must produce the full product AND leave the gathered activations behind; the
    counters are never reset between calls."""
    from ppb200.ops.collective import GatherFlags, allgather_linear_fused
    torch.manual_seed(15)
    w = torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.05
    b = torch.randn(N, device=DEV, dtype=torch.bfloat16)
    flags = [GatherFlags(rows * tp, torch.device(DEV)) for _ in range(tp)]
    for it in range(3):   # fresh inputs each round: stale panels from the previous call would show up as an error
        shards = [torch.randn(rows, K, device=DEV, dtype=torch.bfloat16) for _ in range(tp)]
        ref = R.linear_fwd(torch.cat(shards), w, b, act)
        for rank in range(tp):
            full = torch.full((rows * tp, K), float("nan"), device=DEV, dtype=torch.bfloat16)
            y = allgather_linear_fused([s.data_ptr() for s in shards], rows, K, full, flags[rank], w, rank, b, act)
            assert rel_err(y, ref) < 1e-2, (it, rank)
            assert torch.equal(full, torch.cat(shards)), (it, rank)


def test_tp_comm_and_fused_dp_wiring_world_size_one():
    """The Python side of the fused compute+collective paths (symmetric-memory rendezvous, TPComm buffers and
    barriers, FusedGradAllReduce arming / coverage bookkeeping / finish) driven end to end on ONE GPU with a
    single-rank process group, so the driver's 1-GPU test run covers the wiring and not only the raw kernels."""
    import os
    import torch.distributed as dist
    from ppb200.ops.collective import FusedGradAllReduce, TPComm
    created = False
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", str(29700 + os.getpid() % 200))
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
        created = True
    try:
        torch.manual_seed(16)
        dev = torch.device("cuda", 0)
        rows, d, f = 256, 512, 1024
        tp = TPComm(dist.group.WORLD, rows, d, dev)
        x = torch.randn(rows, d, device=dev, dtype=torch.bfloat16)
        w1 = torch.randn(f, d, device=dev, dtype=torch.bfloat16) * 0.05
        b1 = torch.randn(f, device=dev, dtype=torch.bfloat16)
        w2 = torch.randn(d, f, device=dev, dtype=torch.bfloat16) * 0.05
        for _ in range(3):    # the double-buffered symmetric buffers and the panel counters cycle
            h, full = tp.ag_linear(x, w1, b1, "relu")
            assert rel_err(h, R.linear_fwd(x, w1, b1, "relu")) < 1e-2 and torch.equal(full, x)
            y = tp.linear_rs(h, w2)
            assert rel_err(y, h.float() @ w2.float().t()) < 5e-3
        # fused DP gradient reduction with one replica: final = sum over replicas = the local gradient
        n = 512 * 256 + 512
        grad = torch.zeros(n, device=dev)
        ar = FusedGradAllReduce(grad, n, dist.group.WORLD, 1.0)
        dy = torch.randn(1024, 512, device=dev, dtype=torch.bfloat16)
        xx = torch.randn(1024, 256, device=dev, dtype=torch.bfloat16)
        gw, gb = grad[:512 * 256].view(512, 256), grad[512 * 256:]
        gw.add_(1.0)
        gb.add_(2.0)
        ar.begin_last_backward()
        F.linear_wgrad(dy, xx, gw, None, ar=ar)
        ar.finish()
        torch.cuda.synchronize()
        want = 1.0 + dy.float().t() @ xx.float()
        assert rel_err(ar.final[:512 * 256].view(512, 256), want) < 5e-3
        assert rel_err(ar.final[512 * 256:n], torch.full((512,), 2.0, device=dev)) < 1e-6
        ar.after_optimizer()
    finally:
        if created:
            dist.destroy_process_group()


@pytest.mark.parametrize("tp,M,N,K", [(2, 512, 768, 256), (4, 1024, 1024, 512), (2, 192, 200, 136)])
def test_gemm_reducescatter_single_gpu(tp, M, N, K):
    """K11 numerics: tp 'ranks' with different K-slices push into per-owner buffers on one GPU."""
    from ppb200.ops.collective import linear_reducescatter
    torch.manual_seed(14)
    rows = M // tp
    dsts = [torch.zeros(rows, N, device=DEV) for _ in range(tp)]
    total = torch.zeros(M, N, device=DEV)
    for r in range(tp):
        x = torch.randn(M, K, device=DEV, dtype=torch.bfloat16)
        w = torch.randn(N, K, device=DEV, dtype=torch.bfloat16) * 0.05
        linear_reducescatter(x, w, [d.data_ptr() for d in dsts], N)
        total += x.float() @ w.float().t()
    got = torch.cat(dsts)
    assert rel_err(got, total) < 5e-3


# ---- Llama-style pieces (csrc/llama_ops.cu) --------------------------------------------------------------------
@pytest.mark.parametrize("T,f", [(256, 512), (1000, 1376)])
def test_swiglu_fwd_bwd_matches_fp32_reference(T, f):
    from ppb200.ops import functional as F, reference as ref
    torch.manual_seed(0)
    gu = torch.randn(T, 2 * f, device="cuda").bfloat16()
    dh = torch.randn(T, f, device="cuda").bfloat16()
    h = F.swiglu_fwd(gu)
    want = ref.swiglu_fwd(gu.float().cpu())
    assert (h.float().cpu() - want).abs().max().item() < 2e-2 * want.abs().max().item()
    d = F.swiglu_bwd(dh, gu)
    want = ref.swiglu_bwd(dh.float().cpu(), gu.float().cpu())
    assert (d.float().cpu() - want).abs().max().item() < 2e-2 * want.abs().max().item()


@pytest.mark.parametrize("S,H,dh", [(128, 4, 64), (96, 2, 128), (64, 3, 16)])
def test_rope_inplace_matches_fp32_reference(S, H, dh):
    from ppb200.ops import functional as F, reference as ref
    torch.manual_seed(1)
    B, d = 3, H * dh
    x = torch.randn(B * S, 3 * d, device="cuda").bfloat16()
    want = ref.rope_(x.float().cpu().clone(), 0, S, 2 * H, dh, ref.rope_table(S, dh))
    y = x.clone()
    F.rope_(y, 0, S, 2 * H, dh)
    assert torch.equal(y[:, 2 * d:], x[:, 2 * d:])                        # v columns untouched
    assert (y.float().cpu() - want).abs().max().item() < 2e-2 * want.abs().max().item()
    F.rope_(y, 0, S, 2 * H, dh, inverse=True)
    assert (y.float() - x.float()).abs().max().item() < 4e-2 * x.float().abs().max().item()
