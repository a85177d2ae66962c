"""Fused compute + collective paths over NVLink peer memory.

``FusedGradAllReduce`` implements the data-parallel gradient all-reduce WITHOUT an all-reduce kernel: the
reduction is carried by the epilogue of the last micro-batch's weight-gradient GEMMs
(``ppb_gemm_allreduce_bf16``: tcgen05 tiles whose TMA reduce-adds land in every replica's gradient buffer, the
peers' through peer-mapped symmetric memory) plus one small push kernel for the non-GEMM gradients, bracketed by
two device-side barriers per step.  Symmetric memory allocation / handle exchange is borrowed from
``torch.distributed._symmetric_memory`` (allocator + rendezvous only); every byte that crosses NVLink is moved by
our own kernels.  The NCCL ``all_reduce`` in ``PipelineStage.all_reduce_grads`` is the baseline this replaces.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_longlong, c_void_p
from typing import List, Optional, Tuple

import torch

from . import _lib


def _bind(lib) -> None:
    if getattr(lib, "_ppb_collective_bound", False):
        return
    P, I, L, F = c_void_p, c_int, c_longlong, c_float
    lib.ppb_gemm_allreduce_bf16.argtypes = [P, P, P, P, I, I, I, I, L, L, L, P, L, I, I, P]
    lib.ppb_gemm_allreduce_scratch_bytes.argtypes = [I, I]
    lib.ppb_gemm_allreduce_scratch_bytes.restype = L
    lib.ppb_gemm_allreduce_bf16.restype = I
    lib.ppb_push_add_f32.argtypes = [P, P, I, L, F, P]
    lib.ppb_push_add_f32.restype = I
    lib.ppb_push_add_multimem_f32.argtypes = [P, P, L, F, P]
    lib.ppb_push_add_multimem_f32.restype = I
    lib._ppb_collective_bound = True


_ar_scratch: dict = {}


def gemm_allreduce(dy: torch.Tensor, x: torch.Tensor, accum: torch.Tensor, dst_ptrs: List[int],
                   force_bn: int = 0, force_splits: int = 0) -> None:
    """accum[out, in] += dy^T @ x, and every completed slab of accum is reduce-added into dst_r for every pointer in
    ``dst_ptrs`` (device pointers, possibly peer-mapped) from inside the same kernel.  dy [T, out], x [T, in] bf16;
    accum fp32 [out, in]; the dsts share its leading dimension."""
    lib = _lib.load(required=True)
    _bind(lib)
    T, Nout = dy.shape
    Nin = x.shape[1]
    need = int(lib.ppb_gemm_allreduce_scratch_bytes(Nout, Nin))
    key = (dy.device.index, torch.cuda.current_stream().cuda_stream)
    scratch = _ar_scratch.get(key)
    if scratch is None or scratch.numel() < need:
        scratch = torch.zeros(max(need, 1 << 16), dtype=torch.uint8, device=dy.device)
        _ar_scratch[key] = scratch
    arr = (c_void_p * len(dst_ptrs))(*dst_ptrs)
    rc = lib.ppb_gemm_allreduce_bf16(dy.data_ptr(), x.data_ptr(), accum.data_ptr(), arr, len(dst_ptrs), Nout, Nin, T,
                                     dy.stride(0), x.stride(0), accum.stride(0), scratch.data_ptr(), scratch.numel(),
                                     force_bn, force_splits, torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_gemm_allreduce_bf16")
    _lib.count_launch()


def push_add(src: torch.Tensor, dst_ptrs: List[int], scale: float) -> None:
    """dst_r[i] += scale * src[i] for every pointer in ``dst_ptrs`` (fp32, numel % 4 == 0)."""
    lib = _lib.load(required=True)
    _bind(lib)
    arr = (c_void_p * len(dst_ptrs))(*dst_ptrs)
    rc = lib.ppb_push_add_f32(src.data_ptr(), arr, len(dst_ptrs), src.numel(), float(scale),
                              torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_push_add_f32")
    _lib.count_launch()


def push_add_multimem(src: torch.Tensor, mc_ptr: int, scale: float) -> None:
    """NVLS push: every replica's buffer behind the multicast address ``mc_ptr`` += scale * src (fp32, numel % 4 == 0),
    reduced inside the NVSwitch (``multimem.red``)."""
    lib = _lib.load(required=True)
    _bind(lib)
    rc = lib.ppb_push_add_multimem_f32(src.data_ptr(), mc_ptr, src.numel(), float(scale), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_push_add_multimem_f32")
    _lib.count_launch()


class FusedGradAllReduce:
    """Per-stage state of the fused DP gradient reduction.

    * ``final``: fp32 buffer in symmetric memory, same layout as the stage's gradient arena, that receives the
      already-reduced, already-scaled gradient from every replica (own included);
    * ``begin_last_backward()`` arms the GEMM path; while armed, ``linear_wgrad`` routes through ``wgrad()``;
    * ``finish()`` pushes whatever the GEMMs did not cover and runs the device-side barrier after which
      ``final`` is complete on every replica;
    * ``after_optimizer()`` is the second barrier: nobody starts pushing step k+1 into a buffer that a slower
      replica has not finished consuming/zeroing for step k.
    """

    def __init__(self, grad_arena: torch.Tensor, n_total: int, dp_group, grad_scale: float):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        self.grad = grad_arena
        self.n_total = n_total
        self.grad_scale = grad_scale   # 1/(m*dp): applied by the optimizer kernel, the pushes are unscaled sums
        self.group = dp_group
        dev = grad_arena.device
        n = max(grad_arena.numel(), 8)
        self.final = symm_mem.empty(n, dtype=torch.float32, device=dev)
        self.final.zero_()
        self.hdl = symm_mem.rendezvous(self.final, dp_group)
        self.world = self.hdl.world_size
        self.rank = self.hdl.rank
        self.base_ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        # NVLS (opt-in, PPB_FUSED_DP_NVLS=1): the non-GEMM gradients go through ONE multimem.red stream (in-switch reduction)
        # instead of one push per peer.  Measured standalone at dp = 4 (profiles/fused_paths_n4.json); not the default because
        # a multicast mapping over a SUB-group of a larger job stalled once on this pool.
        import os
        mc = int(getattr(self.hdl, "multicast_ptr", 0) or 0)
        self.mc_ptr = mc if (mc and os.environ.get("PPB_FUSED_DP_NVLS") == "1") else 0
        self.armed = False
        self.covered: List[Tuple[int, int]] = []
        torch.cuda.synchronize(dev)
        dist.barrier(group=dp_group, device_ids=[dev.index])

    def begin_last_backward(self) -> None:
        self.armed = True
        self.covered = []

    def _offset(self, view: torch.Tensor) -> int:
        off = (view.data_ptr() - self.grad.data_ptr()) // 4
        if off < 0 or off + view.numel() > self.grad.numel():
            raise ValueError("gradient view is not inside the stage's gradient arena")
        return off

    def wgrad(self, dy: torch.Tensor, x: torch.Tensor, dw_accum: torch.Tensor) -> None:
        off = self._offset(dw_accum)
        self.covered.append((off, dw_accum.numel()))
        gemm_allreduce(dy, x, dw_accum, [p + 4 * off for p in self.base_ptrs])

    def finish(self) -> None:
        """Push the gradients no GEMM produced (biases, norms, embedding) and synchronise the replicas."""
        self.armed = False
        pos = 0
        for off, n in sorted(self.covered) + [(self.n_total, 0)]:
            if off > pos:
                lo, hi = pos, off
                hi = min(self.n_total, (hi + 3) // 4 * 4)
                if hi > lo:
                    if self.mc_ptr and lo % 4 == 0:
                        push_add_multimem(self.grad[lo:hi], self.mc_ptr + 4 * lo, 1.0)
                    else:
                        push_add(self.grad[lo:hi], [p + 4 * lo for p in self.base_ptrs], 1.0)
            pos = max(pos, off + n)
        self.hdl.barrier(channel=0)

    def after_optimizer(self) -> None:
        self.hdl.barrier(channel=1)


# ------------------------------------------------------------------------------------------------------------
# tensor-parallel fusions (K11 / K12)
# ------------------------------------------------------------------------------------------------------------
def _bind_tp(lib) -> None:
    if getattr(lib, "_ppb_tp_bound", False):
        return
    P, I, L = c_void_p, c_int, c_longlong
    lib.ppb_gemm_allgather_bf16.argtypes = [P, P, P, I, I, I, L, L, L, I, I, I, I, P, P]
    lib.ppb_gemm_allgather_bf16.restype = I
    lib.ppb_gemm_reducescatter_bf16.argtypes = [P, P, P, I, I, I, L, L, L, I, I, P]
    lib.ppb_gemm_reducescatter_bf16.restype = I
    lib.ppb_gemm_allgather_fused_bf16.argtypes = [P, P, P, ctypes.c_uint, P, P, I, I, I, L, L, I, I, I, I, P, P, L, c_float, P]
    lib.ppb_gemm_allgather_fused_bf16.restype = I
    lib.ppb_gemm_allgather_fused_chunks_per_panel.argtypes = [I]
    lib.ppb_gemm_allgather_fused_chunks_per_panel.restype = I
    lib._ppb_tp_bound = True


_EPI = {None: 0, "bias": 1, "relu": 2, "gelu": 3}


def allgather_linear(a_shard_ptrs: List[int], rows_per_rank: int, K: int, lda: int, w: torch.Tensor, rank: int,
                     bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
                     out: Optional[torch.Tensor] = None, w_mn_major: bool = False) -> torch.Tensor:
    """K12: y[M, N] = act(concat_r(A_r)[M, K] @ w^T + bias) with A_r = the [rows_per_rank, K] bf16 shard at device
    pointer ``a_shard_ptrs[r]`` (own buffer or a peer's symmetric buffer).  w [N, K] (or [K, N] if w_mn_major)."""
    lib = _lib.load(required=True)
    _bind_tp(lib)
    tp = len(a_shard_ptrs)
    M = rows_per_rank * tp
    N = w.shape[1] if w_mn_major else w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=w.device)
    epi = 0 if bias is None else _EPI["bias" if act is None else act]
    arr = (c_void_p * tp)(*a_shard_ptrs)
    rc = lib.ppb_gemm_allgather_bf16(arr, w.data_ptr(), out.data_ptr(), M, N, K, lda, w.stride(0), out.stride(0), tp, rank,
                                     int(w_mn_major), epi, _lib.ptr(bias), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_gemm_allgather_bf16")
    _lib.count_launch()
    return out


def _gelu_bwd_ref(dh: torch.Tensor, hpre: torch.Tensor) -> torch.Tensor:
    a = hpre.float()
    cdf = 0.5 * (1.0 + torch.erf(a * 0.70710678118))
    pdf = 0.3989422804 * torch.exp(-0.5 * a * a)
    return (dh.float() * (cdf + a * pdf)).to(dh.dtype)


class GatherFlags:
    """Panel counters of the fused all-gather GEMM: one uint32 per 128-row panel of the gathered activations, zeroed
    once and then only ever incremented by the kernels; ``next_target`` is what every counter reaches after the next
    call (the counters are never reset, so no memset sits between two launches)."""

    def __init__(self, M: int, device: torch.device):
        self.flags = torch.zeros((M + 127) // 128, dtype=torch.int32, device=device)
        self.total = 0

    def next_target(self, K: int) -> int:
        lib = _lib.load(required=True)
        _bind_tp(lib)
        self.total += int(lib.ppb_gemm_allgather_fused_chunks_per_panel(K))
        if self.total >= 2 ** 31:   # ~10^8 calls away; start over rather than wrap
            torch.cuda.current_stream().synchronize()
            self.flags.zero_()
            self.total = int(lib.ppb_gemm_allgather_fused_chunks_per_panel(K))
        return self.total


def fused_allgather_ok(rows_per_rank: int, K: int) -> bool:
    return rows_per_rank % 128 == 0 and K % 64 == 0


def allgather_linear_fused(a_shard_ptrs: List[int], rows_per_rank: int, K: int, full: torch.Tensor, gflags: GatherFlags,
                           w: torch.Tensor, rank: int, bias: Optional[torch.Tensor] = None, act: Optional[str] = None,
                           out: Optional[torch.Tensor] = None, w_mn_major: bool = False,
                           mask_pos_aux: Optional[torch.Tensor] = None, gelu_aux: Optional[torch.Tensor] = None,
                           alpha: float = 1.0) -> torch.Tensor:
    """K12, fused: y[M, N] = act(concat_r(A_r) @ op(w) + bias) in ONE kernel that also gathers: copier warps stream every
    shard (peers over NVLink, each byte once) into ``full`` [M, K] and release 128-row panels through ``gflags`` while
    the tiles of the own shard are already being multiplied.  ``full`` holds the gathered activations afterwards."""
    lib = _lib.load(required=True)
    _bind_tp(lib)
    tp = len(a_shard_ptrs)
    M = rows_per_rank * tp
    N = w.shape[1] if w_mn_major else w.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=w.device)
    epi, aux = (0 if bias is None else _EPI["bias" if act is None else act]), None
    if mask_pos_aux is not None:      # activation backward in the epilogue: . * alpha * (aux > 0)
        epi, aux = 5, mask_pos_aux
    elif gelu_aux is not None:        # . * gelu'(aux)
        epi, aux = 6, gelu_aux
    arr = (c_void_p * tp)(*a_shard_ptrs)
    rc = lib.ppb_gemm_allgather_fused_bf16(arr, full.data_ptr(), gflags.flags.data_ptr(), gflags.next_target(K), w.data_ptr(),
                                           out.data_ptr(), M, N, K, w.stride(0), out.stride(0), tp, rank, int(w_mn_major),
                                           epi, _lib.ptr(bias), _lib.ptr(aux), 0 if aux is None else aux.stride(0),
                                           float(alpha), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_gemm_allgather_fused_bf16")
    _lib.count_launch()
    return out


def linear_reducescatter(x: torch.Tensor, w: torch.Tensor, dst_shard_ptrs: List[int], ld_dst: int,
                         w_mn_major: bool = False) -> None:
    """K11: for every owner o, dst_o[M/tp, N] (fp32) += rows of (x[M, K] @ w^T) owned by o.  The caller zeroes the
    destination buffers beforehand and synchronises the ranks afterwards."""
    lib = _lib.load(required=True)
    _bind_tp(lib)
    tp = len(dst_shard_ptrs)
    M, K = x.shape
    N = w.shape[1] if w_mn_major else w.shape[0]
    arr = (c_void_p * tp)(*dst_shard_ptrs)
    rc = lib.ppb_gemm_reducescatter_bf16(x.data_ptr(), w.data_ptr(), arr, M, N, K, x.stride(0), w.stride(0), ld_dst, tp,
                                         int(w_mn_major), torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "ppb_gemm_reducescatter_bf16")
    _lib.count_launch()


class SymmetricBuffer:
    """A tensor in symmetric memory + the table of its peer-mapped addresses on every rank of ``group``."""

    def __init__(self, shape, dtype, device, group):
        import torch.distributed._symmetric_memory as symm_mem
        self.tensor = symm_mem.empty(*shape, dtype=dtype, device=device)
        self.tensor.zero_()
        self.hdl = symm_mem.rendezvous(self.tensor, group)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.rank, self.world = self.hdl.rank, self.hdl.world_size

    def barrier(self, channel: int = 0) -> None:
        self.hdl.barrier(channel=channel)


class TensorParallelMLP:
    """Megatron-style tensor + sequence parallel feed-forward block built from the two fused kernels:

        x_shard [T/tp, d]  --(K12 all-gather -> GEMM, bias+act epilogue)-->  h [T, f/tp]
        h                  --(K11 GEMM -> reduce-scatter)----------------->  y_shard [T/tp, d] (fp32 sums)

    ``w1`` is the [f/tp, d] row slice of linear1 and ``w2`` the [d, f/tp] column slice of linear2 owned by this
    rank.  Activations stay sharded on the sequence dimension between the blocks."""

    def __init__(self, w1: torch.Tensor, b1: torch.Tensor, w2: torch.Tensor, T: int, group, act: str = "relu"):
        dev = w1.device
        self.w1, self.b1, self.w2, self.act = w1, b1, w2, act
        self.d = w1.shape[1]
        import torch.distributed as dist
        self.tp = dist.get_world_size(group)
        self.rows = T // self.tp
        self.x_sym = SymmetricBuffer((self.rows, self.d), torch.bfloat16, dev, group)
        self.y_sym = SymmetricBuffer((self.rows, self.d), torch.float32, dev, group)
        self.rank = self.x_sym.rank
        self.fused_gather = fused_allgather_ok(self.rows, self.d)
        self.full = torch.empty((T, self.d), dtype=torch.bfloat16, device=dev)
        self.gflags = GatherFlags(T, dev)

    def forward(self, x_shard: torch.Tensor) -> torch.Tensor:
        self.x_sym.tensor.copy_(x_shard)
        self.y_sym.tensor.zero_()
        self.x_sym.barrier(0)                       # every rank's input shard and zeroed output are in place
        if self.fused_gather:
            h = allgather_linear_fused(self.x_sym.ptrs, self.rows, self.d, self.full, self.gflags, self.w1, self.rank,
                                       self.b1, self.act)
        else:
            h = allgather_linear(self.x_sym.ptrs, self.rows, self.d, self.d, self.w1, self.rank, self.b1, self.act)
        linear_reducescatter(h, self.w2, self.y_sym.ptrs, self.d)
        self.y_sym.barrier(1)                       # all partial sums have landed
        return self.y_sym.tensor


class TPComm:
    """Tensor-parallel communicator of one pipeline stage: the two fused compute+collective ops every sharded
    linear layer needs, with sequence-parallel activations ([T/tp, d] row shards) on both sides.

        ag_linear : [T/tp, d] shard  --all-gather -> GEMM (K12)-->  [T, N_local]      (column-parallel weights)
        linear_rs : [T, K_local]     --GEMM -> reduce-scatter (K11)-->  [T/tp, d] fp32 (row-parallel weights)

    On CUDA the shards live in symmetric memory and the kernels read / reduce-add peer memory over NVLink; buffers
    are double-buffered so ONE device-side barrier per op is enough (all shards written / all partial sums landed):
    passing the barrier of op k+1 proves every rank finished op k, so the buffer of op k may be rewritten at k+2.
    On CPU (gloo) the same interface runs all_gather / all_reduce + matmul, which is what the multi-process CPU tests
    check the sharded layer program against.  Every AG input and every RS output of a transformer layer has d
    columns, so two [T/tp, d] buffers per direction serve the whole stage."""

    def __init__(self, group, rows: int, d: int, device: torch.device):
        import torch.distributed as dist
        self.group = group
        self.tp = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.rows, self.d = int(rows), int(d)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        if self.cuda:
            self.x_sym = [SymmetricBuffer((self.rows, self.d), torch.bfloat16, self.device, group) for _ in range(2)]
            self.y_sym = [SymmetricBuffer((self.rows, self.d), torch.float32, self.device, group) for _ in range(2)]
            self.kx = self.ky = 0
            import os
            self.direct_peer = os.environ.get("PPB_TP_DIRECT_PEER", "0") == "1"
            self.fused_gather = fused_allgather_ok(self.rows, self.d) and os.environ.get("PPB_TP_FUSED_GATHER", "1") != "0"
            self.gflags = GatherFlags(self.rows * self.tp, self.device)
            self.x_sym[0].barrier(0)   # every rank's zeroed reduce-scatter buffers exist before the first push

    # ---- all-gather -> GEMM -------------------------------------------------------------------------------------
    def ag_linear(self, x_shard: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
                  act: Optional[str] = None, w_mn_major: bool = False, need_full: bool = True,
                  mask_pos_aux: Optional[torch.Tensor] = None, gelu_aux: Optional[torch.Tensor] = None, alpha: float = 1.0):
        """(act(gather(x_shard) @ op(w) + bias), gather(x_shard) or None).  w is [N, d] (or [d, N] when w_mn_major).
        ``mask_pos_aux`` / ``gelu_aux`` (+ ``alpha``) apply an activation derivative in the GEMM epilogue instead
        (the input-gradient GEMM of a row-parallel layer): y = alpha * (gather(x) @ op(w)) * act'(aux)."""
        import torch.distributed as dist
        if not self.cuda:
            full = torch.empty((self.rows * self.tp, self.d), dtype=x_shard.dtype)
            dist.all_gather_into_tensor(full, x_shard.contiguous(), group=self.group)
            y = full @ (w if w_mn_major else w.t())
            if bias is not None:
                y = y + bias
            if act == "relu":
                y = torch.relu(y)
            elif act == "gelu":
                y = torch.nn.functional.gelu(y)
            if mask_pos_aux is not None:
                y = torch.where(mask_pos_aux > 0, y * alpha, torch.zeros_like(y))
            elif gelu_aux is not None:
                y = _gelu_bwd_ref(y * alpha, gelu_aux)
            return y, (full if need_full else None)
        buf = self.x_sym[self.kx]
        self.kx ^= 1
        buf.tensor.copy_(x_shard)
        buf.barrier(0)
        if self.direct_peer and not need_full:
            # the GEMM's TMA producer loads remote A tiles straight from the peers' buffers.  Every N tile re-fetches
            # its A tiles and peer memory is not cached in the local L2, so a remote tile crosses NVLink N/256 times
            # (43x at Llama-7B FFN width: measured 5.4 ms vs 1.5 ms of GEMM time per layer) -- only worth it for narrow N
            return allgather_linear(buf.ptrs, self.rows, self.d, self.d, w, self.rank, bias, act, w_mn_major=w_mn_major), None
        full = torch.empty((self.rows * self.tp, self.d), dtype=torch.bfloat16, device=self.device)
        if self.fused_gather:
            # ONE kernel: copier warps bring every shard over NVLink exactly once into `full` (which the weight-gradient
            # GEMM of the backward pass wants anyway) and release panels to the tiles as they land
            out = allgather_linear_fused(buf.ptrs, self.rows, self.d, full, self.gflags, w, self.rank, bias, act,
                                         w_mn_major=w_mn_major, mask_pos_aux=mask_pos_aux, gelu_aux=gelu_aux, alpha=alpha)
            return out, (full if need_full else None)
        # shapes the fused kernel does not take (rows per rank not a multiple of 128, K not of 64): pull the peers'
        # shards next to ours, then run the gather kernel with every shard pointer local
        for r in range(self.tp):
            src = buf.tensor if r == self.rank else buf.hdl.get_buffer(r, (self.rows, self.d), torch.bfloat16)
            full[r * self.rows:(r + 1) * self.rows].copy_(src)
        base = full.data_ptr()
        ptrs = [base + r * self.rows * self.d * 2 for r in range(self.tp)]
        out = allgather_linear(ptrs, self.rows, self.d, self.d, w, self.rank, bias, act, w_mn_major=w_mn_major)
        if mask_pos_aux is not None:      # unusual shapes only (the fused kernel applies these in its epilogue)
            out = torch.where(mask_pos_aux > 0, out * alpha, torch.zeros_like(out))
        elif gelu_aux is not None:
            out = _gelu_bwd_ref(out * alpha, gelu_aux)
        return out, (full if need_full else None)

    # ---- GEMM -> reduce-scatter ---------------------------------------------------------------------------------
    def linear_rs(self, x: torch.Tensor, w: torch.Tensor, w_mn_major: bool = False) -> torch.Tensor:
        """Rows of sum_over_ranks(x_r @ op(w_r)) owned by this rank, fp32 [T/tp, d].  The result lives in a symmetric buffer
        that the NEXT call re-zeroes: copy or convert it right away.  w is [d, K_local] (or [K_local, d] when w_mn_major)."""
        import torch.distributed as dist
        if not self.cuda:
            y = (x @ (w if w_mn_major else w.t())).float()
            dist.all_reduce(y, group=self.group)
            return y[self.rank * self.rows:(self.rank + 1) * self.rows]
        cur, other = self.y_sym[self.ky], self.y_sym[self.ky ^ 1]
        self.ky ^= 1
        other.tensor.zero_()   # the previous call's result (already copied out); peers add to it only after the barrier below
        linear_reducescatter(x, w, cur.ptrs, self.d, w_mn_major=w_mn_major)
        cur.barrier(1)
        return cur.tensor

    def rs_finish(self, y32: torch.Tensor, bias: Optional[torch.Tensor], dtype: torch.dtype) -> torch.Tensor:
        """fp32 reduce-scatter result (+ the replicated bias of a row-parallel layer) -> a fresh tensor in the activation
        dtype, one pass (``ppb_bias_cast_f32_to_bf16``); the symmetric reduce-scatter buffer is free for reuse afterwards."""
        if not (self.cuda and dtype == torch.bfloat16 and y32.shape[1] % 4 == 0):
            return ((y32 + bias.float()) if bias is not None else y32).to(dtype).clone()
        lib = _lib.load(required=True)
        out = torch.empty(y32.shape, dtype=torch.bfloat16, device=y32.device)
        lib.ppb_bias_cast_f32_to_bf16.argtypes = [c_void_p, c_void_p, c_void_p, c_longlong, c_int, c_void_p]
        lib.ppb_bias_cast_f32_to_bf16.restype = c_int
        _lib.check(lib.ppb_bias_cast_f32_to_bf16(y32.data_ptr(), _lib.ptr(bias), out.data_ptr(), y32.shape[0], y32.shape[1],
                                                 torch.cuda.current_stream().cuda_stream), "ppb_bias_cast_f32_to_bf16")
        _lib.count_launch()
        return out

    def all_reduce(self, t: torch.Tensor) -> None:
        import torch.distributed as dist
        dist.all_reduce(t, group=self.group)
