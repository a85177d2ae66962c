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
                    push_add(self.grad[lo:hi], [p + 4 * lo for p in self.base_ptrs], 1.0)
            pos = max(pos, off + n)
        self.hdl.barrier(channel=0)

    def after_optimizer(self) -> None:
        self.hdl.barrier(channel=1)
