"""The allocation-free LayerNorm-backward entry point, called straight through the C ABI.

``ppb_norm_bwd_noalloc`` exists for hosts that are not allowed to create device memory (a runtime that owns the
whole HBM arena, a plug-in hook inside another framework, a thread that must not hit cudaMalloc's implicit
synchronisation): its scratch is a fixed array inside the library instead of a caller-supplied buffer.  The test
feeds the same inputs to it and to the regular ``ppb_norm_bwd`` and requires the same outputs (within the
tolerance the other kernel tests use), twice in a row -- the second call runs on an array that still holds the
first call's partial sums -- and checks that an over-sized problem is refused instead of overrunning the array.

This file sorts last on purpose: with ``pytest -x`` every other test has already run if this one fails.
"""
import ctypes
from ctypes import c_float, c_int, c_longlong, c_ulonglong, c_void_p

import pytest
import torch

from ppb200.ops import _lib

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _bind():
    lib = _lib.load(required=True)
    P, I, L, F, U = c_void_p, c_int, c_longlong, c_float, c_ulonglong
    lib.ppb_norm_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, P, P, P, P, I, I, F, U, I, P]
    lib.ppb_norm_bwd_noalloc.restype = I
    lib.ppb_norm_bwd_noalloc_max_d.argtypes = []
    lib.ppb_norm_bwd_noalloc_max_d.restype = L
    return lib


def _p(t):
    return 0 if t is None else t.data_ptr()


@pytest.mark.parametrize("T,d,p,rms", [(4096, 1024, 0.1, False), (333, 768, 0.0, False), (1024, 256, 0.1, True)])
def test_noalloc_matches_regular_entry_point(T, d, p, rms):
    lib = _bind()
    dev = "cuda"
    torch.manual_seed(11)
    dys = [torch.randn(T, d, device=dev, dtype=torch.bfloat16) for _ in range(3)]
    z = torch.randn(T, d, device=dev, dtype=torch.bfloat16)
    mean = z.float().mean(-1).contiguous()
    rstd = torch.rsqrt(z.float().var(-1, unbiased=False) + 1e-5).contiguous()
    gamma = (1 + 0.1 * torch.randn(d, device=dev)).to(torch.bfloat16)
    stream = torch.cuda.current_stream().cuda_stream
    seed = 2024

    # regular entry point (caller-supplied scratch)
    need = int(lib.ppb_norm_bwd_scratch_bytes(T, d))
    scratch = torch.zeros(need, dtype=torch.uint8, device=dev)
    dz_ref, da_ref = torch.empty_like(z), torch.empty_like(z)
    dg_ref = torch.zeros(d, device=dev)
    db_ref = None if rms else torch.zeros(d, device=dev)
    for _ in range(2):
        rc = lib.ppb_norm_bwd(_p(dys[0]), _p(dys[1]), _p(dys[2]), _p(z), _p(None if rms else mean), _p(rstd), _p(gamma),
                              _p(dz_ref), _p(da_ref), _p(dg_ref), _p(db_ref), _p(scratch), need, T, d, p, seed, int(rms),
                              stream)
        _lib.check(rc, "ppb_norm_bwd")

    # allocation-free entry point, two calls in a row on the shared built-in array
    dz, da = torch.empty_like(z), torch.empty_like(z)
    dg = torch.zeros(d, device=dev)
    db = None if rms else torch.zeros(d, device=dev)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for call in range(2):
        rc = lib.ppb_norm_bwd_noalloc(_p(dys[0]), _p(dys[1]), _p(dys[2]), _p(z), _p(None if rms else mean), _p(rstd),
                                      _p(gamma), _p(dz), _p(da), _p(dg), _p(db), T, d, p, seed, int(rms), stream)
        _lib.check(rc, "ppb_norm_bwd_noalloc")
        torch.cuda.synchronize()
        assert rel_err(dz, dz_ref) < 1e-2, call
        assert rel_err(da, da_ref) < 1e-2, call
    assert torch.cuda.memory_allocated() == before  # nothing was allocated through torch's allocator either
    # both paths accumulated twice into their fp32 gradient accumulators
    assert rel_err(dg, dg_ref) < 1e-2
    if not rms:
        assert rel_err(db, db_ref) < 1e-2


def test_noalloc_refuses_oversized_problem():
    lib = _bind()
    max_d = int(lib.ppb_norm_bwd_noalloc_max_d())
    d = max_d + 8
    T = 16
    t = torch.zeros(T, d, device="cuda", dtype=torch.bfloat16)
    f = torch.zeros(max(T, d), device="cuda")
    rc = lib.ppb_norm_bwd_noalloc(_p(t), 0, 0, _p(t), _p(f), _p(f), _p(t), _p(t), 0, _p(f), _p(f), T, d, 0.0, 0, 0,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.ppb_last_error().decode()
    assert "exceeds" in msg and str(max_d) in msg
