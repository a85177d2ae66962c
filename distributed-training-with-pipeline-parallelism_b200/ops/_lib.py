"""ctypes loader for libppb200.so (C ABI in include/ppb200.h).

The library is built in-tree by ``build.py`` (nvcc, sm_100a) and travels to the GPU box with the repo
snapshot.  On a machine with a GPU a missing/unloadable library is a hard error -- there is no silent
PyTorch fallback for CUDA tensors (CPU tensors use the fp32 references in ``ops.reference``).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_ulonglong, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libppb200.so")

_lib = None
_load_error: str | None = None


class KernelError(RuntimeError):
    pass


def _declare(lib) -> None:
    P, I, L, F, U = c_void_p, c_int, c_longlong, c_float, c_ulonglong
    lib.ppb_last_error.restype = c_char_p
    lib.ppb_last_error.argtypes = []
    lib.ppb_version.restype = I
    sigs = {
        "ppb_set_dropout_offset_ptr": [P],
        "ppb_gemm_bf16": [P, P, P, I, I, I, L, L, L, I, I, I, I, F, P, P, L, U, F, I, I, P],
        "ppb_add_dropout_norm_fwd": [P, P, P, P, P, P, P, P, I, I, F, F, U, I, P],
        "ppb_norm_bwd": [P, P, P, P, P, P, P, P, P, P, P, P, L, I, I, F, U, I, P],
        "ppb_xent_fwd_bwd": [P, L, P, P, I, I, F, I, P],
        "ppb_reduce_sum_f32": [P, L, P, F, I, P],
        "ppb_adamw_step": [P, P, P, P, P, L, F, F, F, F, F, I, F, P, I, P],
        "ppb_sumsq_f32": [P, L, P, P],
        "ppb_clip_coef": [P, F, F, P, P],
        "ppb_embedding_fwd": [P, P, P, I, I, I, P],
        "ppb_embedding_bwd": [P, P, P, P, P, I, I, I, P],
        "ppb_colsum_accum": [P, L, P, I, I, P],
        "ppb_add_n_bf16": [P, P, P, P, L, P],
        "ppb_gelu_fwd": [P, P, L, P],
        "ppb_add_dropout_fwd": [P, P, P, L, F, U, P],
        "ppb_sum_dropout_bwd": [P, P, P, P, P, L, F, U, P],
        "ppb_cast_f32_to_bf16": [P, P, L, P],
        "ppb_swiglu_fwd": [P, L, P, L, L, I, P],
        "ppb_swiglu_bwd": [P, L, P, L, P, L, L, I, P],
        "ppb_rope_inplace": [P, L, I, L, I, I, I, P, I, P],
        "ppb_attn_fwd": [P, P, P, P, P, I, I, I, I, I, L, L, L, L, I, I, I, F, F, U, P],
        "ppb_attn_bwd": [P, P, P, P, P, P, P, P, P, P, P, I, I, I, I, I, L, L, L, L, I, I, I, F, F, U, P],
        "ppb_attn_bwd_scratch_bytes": [I, I, I, I],
        "ppb_gemm_allreduce_bf16": None,  # declared lazily in ops.collective (optional symbols)
    }
    for name, args in sigs.items():
        if args is None or not hasattr(lib, name):
            continue
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = I
    if hasattr(lib, "ppb_norm_bwd_scratch_bytes"):
        lib.ppb_norm_bwd_scratch_bytes.argtypes = [I, I]
        lib.ppb_norm_bwd_scratch_bytes.restype = L
    if hasattr(lib, "ppb_attn_bwd_scratch_bytes"):
        lib.ppb_attn_bwd_scratch_bytes.restype = L


def load(required: bool = False):
    """Return the ctypes handle, or None when the library is not built (only tolerated without a GPU)."""
    global _lib, _load_error
    if _lib is not None:
        return _lib
    if _load_error is None:
        try:
            lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
            _declare(lib)
            _lib = lib
            return _lib
        except OSError as e:  # not built, or libcudart missing
            _load_error = f"{LIB_PATH}: {e}"
    if required:
        raise KernelError(
            f"libppb200.so could not be loaded ({_load_error}); run `python build.py`. "
            "CUDA tensors have no PyTorch fallback in this framework.")
    return None


def available() -> bool:
    return load() is not None


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = _lib.ppb_last_error().decode("utf-8", "replace") if _lib is not None else "?"
        raise KernelError(f"{what} failed (rc={rc}): {msg}")


def ptr(t) -> int:
    """Device pointer of a tensor (0 for None)."""
    return 0 if t is None else t.data_ptr()


# ------------------------------------------------------------------------------------------------------------
# launch counter: bench.py reports how many of OUR kernels were launched inside the timed region
# ------------------------------------------------------------------------------------------------------------
_launches = 0


def count_launch(n: int = 1) -> None:
    global _launches
    _launches += n


def launches() -> int:
    return _launches


def reset_launches() -> None:
    global _launches
    _launches = 0
