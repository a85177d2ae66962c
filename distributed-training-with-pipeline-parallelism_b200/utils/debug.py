"""Debug tooling for the kernel library: the in-repo stand-in for a memory checker.

``compute-sanitizer`` is not always available where the kernels run, and the kernels signal across CTAs / GPUs, so
the framework carries its own checks:

* ``guarded_allocations()``: every tensor the stage program allocates (torch.empty / empty_like / zeros) gets guard
  bands of a canary byte on both sides and a NaN-pattern (0xFF) fill, so an out-of-bounds write shows up in
  ``check()`` with the allocation site, and a read of memory nobody wrote shows up as NaN in the results;
* ``determinism_check()``: run a closure repeatedly while perturbing the caching allocator (so buffers land at
  different addresses) and compare the outputs bit for bit -- separates "order of fp32 atomics" from "stale or
  clobbered operand" when two runs of the same program disagree;
* ``poison_``: fill a tensor with the NaN pattern (cross-GPU flag protocols: poison the receive buffers between epochs).

These found the cross-slot scratch aliasing of the first CUDA-graph pools.
"""
from __future__ import annotations

import contextlib
import os
import random
import traceback
from typing import Callable, List, Optional, Sequence

import torch

CANARY = 0xA5
PAD = 4096


class GuardReport:
    def __init__(self):
        self.entries: List[tuple] = []     # (raw uint8 tensor, payload bytes, allocation site)
        self.violations: List[str] = []
        self.checked = 0

    def check(self, tag: str = "") -> int:
        """Verify the guard bands of every live guarded allocation; returns the number of violated buffers."""
        if self.entries and self.entries[0][0].is_cuda:
            torch.cuda.synchronize()
        bad = 0
        for raw, nbytes, site in self.entries:
            lo = (raw[:PAD] != CANARY).nonzero()
            hi = (raw[PAD + nbytes:] != CANARY).nonzero()
            if lo.numel() or hi.numel():
                bad += 1
                self.violations.append(
                    f"{tag}: {nbytes}-byte buffer allocated at {site}: {lo.numel()} bytes clobbered before "
                    f"(nearest at -{PAD - int(lo.max()) if lo.numel() else 0}), {hi.numel()} after "
                    f"(nearest at +{int(hi.min()) if hi.numel() else 0})")
        self.checked += len(self.entries)
        self.entries.clear()
        return bad


def _site() -> str:
    frames = traceback.extract_stack(limit=8)[:-3]
    return " <- ".join(f"{os.path.basename(f.filename)}:{f.lineno}" for f in reversed(frames[-4:]))


@contextlib.contextmanager
def guarded_allocations(device_type: str = "cuda", poison: bool = True):
    """Context manager: patch torch.empty / torch.empty_like / torch.zeros for ``device_type`` tensors."""
    rep = GuardReport()
    o_empty, o_empty_like, o_zeros = torch.empty, torch.empty_like, torch.zeros

    def guard(shape, dtype, device, zero):
        if isinstance(shape, int):
            shape = (shape,)
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * dtype.itemsize
        nb = (nbytes + 255) // 256 * 256
        raw = o_empty(nb + 2 * PAD, dtype=torch.uint8, device=device)
        raw.fill_(CANARY)
        body = raw[PAD:PAD + nbytes]
        if zero:
            body.zero_()
        elif poison:
            body.fill_(0xFF)
        rep.entries.append((raw, nbytes, _site()))
        return body.view(dtype).view(tuple(int(s) for s in shape))

    def capturing() -> bool:
        return device_type == "cuda" and torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()

    def norm_size(size):
        if len(size) == 1 and isinstance(size[0], (tuple, list, torch.Size)):
            return tuple(size[0])
        return size

    def g_empty(*size, dtype=None, device=None, **kw):
        dev = torch.device(device) if device is not None else torch.device("cpu")
        if dev.type != device_type or kw or capturing():
            return o_empty(*size, dtype=dtype, device=device, **kw)
        return guard(norm_size(size), dtype or torch.get_default_dtype(), dev, False)

    def g_zeros(*size, dtype=None, device=None, **kw):
        dev = torch.device(device) if device is not None else torch.device("cpu")
        if dev.type != device_type or kw or capturing():
            return o_zeros(*size, dtype=dtype, device=device, **kw)
        return guard(norm_size(size), dtype or torch.get_default_dtype(), dev, True)

    def g_empty_like(t, **kw):
        if t.device.type != device_type or kw or capturing():
            return o_empty_like(t, **kw)
        return guard(tuple(t.shape), t.dtype, t.device, False)

    torch.empty, torch.zeros, torch.empty_like = g_empty, g_zeros, g_empty_like
    try:
        yield rep
    finally:
        torch.empty, torch.zeros, torch.empty_like = o_empty, o_zeros, o_empty_like


def poison_(t: torch.Tensor) -> torch.Tensor:
    """Fill with the all-ones byte pattern (NaN for every floating type, -1 for integers)."""
    t.view(torch.uint8).fill_(0xFF) if t.is_contiguous() else t.fill_(float("nan"))
    return t


def determinism_check(fn: Callable[[], Sequence[Optional[torch.Tensor]]], reps: int = 20, device: str = "cuda",
                      seed: int = 0) -> dict:
    """Call ``fn`` ``reps`` times with the allocator perturbed in between; compare every output with the first call.
    Returns {"mismatching": runs that differ, "worst_abs": largest absolute difference, "nan": any NaN seen}."""
    rng = random.Random(seed)
    junk: List[torch.Tensor] = []
    ref = None
    worst, nbad, nan = 0.0, 0, False
    for _ in range(reps):
        junk.append(torch.empty(rng.randint(1, 64) * 512 * rng.choice([1, 7, 33, 257]), dtype=torch.uint8, device=device))
        if len(junk) > 12:
            del junk[rng.randrange(len(junk))]
        outs = [o.detach().clone() for o in fn() if o is not None]
        if device == "cuda":
            torch.cuda.synchronize()
        if ref is None:
            ref = outs
            continue
        bad = False
        for a, b in zip(ref, outs):
            d = (a.double() - b.double()).abs()
            if torch.isnan(d).any():
                nan = bad = True
                continue
            m = d.max().item() if d.numel() else 0.0
            if m != 0.0:
                bad = True
                worst = max(worst, m)
        nbad += bad
    return {"reps": reps, "mismatching": nbad, "worst_abs": worst, "nan": nan}
