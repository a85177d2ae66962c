"""Build check on the machine code itself (no GPU needed): the in-tree library must contain the Blackwell instructions the
design claims -- tcgen05.mma (UTCHMMA; `.2CTA` in the CTA-pair kernels), TMA loads / stores / reduces (UTMALDG / UTMASTG /
UTMAREDG), TMEM loads (LDTM) and, for the TMEM-resident attention forward, TMEM stores (STTM) -- and no legacy
mma.sync (HMMA) in the hot kernels.  Mnemonics: /opt/skills/guides/B200_PROFILING.md."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "distributed-training-with-pipeline-parallelism_b200", "ops", "libppb200.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"

pytestmark = pytest.mark.skipif(not (os.path.exists(LIB) and os.path.exists(CUOBJDUMP)),
                                reason="needs the built library (python build.py) and cuobjdump")


def _functions():
    out = subprocess.run([CUOBJDUMP, "-elf", LIB], capture_output=True, text=True, timeout=300).stdout
    return sorted(set(re.findall(r"\.text\.(_ZN3ppb[0-9A-Za-z_]+)", out)))


def _ops(fn: str):
    out = subprocess.run([CUOBJDUMP, "-sass", "-fun", fn, LIB], capture_output=True, text=True, timeout=300).stdout
    ops = []
    for line in out.splitlines():
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m:
            ops.append(m.group(1))
    return ops


def _pick(funcs, *needles):
    hits = [f for f in funcs if all(n in f for n in needles)]
    assert hits, f"no kernel matching {needles} in the library"
    return hits[0]


def test_library_is_sm100a_tcgen05_code():
    funcs = _functions()
    count = lambda ops, p: sum(1 for o in ops if o.startswith(p))   # noqa: E731
    gemm = _ops(_pick(funcs, "gemm_tcgen05_kernelILi256ELb0ELb0ELi0EE"))
    assert count(gemm, "UTCHMMA") >= 1 and count(gemm, "UTMALDG") >= 2 and count(gemm, "UTMASTG") >= 1 and count(gemm, "LDTM") >= 1
    assert count(gemm, "HMMA") == 0
    pair = _ops(_pick(funcs, "gemm_tcgen05_pair_kernelILi256ELb0ELb0ELi0EE"))
    assert any(o.startswith("UTCHMMA") and "2CTA" in o for o in pair), "CTA-pair GEMM without cta_group::2 MMAs"
    assert any(o.startswith("UTMALDG") and "2CTA" in o for o in pair)
    wgrad = _ops(_pick(funcs, "gemm_tcgen05_kernelILi256ELb1ELb1ELi2EE"))
    assert count(wgrad, "UTMAREDG") >= 1                       # fp32 accumulation through TMA reduce-add
    fused = _ops(_pick(funcs, "gemm_tcgen05_agfused_kernelILi256ELb0EE"))
    assert count(fused, "UBLKCP") >= 1 and count(fused, "UTCHMMA") >= 1   # bulk copies from the peers + MMAs in ONE kernel
    fwd = _ops(_pick(funcs, "attn_fwd_kernelILi64ELi64ELi2ELi1EE"))
    assert count(fwd, "UTCHMMA") >= 16 and count(fwd, "LDTM") >= 1 and count(fwd, "STTM") >= 1 and count(fwd, "HMMA") == 0
    for part in ("attn_bwd_kernelILi64ELi64ELi0EE", "attn_bwd_kernelILi192ELi192ELi1EE", "attn_bwd_kernelILi192ELi192ELi2EE"):
        bwd = _ops(_pick(funcs, part))
        assert count(bwd, "UTCHMMA") >= 16 and count(bwd, "HMMA") == 0, part
    assert count(_ops(_pick(funcs, "attn_bwd_kernelILi192ELi192ELi2EE")), "UTMAREDG") >= 1   # dQ leaves through TMA reduce-add
    wide = _ops(_pick(funcs, "attn_fwd_kernelILi192ELi192ELi1ELi2EE"))
    assert count(wide, "UTCHMMA") >= 16 and count(wide, "UTMALDG") >= 9    # three 64-column chunks for Q, K and V
