#!/usr/bin/env python
"""Build libppb200.so (every sm_100a kernel of the framework) in-tree with nvcc.

    python build.py            # incremental: only re-compiles .cu files newer than their .o
    python build.py --force    # rebuild everything
    python build.py --sass     # also dump cuobjdump -sass summaries into docs/sass/

The library is plain C ABI (include/ppb200.h), has no dependency on libtorch, and is loaded by
`ppb200.ops._lib` through ctypes.  nvcc cross-compiles for sm_100a without a GPU present.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "distributed-training-with-pipeline-parallelism_b200", "ops")
BUILD = os.path.join(ROOT, "build")
LIB = os.path.join(OUT_DIR, "libppb200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-I", os.path.join(ROOT, "include"),
    "-I", CSRC,
]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found")
    return p


def sources() -> list[str]:
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _fingerprint(src: str) -> str:
    """Content hash of everything an object file depends on: its source, every header, the flags and the compiler.
    Staleness is decided by CONTENT, not by mtimes, so a stale .o/.so pair (fresh checkout, copied tree, clock skew)
    can never pass as up to date."""
    import hashlib
    h = hashlib.sha256()
    files = [os.path.join(CSRC, src)] + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    files.append(os.path.join(ROOT, "include", "ppb200.h"))
    for f in files:
        h.update(f.encode())
        h.update(open(f, "rb").read())
    h.update(" ".join(NVCC_FLAGS).encode())
    h.update(nvcc_path().encode())
    return h.hexdigest()


def compile_one(src: str, force: bool, verbose: bool) -> tuple[str, str]:
    obj = os.path.join(BUILD, src.replace(".cu", ".o"))
    stamp = obj + ".sha256"
    spath = os.path.join(CSRC, src)
    fp = _fingerprint(src)
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read().strip() == fp:
        return obj, ""
    cmd = [nvcc_path(), *NVCC_FLAGS, "-Xptxas", "-v" if verbose else "-warn-spills", "-c", spath, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    with open(stamp, "w") as f:
        f.write(fp)
    return obj, r.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(OUT_DIR, exist_ok=True)
    srcs = sources()
    logs = []
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: compile_one(s, force, verbose), srcs))
    objs = [o for o, _ in results]
    logs = [l for _, l in results if l]
    # the library is stamped with the fingerprints of the objects it was linked from
    lib_fp = "\n".join(open(o + ".sha256").read().strip() for o in objs)
    lib_stamp = os.path.join(BUILD, "libppb200.so.sha256")
    if (force or not os.path.exists(LIB) or not os.path.exists(lib_stamp) or open(lib_stamp).read() != lib_fp):
        cmd = [nvcc_path(), "-shared", "-o", LIB, *objs, "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        with open(lib_stamp, "w") as f:
            f.write(lib_fp)
    if verbose:
        print("\n".join(logs))
    return LIB


def dump_sass() -> None:
    """Write per-kernel SASS opcode histograms + the tcgen05/TMA evidence lines to docs/sass/."""
    out_dir = os.path.join(ROOT, "docs", "sass")
    os.makedirs(out_dir, exist_ok=True)
    r = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(r.stderr)
    kernels: dict[str, dict[str, int]] = {}
    cur = None
    for line in r.stdout.splitlines():
        line = line.strip()
        if line.startswith("Function :"):
            cur = line.split(":", 1)[1].strip()
            kernels[cur] = {}
        elif cur and line.startswith("/*") and "*/" in line:
            body = line.split("*/", 1)[1].strip()
            if not body or body.startswith("/*"):
                continue
            op = body.split()[0].rstrip(";")
            if op.startswith("@"):
                parts = body.split()
                op = parts[1].rstrip(";") if len(parts) > 1 else op
            base = op.split(".")[0]
            kernels[cur][base] = kernels[cur].get(base, 0) + 1
    keys = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "SYNCS", "HMMA", "REDG", "ATOMG"]
    with open(os.path.join(out_dir, "summary.md"), "w") as f:
        f.write("# SASS evidence per kernel (cuobjdump -sass libppb200.so)\n\n")
        f.write("Counts of the Blackwell-specific opcodes per kernel. `UTCHMMA` = tcgen05.mma, `LDTM`/`STTM` = "
                "tcgen05.ld/st, `UTMALDG`/`UTMASTG`/`UTMAREDG` = TMA load/store/reduce, `SYNCS` = mbarrier.\n\n")
        f.write("| kernel | instrs | " + " | ".join(keys) + " |\n|---|---|" + "---|" * len(keys) + "\n")
        for k, hist in sorted(kernels.items()):
            tot = sum(hist.values())
            f.write(f"| `{k[:110]}` | {tot} | " + " | ".join(str(hist.get(x, 0)) for x in keys) + " |\n")
    print(f"wrote {out_dir}/summary.md ({len(kernels)} kernels)")
    dump_listings(out_dir)


# one representative instantiation per kernel family: full `cuobjdump -sass` listing, committed under docs/sass/
LISTINGS = {
    "gemm_fwd_bn256": "gemm_tcgen05_kernelILi256ELb0ELb0ELi0EE",
    "gemm_wgrad_bn256_accum": "gemm_tcgen05_kernelILi256ELb1ELb1ELi2EE",
    "gemm_pair_fwd_bn256": "gemm_tcgen05_pair_kernelILi256ELb0ELb0ELi0EE",
    "gemm_pair_dgrad_bn256": "gemm_tcgen05_pair_kernelILi256ELb0ELb1ELi0EE",
    "k10_gemm_allreduce_bn256": "gemm_tcgen05_allreduce_kernelILi256EE",
    "k11_gemm_reducescatter_bn256": "gemm_tcgen05_rs_kernelILi256ELb0EE",
    "k11_gemm_reducescatter_pair_bn256": "gemm_tcgen05_rs_pair_kernelILi256ELb0EE",
    "k12_allgather_fused_bn256": "gemm_tcgen05_agfused_kernelILi256ELb0EE",
    "k12_allgather_fused_pair_bn256": "gemm_tcgen05_agfused_pair_kernelILi256ELb0EE",
    "attn_fwd_dh64": "attn_fwd_kernelILi64ELi64ELi2ELi1EE",
    "attn_bwd_dh64": "attn_bwd_kernelILi64ELi64ELi0EE",
    "attn_fwd_dh192": "attn_fwd_kernelILi192ELi192ELi1ELi2EE",
    "attn_bwd_dh192_dv": "attn_bwd_kernelILi192ELi192ELi1EE",
    "attn_bwd_dh192_dk_dq": "attn_bwd_kernelILi192ELi192ELi2EE",
    "ln_bwd_d1024": "ln_bwd_kernelILi128EE",
    "ln_fwd_d1024": "ln_fwd_kernelILi4EE",
    "xent": "xent_fwd_bwd_kernelILb1EE",
    "adamw": "adamw_kernel",
}


def dump_listings(out_dir: str) -> None:
    names = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs = [l.split(":", 1)[1].strip() for l in names.splitlines() if l.strip().startswith("Function :")]
    index = ["# Full SASS listings (cuobjdump -sass), one representative instantiation per kernel family\n",
             "Regenerate with `python build.py --sass`.  `UTCHMMA` = tcgen05.mma (`.2CTA` = cta_group::2), `UTMALDG` / `UTMASTG` /",
             "`UTMAREDG` = TMA load / store / reduce, `UBLKCP` = 1-D bulk copy (the all-gather copier), `LDTM` / `STTM` = tcgen05.ld / st,",
             "`UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier.\n", "| listing | kernel | instructions | UTCHMMA | UTMALDG | UTMASTG | UTMAREDG | UBLKCP | LDTM | HMMA |", "|---|---|---|---|---|---|---|---|---|---|"]
    for short, pat in LISTINGS.items():
        match = [f for f in funcs if pat in f]
        if not match:
            continue
        r = subprocess.run(["cuobjdump", "-sass", "-fun", match[0], LIB], capture_output=True, text=True)
        # keep the function's instruction stream only: address + instruction, without the encoding words and fatbin headers
        import re
        lines, on = [], False
        for l in r.stdout.splitlines():
            if "Function :" in l:
                on = True
                lines.append(l.strip())
                continue
            if not on:
                continue
            mm = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\*", l)
            if mm:
                lines.append(f"/*{mm.group(1)}*/  {mm.group(2)} ;")
            elif l.strip().startswith(".") and ":" in l:
                lines.append(l.strip())
        with open(os.path.join(out_dir, short + ".sass"), "w") as f:
            f.write("\n".join(lines) + "\n")
        ops = [re.sub(r"^@!?U?P\d+\s+", "", l.split("*/", 1)[1].strip()).split()[0] for l in lines[1:] if l.startswith("/*")]
        cnt = lambda op: sum(1 for o in ops if o.startswith(op))   # noqa: E731
        n_instr = len(ops)
        index.append(f"| [{short}.sass]({short}.sass) | `{match[0][:70]}` | {n_instr} | {cnt('UTCHMMA')} | {cnt('UTMALDG')} | "
                     f"{cnt('UTMASTG')} | {cnt('UTMAREDG')} | {cnt('UBLKCP')} | {cnt('LDTM')} | {cnt('HMMA')} |")
    with open(os.path.join(out_dir, "README.md"), "w") as f:
        f.write("\n".join(index) + "\n")
    print(f"wrote {len(LISTINGS)} listings to {out_dir}")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", "-v", action="store_true")
    ap.add_argument("--sass", action="store_true")
    a = ap.parse_args()
    lib = build(force=a.force, verbose=a.verbose)
    print(lib)
    if a.sass:
        dump_sass()
