"""The stock-torch tensor-parallel baseline behind `bench.py --impl reference --tp N` (benchmarks/tp_baseline.py) must be a
CORRECT model before its speed means anything: sharded over 2 tensor-parallel ranks, and over 2 pipeline stages x 2
tensor-parallel ranks, it has to follow the same loss trajectory (forward, backward, optimizer) as on one rank."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(world: int, tp: int, block: str, sched: str = "1F1B"):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(29700 + (os.getpid() + 7 * world + tp) % 200), os.path.join(ROOT, "tests", "_tp_baseline_worker.py"),
           str(tp), block, sched]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])["losses"]


@pytest.mark.parametrize("block", ["gpt", "llama"])
def test_tp_baseline_matches_its_single_rank_model(block):
    ref = _run(1, 1, block)
    assert ref[-1] < ref[0]                                  # it trains
    tp2 = _run(2, 2, block)
    pp2tp2 = _run(4, 2, block, "GPipe" if block == "gpt" else "1F1B")
    for got in (tp2, pp2tp2):
        assert len(got) == len(ref)
        for a, b in zip(got, ref):
            assert abs(a - b) < 2e-4 * abs(b), (got, ref)
