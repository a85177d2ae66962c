"""The reference arm's data-parallel composition (`bench.py --impl reference --dp N`): the unmodified helper's pruned stage
module under DistributedDataParallel inside torch's own pipeline runtime.  Run here on CPU / gloo (2 stages x 2
replicas): the helper's split must work on a pipeline group smaller than the world, the replicas must stay bit-identical
(the gradient all-reduce really happens under the 1F1B schedule) and the loss must go down."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "LLMsDistributedTrainingHelper.py")


@pytest.mark.skipif(not os.path.exists(REF), reason="baseline/_ref not installed (DESIGN.md: reference arm)")
def test_reference_arm_with_data_parallel_replicas():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=4", "--master-addr", "127.0.0.1",
           "--master-port", str(29900 + os.getpid() % 90), os.path.join(ROOT, "tests", "_ref_dp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    res = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert res["replicas_equal"] is True
    assert len(res["losses"]) == 2                     # the last stage of each replica
    for h in res["losses"]:
        assert len(h) == 3 and h[-1] < h[0], res


@pytest.mark.skipif(not os.path.exists(REF), reason="baseline/_ref not installed (DESIGN.md: reference arm)")
@pytest.mark.parametrize("sched", ["GPipe", "1F1B", "Interleaved1F1B"])
def test_pipeline_losses_match_the_unmodified_helper(sched):
    """Oracle vs. reference at the PIPELINE level: same weights, same batch, two ranks."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(29800 + (os.getpid() + len(sched)) % 90), os.path.join(ROOT, "tests", "_ref_oracle_worker.py"), sched]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    ranks = json.loads([l for l in r.stdout.splitlines() if l.startswith("[")][-1])
    last = ranks[-1]
    assert len(last["ref_losses"]) == len(last["our_losses"]) == 4
    for a, b in zip(last["our_losses"], last["ref_losses"]):
        assert abs(a - b) < 1e-4 * abs(b), last
    for rk in ranks:
        assert rk["n_grads"] > 0 and rk["grad_rel_err"] < 1e-3, rk
