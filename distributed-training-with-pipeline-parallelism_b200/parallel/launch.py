"""Process launcher: N ranks on one host via multiprocessing 'spawn' (nb:311-325), with what the reference lacks
(SURVEY.md section 5.3): an ephemeral rendezvous port, a timeout, kill-all on the first failure, and the same
error-dict contract (``{'error': ...}``, ``'No results returned'``)."""
from __future__ import annotations

import multiprocessing as mp
import queue as _queue
import socket
import time
from typing import Callable, Optional, Sequence


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch_workers(target: Callable, world_size: int, args: Sequence, kwargs: Optional[dict] = None,
                   timeout: float = 1800.0, poll: float = 0.2) -> dict:
    """Run ``target(rank, world_size, *args, results_queue, **kwargs)`` on ``world_size`` spawned processes and
    return the first result dict posted to the queue."""
    kwargs = dict(kwargs or {})
    kwargs.setdefault("master_port", free_port())
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = []
    for rank in range(world_size):
        p = ctx.Process(target=target, args=(rank, world_size, *args, q), kwargs=kwargs)
        p.start()
        procs.append(p)
    result: Optional[dict] = None
    deadline = time.time() + timeout
    try:
        while time.time() < deadline:
            try:
                item = q.get(timeout=poll)
                if result is None or "error" in item:
                    result = item
                if "error" in item:
                    break
            except _queue.Empty:
                pass
            alive = [p for p in procs if p.is_alive()]
            if not alive:
                break
            dead_bad = [p for p in procs if not p.is_alive() and p.exitcode not in (0, None)]
            if dead_bad and result is None:
                result = {"error": f"rank process exited with code {dead_bad[0].exitcode}"}
                break
        else:
            result = {"error": f"timeout after {timeout:.0f} s"}
        # drain anything posted while we were checking liveness
        try:
            while True:
                item = q.get_nowait()
                if result is None or ("error" in item and "error" not in result):
                    result = item
        except _queue.Empty:
            pass
    finally:
        grace = time.time() + (5.0 if result is not None and "error" not in result else 1.0)
        for p in procs:
            p.join(timeout=max(0.0, grace - time.time()))
        for p in procs:
            if p.is_alive():
                p.terminate()
        for p in procs:
            p.join(timeout=5.0)
            if p.is_alive():
                p.kill()
    return result if result is not None else {"error": "No results returned"}
