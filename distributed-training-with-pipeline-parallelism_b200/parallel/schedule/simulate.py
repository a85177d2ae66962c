"""Discrete-event simulator for lowered schedules: deadlock check + predicted timeline / bubble.

Each rank executes its action list in order on one compute engine.  Communication is asynchronous:
a SEND becomes available when the compute that produced it ends and arrives ``comm_latency`` later; a RECV is a
no-cost placeholder; the compute that consumes a received tensor waits for the arrival.  This mirrors what the
CUDA runner does with events between the compute stream and the comm streams, so the simulated bubble is the
target the measured bubble is compared against.  (torch ships a toy simulator that raises on the UNSHARD
actions its own lowering inserts, schedules.py:3255/:3338 -- SURVEY.md section 3.5; this one is independent.)
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from .ir import Action, ActionType as T


@dataclass
class SimResult:
    makespan: float
    busy: Dict[int, float]                      # per rank: sum of compute durations
    bubble: Dict[int, float]                    # per rank: 1 - busy / makespan
    timeline: Dict[int, List[Tuple[Action, float, float]]] = field(default_factory=dict)

    @property
    def bubble_max(self) -> float:
        return max(self.bubble.values())

    @property
    def bubble_mean(self) -> float:
        return sum(self.bubble.values()) / len(self.bubble)


class DeadlockError(RuntimeError):
    pass


def simulate(order: Dict[int, Sequence[Action]], num_stages: int, cost: Optional[Callable[[Action], float]] = None,
             comm_latency: float = 0.0, f_cost: float = 1.0, b_cost: float = 2.0) -> SimResult:
    """Simulate ``order``.  ``cost(action)`` overrides the default F = f_cost, B = b_cost, I = W = b_cost / 2,
    everything else 0.  Raises DeadlockError when no rank can progress."""

    def dur(a: Action) -> float:
        if cost is not None:
            return cost(a)
        if a.type == T.F:
            return f_cost
        if a.type == T.B:
            return b_cost
        if a.type in (T.I, T.W):
            return b_cost / 2
        return 0.0

    ranks = sorted(order)
    idx = {r: 0 for r in ranks}
    clock = {r: 0.0 for r in ranks}
    busy = {r: 0.0 for r in ranks}
    end_of: Dict[Action, float] = {}      # completion time of compute actions (global: stage ids are unique)
    arrival: Dict[Action, float] = {}     # RECV action -> time its payload is available
    timeline: Dict[int, List[Tuple[Action, float, float]]] = {r: [] for r in ranks}
    local = {r: {a for a in order[r]} for r in ranks}

    def input_ready(r: int, a: Action) -> Optional[float]:
        """Time at which the inputs of compute action ``a`` are available, None if not yet known."""
        if a.type == T.F:
            if a.stage == 0:
                return 0.0
            rv = Action(a.stage, T.RECV_F, a.mb)
            if rv in local[r]:
                return arrival.get(rv)
            return end_of.get(Action(a.stage - 1, T.F, a.mb))
        if a.type in (T.B, T.I):
            own_f = end_of.get(Action(a.stage, T.F, a.mb))
            if own_f is None:
                return None
            if a.stage == num_stages - 1:
                return own_f
            rv = Action(a.stage, T.RECV_B, a.mb)
            if rv in local[r]:
                t = arrival.get(rv)
            else:
                t = end_of.get(Action(a.stage + 1, T.B, a.mb), end_of.get(Action(a.stage + 1, T.I, a.mb)))
            return None if t is None else max(t, own_f)
        if a.type == T.W:
            return end_of.get(Action(a.stage, T.I, a.mb))
        return 0.0

    remaining = sum(len(order[r]) for r in ranks)
    while remaining:
        progressed = False
        for r in ranks:
            while idx[r] < len(order[r]):
                a = order[r][idx[r]]
                if a.type in (T.SEND_F, T.SEND_B):
                    src = Action(a.stage, T.F, a.mb) if a.type == T.SEND_F else Action(a.stage, T.B, a.mb)
                    t = end_of.get(src, end_of.get(Action(a.stage, T.I, a.mb)))
                    if t is None:
                        raise DeadlockError(f"rank {r}: {a} precedes the compute that produces it")
                    peer = Action(a.stage + 1, T.RECV_F, a.mb) if a.type == T.SEND_F else Action(a.stage - 1, T.RECV_B, a.mb)
                    arrival[peer] = t + comm_latency
                elif a.type in (T.RECV_F, T.RECV_B):
                    pass  # posting a receive never blocks; the consumer waits for `arrival`
                elif a.type.is_compute:
                    t_in = input_ready(r, a)
                    if t_in is None:
                        break
                    start = max(clock[r], t_in)
                    d = dur(a)
                    end = start + d
                    clock[r] = end
                    busy[r] += d
                    end_of[a] = end
                    timeline[r].append((a, start, end))
                else:  # REDUCE_GRAD / ALLREDUCE / OPTIM
                    d = dur(a)
                    start = clock[r]
                    clock[r] = start + d
                    if d > 0:
                        timeline[r].append((a, start, start + d))
                idx[r] += 1
                remaining -= 1
                progressed = True
        if not progressed:
            stuck = {r: str(order[r][idx[r]]) for r in ranks if idx[r] < len(order[r])}
            raise DeadlockError(f"schedule deadlocks; blocked actions: {stuck}")
    makespan = max(clock.values()) if clock else 0.0
    bubble = {r: (1.0 - busy[r] / makespan) if makespan > 0 else 0.0 for r in ranks}
    return SimResult(makespan, busy, bubble, timeline)
