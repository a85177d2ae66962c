"""Schedule generators and lowering passes (pure Python, no process group needed).

    compute order  ->  + REDUCE_GRAD  ->  (+ ALLREDUCE)  ->  + SEND/RECV  ->  (+ OPTIM)  ->  validate

Behavioural spec: GPipe = torch schedules.py:684-800 (T4), 1F1B = :803-1044 (T5),
Interleaved-1F1B = :2351-2618 (T10), lowering passes = :1047-1451 (T7).  The reference picks between them by
name, /root/reference/LLMsDistributedTrainingHelper.py:215-220.  Everything here is an independent
implementation: the generators are closed-form, the comm lowering is a round-robin list scheduler in which the
SENDER places the receiver's RECV (that is what makes the resulting per-rank orders deadlock-free).
"""
from __future__ import annotations

from collections import defaultdict
from typing import Callable, Dict, List, Optional, Sequence

from .ir import Action, ActionType as T

Order = Dict[int, List[Action]]

SCHEDULES = ("GPipe", "1F1B", "Interleaved1F1B")   # the reference's three (helper:215-219)
EXTRA_SCHEDULES = ("ZeroBubble1F1B", "ZeroBubbleInterleaved")


def stage_to_rank_loop(pp_size: int) -> Callable[[int], int]:
    """'Loop' placement used by the reference for virtual stages: stage s lives on rank s % P
    (helper:208 ``stage_idx = rank + world_size * i``; torch schedules.py T8)."""
    return lambda s: s % pp_size


# ------------------------------------------------------------------------------------------------------------
# compute orders
# ------------------------------------------------------------------------------------------------------------
def gpipe_order(pp_size: int, n_microbatches: int) -> Order:
    """All forwards, then all backwards in micro-batch order 0..m-1 (torch ScheduleGPipe runs the backward loop
    in the same order as the forward loop)."""
    m = n_microbatches
    return {r: [Action(r, T.F, i) for i in range(m)] + [Action(r, T.B, i) for i in range(m)] for r in range(pp_size)}


def one_f_one_b_order(pp_size: int, n_microbatches: int, extra_warmup: int = 0) -> Order:
    """Warm-up of min(m, P - rank) forwards, then strictly alternating 1B-1F, then the backward drain.

    ``extra_warmup`` = additional forwards every rank but the last keeps in flight ("eager" 1F1B).  With perfectly
    balanced stages the torch warm-up count leaves zero slack in the steady state, so every stage-to-stage transfer
    latency lands on the critical path once per micro-batch; one more forward in flight gives the neighbours' messages
    a whole action of slack at the price of one more stashed micro-batch per rank.  0 reproduces torch exactly."""
    m = n_microbatches
    if m < pp_size:
        raise ValueError(f"1F1B needs n_microbatches ({m}) >= number of stages ({pp_size})")
    order: Order = {}
    for r in range(pp_size):
        warm = min(m, pp_size - r + (extra_warmup if r < pp_size - 1 else 0))
        acts = [Action(r, T.F, i) for i in range(warm)]
        f, b = warm, 0
        while b < m:
            acts.append(Action(r, T.B, b))
            b += 1
            if f < m:
                acts.append(Action(r, T.F, f))
                f += 1
        order[r] = acts
    return order


def interleaved_order(pp_size: int, n_virtual: int, n_microbatches: int) -> Order:
    """Depth-first interleaved 1F1B with ``n_virtual`` stages per rank placed loop-wise.

    rounds = max(1, m // P); micro-batches are processed in groups of m // rounds per virtual stage.
    Rank r warms up with (v-1)*mb_per_round + 2*(P-1-r) forwards (capped at m*v), then runs (F, B) pairs and
    drains the remaining backwards."""
    P, v, m = pp_size, n_virtual, n_microbatches
    rounds = max(1, m // P)
    if m % rounds != 0:
        raise ValueError(f"Interleaved1F1B needs n_microbatches ({m}) divisible by the number of rounds ({rounds})")
    per_round = m // rounds
    total = m * v
    order: Order = {}
    for r in range(P):
        warm = min((v - 1) * per_round + 2 * (P - 1 - r), total)
        fwd_mb: Dict[int, int] = defaultdict(int)
        bwd_mb: Dict[int, int] = defaultdict(int)

        def fwd_stage(step: int) -> int:
            return ((step // per_round) % v) * P + r

        def bwd_stage(step: int) -> int:
            return (v - 1 - ((step - warm) // per_round) % v) * P + r

        acts: List[Action] = []

        def emit_f(step: int) -> None:
            s = fwd_stage(step)
            acts.append(Action(s, T.F, fwd_mb[s]))
            fwd_mb[s] += 1

        def emit_b(step: int) -> None:
            s = bwd_stage(step)
            acts.append(Action(s, T.B, bwd_mb[s]))
            bwd_mb[s] += 1

        for step in range(warm):
            emit_f(step)
        for step in range(warm, total):
            emit_f(step)
            emit_b(step)
        for step in range(total, total + warm):
            emit_b(step)
        order[r] = acts
    return order


def zero_bubble_order(pp_size: int, n_microbatches: int, max_in_flight: Optional[int] = None,
                      max_deferred_w: Optional[int] = None) -> Order:
    """ZB-H1-style schedule: 1F1B with every backward split into I (input gradient, on the critical path between
    ranks) and W (weight gradient, local), the W's greedily moved into what would otherwise be idle slots.

    Built by list-scheduling on a unit-cost clock (F = I = W = 1): each tick a rank runs, in priority order, a W when
    ``max_deferred_w`` (default ``pp_size``) weight-gradient passes are already waiting, else a ready I, else a ready
    F while fewer than ``max_in_flight`` (default ``pp_size - rank``, the 1F1B bound on full activation stashes, F
    until I) micro-batches are live, else a pending W, else it idles.  A deferred W keeps the GEMM operands of its
    micro-batch (layer inputs + output gradients) alive, so ``max_deferred_w`` is the memory knob.  Not one of the reference's
    three schedules (helper:215-219); it exercises the I/W actions of the IR end to end (SURVEY.md section 2.2 T9/T13
    "split I/W variants")."""
    p, m = pp_size, n_microbatches
    f_done: Dict[tuple, int] = {}    # (stage, mb) -> tick at which the result is available
    i_done: Dict[tuple, int] = {}
    nf = [0] * p
    ni = [0] * p
    nw = [0] * p
    order: Order = {r: [] for r in range(p)}
    t = 0
    while any(nw[r] < m for r in range(p)):
        picks = []
        for r in range(p):
            cap = max_in_flight if max_in_flight is not None else p - r
            wcap = max_deferred_w if max_deferred_w is not None else p
            k = ni[r]
            i_ready = (k < m and f_done.get((r, k), 1 << 60) <= t and (r == p - 1 or i_done.get((r + 1, k), 1 << 60) <= t))
            k2 = nf[r]
            f_ready = (k2 < m and (r == 0 or f_done.get((r - 1, k2), 1 << 60) <= t) and nf[r] - ni[r] < cap)
            if ni[r] - nw[r] >= wcap:
                picks.append((r, T.W, nw[r]))
            elif i_ready:
                picks.append((r, T.I, k))
            elif f_ready:
                picks.append((r, T.F, k2))
            elif nw[r] < ni[r]:
                picks.append((r, T.W, nw[r]))
        for r, typ, k in picks:
            order[r].append(Action(r, typ, k))
            if typ == T.F:
                f_done[(r, k)] = t + 1
                nf[r] += 1
            elif typ == T.I:
                i_done[(r, k)] = t + 1
                ni[r] += 1
            else:
                nw[r] += 1
        t += 1
        if t > 8 * (m + p) + 64:
            raise RuntimeError("zero_bubble_order did not converge")
    return order


def zero_bubble_interleaved_order(pp_size: int, n_virtual: int, n_microbatches: int,
                                 max_deferred_w: Optional[int] = None) -> Order:
    """Split-backward list schedule for ``n_virtual`` stages per rank (loop placement, stage = chunk * pp_size + rank):
    the interleaved-1F1B visiting order (rounds of ``pp_size`` micro-batches per chunk, forward chunks ascending,
    backward chunks descending) with every backward split into I and W and the W's filling idle ticks.

    Same construction as ``zero_bubble_order``: one action per rank and unit-cost tick, priorities  forced W (more than
    ``max_deferred_w`` waiting, default ``pp_size * n_virtual``)  >  ready I  >  ready F while the rank holds fewer
    full activation stashes than interleaved 1F1B would (its warm-up count + 1)  >  pending W.  Every action is issued
    only after its producers finished in an earlier tick, so the order cannot deadlock."""
    p, v, m = pp_size, n_virtual, n_microbatches
    S = p * v
    done_f: Dict[tuple, int] = {}
    done_i: Dict[tuple, int] = {}
    nf = {s: 0 for s in range(S)}          # next micro-batch index per stage, per action type (in-order per stage)
    ni = {s: 0 for s in range(S)}
    nw = {s: 0 for s in range(S)}
    order: Order = {r: [] for r in range(p)}
    wcap = max_deferred_w if max_deferred_w is not None else p * v
    BIG = 1 << 60
    t = 0
    while any(nw[s] < m for s in range(S)):
        picks = []
        for r in range(p):
            mine = [k * p + r for k in range(v)]
            cap = min((v - 1) * p + 2 * (p - 1 - r), m * v) + 1
            live = sum(nf[s] - ni[s] for s in mine)
            waiting = sum(ni[s] - nw[s] for s in mine)
            # the next forward / input-gradient of a rank is fixed by the interleaved visiting order (no skipping ahead:
            # a rank that ran later rounds of chunk 0 early would fill its stash budget and starve the deeper chunks)
            todo_i = [s for s in mine if ni[s] < m]
            todo_f = [s for s in mine if nf[s] < m]
            ready_i, ready_f = [], []
            if todo_i:
                s = min(todo_i, key=lambda s: (ni[s] // p, -(s // p), ni[s]))          # backward: last chunk first
                if done_f.get((s, ni[s]), BIG) <= t and (s == S - 1 or done_i.get((s + 1, ni[s]), BIG) <= t):
                    ready_i = [s]
            if todo_f:
                s = min(todo_f, key=lambda s: (nf[s] // p, s // p, nf[s]))             # forward: first chunk first
                if s == 0 or done_f.get((s - 1, nf[s]), BIG) <= t:
                    ready_f = [s]
            pending_w = [s for s in mine if nw[s] < ni[s]]
            if waiting >= wcap and pending_w:
                s = min(pending_w, key=lambda s: (nw[s] // p, -(s // p), nw[s]))
                picks.append((r, T.W, s, nw[s]))
            elif ready_i:
                s = min(ready_i, key=lambda s: (ni[s] // p, -(s // p), ni[s]))      # backward: last chunk first
                picks.append((r, T.I, s, ni[s]))
            elif ready_f and live < cap:
                s = min(ready_f, key=lambda s: (nf[s] // p, s // p, nf[s]))         # forward: first chunk first
                picks.append((r, T.F, s, nf[s]))
            elif pending_w:
                s = min(pending_w, key=lambda s: (nw[s] // p, -(s // p), nw[s]))
                picks.append((r, T.W, s, nw[s]))
        for r, typ, s, k in picks:
            order[r].append(Action(s, typ, k))
            if typ == T.F:
                done_f[(s, k)] = t + 1
                nf[s] += 1
            elif typ == T.I:
                done_i[(s, k)] = t + 1
                ni[s] += 1
            else:
                nw[s] += 1
        t += 1
        if t > 16 * (m * v + p) + 256:
            raise RuntimeError("zero_bubble_interleaved_order did not converge")
    return order


# ------------------------------------------------------------------------------------------------------------
# lowering passes
# ------------------------------------------------------------------------------------------------------------
def split_bw(order: Order) -> Order:
    """Replace every full backward ``B`` by the adjacent pair ``I``, ``W``."""
    return {r: [x for a in acts for x in ((Action(a.stage, T.I, a.mb), Action(a.stage, T.W, a.mb)) if a.type == T.B else (a,))]
            for r, acts in order.items()}


def overlap_f_b(order: Order) -> Order:
    """Group every adjacent (F of one micro-batch, B of another) pair of the same rank into one ``OVERLAP_F_B`` action
    (the unit torch's dual-pipe schedules emit, schedules.py:45-113).  Apply to compute-only orders."""
    out: Order = {}
    for r, acts in order.items():
        res: List[Action] = []
        i = 0
        while i < len(acts):
            a = acts[i]
            nxt = acts[i + 1] if i + 1 < len(acts) else None
            if nxt is not None and {a.type, nxt.type} == {T.F, T.B} and a.mb != nxt.mb:
                res.append(Action(a.stage, T.OVERLAP_F_B, a.mb, (a, nxt)))
                i += 2
            else:
                res.append(a)
                i += 1
        out[r] = res
    return out


def expand_overlap(order: Order) -> Order:
    """Replace every ``OVERLAP_F_B`` by its sub-actions, in order (the runner drives one compute stream per rank, so
    the grouping is an IR-level statement; lowering, validation and simulation work on the expanded lists)."""
    return {r: [x for a in acts for x in (a.sub if a.sub is not None else (a,))] for r, acts in order.items()}


def merge_bw(order: Order) -> Order:
    """Fuse an ``I`` immediately followed by the ``W`` of the same stage and micro-batch back into one ``B``
    (torch _merge_bw, schedules.py:1204): the fused backward shares loads between dgrad and wgrad and skips the
    deferral bookkeeping."""
    out: Order = {}
    for r, acts in order.items():
        res: List[Action] = []
        i = 0
        while i < len(acts):
            a = acts[i]
            if (a.type == T.I and i + 1 < len(acts) and acts[i + 1].type == T.W and acts[i + 1].stage == a.stage
                    and acts[i + 1].mb == a.mb):
                res.append(Action(a.stage, T.B, a.mb))
                i += 2
            else:
                res.append(a)
                i += 1
        out[r] = res
    return out


def add_reduce_grad(order: Order, n_microbatches: int, also_allreduce: bool = False) -> Order:
    """Insert ``{s}REDUCE_GRAD`` right after a stage's m-th backward (B or W) -- and, for data-parallel replicas,
    ``{s}ALLREDUCE`` right behind it, which is what lets the gradient all-reduce of the later stages hide inside
    the 1F1B cool-down of the earlier ones."""
    out: Order = {}
    for r, acts in order.items():
        seen: Dict[int, int] = defaultdict(int)
        res: List[Action] = []
        for a in acts:
            res.append(a)
            if a.type in (T.B, T.W):
                seen[a.stage] += 1
                if seen[a.stage] == n_microbatches:
                    res.append(Action(a.stage, T.REDUCE_GRAD))
                    if also_allreduce:
                        res.append(Action(a.stage, T.ALLREDUCE))
        out[r] = res
    return out


def add_unshard_reshard(order: Order, max_active_stages: int = 3) -> Order:
    """Insert UNSHARD / RESHARD slots with a look-ahead residency policy: before each compute action the next
    ``max_active_stages`` distinct stages must be resident; resident stages outside that window are released
    first, everything is released at the end.  Without FSDP the slots do nothing at run time, but they take a
    turn in the comm-lowering round-robin, which keeps the lowered order string-identical to the golden vectors
    of SURVEY.md section 3.5 (torch inserts the same slots, schedules.py:1084-1164)."""
    out: Order = {}
    for r, acts in order.items():
        active: List[int] = []
        res: List[Action] = []
        for i, a in enumerate(acts):
            window: List[int] = []
            for nxt in acts[i:]:
                if nxt.stage not in window:
                    window.append(nxt.stage)
                    if len(window) == max_active_stages:
                        break
            for s in [s for s in active if s not in window]:
                active.remove(s)
                res.append(Action(s, T.RESHARD))
            for s in window:
                if s not in active:
                    active.append(s)
                    res.append(Action(s, T.UNSHARD))
            res.append(a)
        for s in list(active):
            res.append(Action(s, T.RESHARD))
        out[r] = res
    return out


def strip_fsdp_slots(order: Order) -> Order:
    return {r: [a for a in acts if a.type not in (T.UNSHARD, T.RESHARD)] for r, acts in order.items()}


def add_optimizer(order: Order) -> Order:
    """Append one ``{s}OPTIM`` per local stage at the end of each rank's list."""
    out: Order = {}
    for r, acts in order.items():
        stages = sorted({a.stage for a in acts if a.type in (T.F, T.B, T.I, T.W)})
        out[r] = list(acts) + [Action(s, T.OPTIM) for s in stages]
    return out


def add_send_recv(order: Order, stage_to_rank: Callable[[int], int], num_stages: int) -> Order:
    """Insert SEND/RECV actions.  Ranks are visited round-robin; a rank may emit its next compute action only
    when the data it consumes has been produced (its RECV is already in its list, or the producing stage is
    local).  When an action with a remote consumer is emitted, its SEND goes right behind it and the matching
    RECV is appended to the CONSUMER's list at that moment.  Adjacent stages on the same rank exchange tensors
    directly (no comm actions)."""
    pending = {r: list(a) for r, a in order.items() if a}
    out: Order = {r: [] for r in order}
    done: Dict[int, set] = {r: set() for r in order}

    def remote_consumer(a: Action) -> Optional[Action]:
        if a.type == T.F and a.stage != num_stages - 1 and stage_to_rank(a.stage + 1) != stage_to_rank(a.stage):
            return Action(a.stage + 1, T.RECV_F, a.mb)
        if a.type in (T.B, T.I) and a.stage != 0 and stage_to_rank(a.stage - 1) != stage_to_rank(a.stage):
            return Action(a.stage - 1, T.RECV_B, a.mb)
        return None

    def ready(a: Action, have: set) -> bool:
        if a.type == T.F and a.stage != 0:
            return Action(a.stage, T.RECV_F, a.mb) in have or Action(a.stage - 1, T.F, a.mb) in have
        if a.type in (T.B, T.I) and a.stage != num_stages - 1:
            return (Action(a.stage, T.RECV_B, a.mb) in have or Action(a.stage + 1, T.B, a.mb) in have
                    or Action(a.stage + 1, T.I, a.mb) in have)
        return True

    while pending:
        progress = False
        for r in sorted(pending):
            a = pending[r][0]
            if not ready(a, done[r]):
                continue
            out[r].append(a)
            done[r].add(a)
            recv = remote_consumer(a)
            if recv is not None:
                send = Action(a.stage, T.SEND_F if a.type == T.F else T.SEND_B, a.mb)
                out[r].append(send)
                done[r].add(send)
                dst = stage_to_rank(recv.stage)
                out[dst].append(recv)
                done[dst].add(recv)
            pending[r].pop(0)
            if not pending[r]:
                del pending[r]
            progress = True
        if not progress:
            raise RuntimeError("malformed compute schedule: no rank can make progress while placing sends/recvs\n"
                               + "\n".join(f"rank{r}: next={pending[r][0]}" for r in sorted(pending)))
    return out


# ------------------------------------------------------------------------------------------------------------
# validation + theory
# ------------------------------------------------------------------------------------------------------------
def validate_schedule(order: Order, num_stages: int, n_microbatches: int,
                      stage_to_rank: Optional[Callable[[int], int]] = None) -> None:
    """Raise ValueError unless every stage runs every micro-batch exactly once forward and once backward, on the
    right rank, forward before backward, and every SEND has exactly one matching RECV on the consumer's rank."""
    f_seen: Dict[tuple, int] = {}
    b_seen: Dict[tuple, int] = {}
    sends, recvs = set(), set()
    for r, acts in order.items():
        pos: Dict[Action, int] = {}
        for i, a in enumerate(acts):
            if a in pos and a.type not in (T.REDUCE_GRAD, T.ALLREDUCE, T.OPTIM):
                raise ValueError(f"rank {r}: duplicate action {a}")
            pos[a] = i
            if stage_to_rank is not None and stage_to_rank(a.stage) != r:
                raise ValueError(f"rank {r}: action {a} belongs to rank {stage_to_rank(a.stage)}")
            if a.type == T.F:
                f_seen[(a.stage, a.mb)] = f_seen.get((a.stage, a.mb), 0) + 1
            elif a.type in (T.B, T.I):
                b_seen[(a.stage, a.mb)] = b_seen.get((a.stage, a.mb), 0) + 1
                if Action(a.stage, T.F, a.mb) not in pos:
                    raise ValueError(f"rank {r}: {a} before its forward")
            elif a.type == T.W:
                if Action(a.stage, T.I, a.mb) not in pos:
                    raise ValueError(f"rank {r}: {a} before its input-gradient pass")
            elif a.type in (T.SEND_F, T.SEND_B):
                sends.add((a.type, a.stage, a.mb))
                src = Action(a.stage, T.F if a.type == T.SEND_F else T.B, a.mb)
                if src not in pos and Action(a.stage, T.I, a.mb) not in pos:
                    raise ValueError(f"rank {r}: {a} before the compute that produces it")
            elif a.type in (T.RECV_F, T.RECV_B):
                recvs.add((a.type, a.stage, a.mb))
        for a in pos:
            if a.type == T.I and Action(a.stage, T.W, a.mb) not in pos:
                raise ValueError(f"rank {r}: {a} has no weight-gradient pass")
        for a, i in pos.items():  # a consumer must come after its RECV
            if a.type == T.F and Action(a.stage, T.RECV_F, a.mb) in pos and pos[Action(a.stage, T.RECV_F, a.mb)] > i:
                raise ValueError(f"rank {r}: {a} before its RECV_F")
            if a.type in (T.B, T.I) and Action(a.stage, T.RECV_B, a.mb) in pos and pos[Action(a.stage, T.RECV_B, a.mb)] > i:
                raise ValueError(f"rank {r}: {a} before its RECV_B")
    for s in range(num_stages):
        for mb in range(n_microbatches):
            if f_seen.get((s, mb), 0) != 1:
                raise ValueError(f"stage {s} micro-batch {mb}: {f_seen.get((s, mb), 0)} forwards (want 1)")
            if b_seen.get((s, mb), 0) != 1:
                raise ValueError(f"stage {s} micro-batch {mb}: {b_seen.get((s, mb), 0)} backwards (want 1)")
    for typ, s, mb in sends:
        want = (T.RECV_F, s + 1, mb) if typ == T.SEND_F else (T.RECV_B, s - 1, mb)
        if want not in recvs:
            raise ValueError(f"{s}{typ.value}{mb} has no matching receive")
    if len(sends) != len(recvs):
        raise ValueError(f"{len(sends)} sends vs {len(recvs)} receives")


def theoretical_bubble(pp_size: int, n_microbatches: int, n_virtual: int = 1) -> float:
    """(p-1)/(v*m+p-1): GPipe and 1F1B have v = 1 (BASELINE.json bound)."""
    p, m, v = pp_size, n_microbatches, n_virtual
    return (p - 1) / (v * m + p - 1)


def build_schedule(name: str, pp_size: int, n_microbatches: int, n_virtual: int = 1, dp_allreduce: bool = False,
                   with_optimizer: bool = True, keep_fsdp_slots: bool = False, extra_warmup: int = 0) -> Order:
    """Full lowering for one of 'GPipe' | '1F1B' | 'Interleaved1F1B' (names of helper:215-219)."""
    if name == "GPipe":
        if n_virtual != 1:
            raise ValueError("GPipe runs one stage per rank")
        order = gpipe_order(pp_size, n_microbatches)
    elif name == "1F1B":
        if n_virtual != 1:
            raise ValueError("1F1B runs one stage per rank")
        order = one_f_one_b_order(pp_size, n_microbatches, extra_warmup)
    elif name == "Interleaved1F1B":
        order = interleaved_order(pp_size, n_virtual, n_microbatches)
    elif name == "ZeroBubble1F1B":
        if n_virtual != 1:
            raise ValueError("ZeroBubble1F1B runs one stage per rank")
        order = merge_bw(zero_bubble_order(pp_size, n_microbatches))
    elif name == "ZeroBubbleInterleaved":
        order = merge_bw(zero_bubble_interleaved_order(pp_size, n_virtual, n_microbatches))
    else:
        raise ValueError(f"unknown schedule {name!r}; expected one of {SCHEDULES + EXTRA_SCHEDULES}")
    num_stages = pp_size * n_virtual
    s2r = stage_to_rank_loop(pp_size)
    order = add_unshard_reshard(order)
    order = add_reduce_grad(order, n_microbatches, also_allreduce=dp_allreduce)
    order = add_send_recv(order, s2r, num_stages)
    if not keep_fsdp_slots:
        order = strip_fsdp_slots(order)
    if with_optimizer:
        order = add_optimizer(order)
    validate_schedule(order, num_stages, n_microbatches, s2r)
    return order
