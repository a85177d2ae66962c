"""Schedule IR: one ``Action`` = (stage index, action type, micro-batch index).

Text form is ``"{stage}{TYPE}{microbatch}"`` -- ``2F0``, ``3SEND_B1``, ``4REDUCE_GRAD`` -- the same strings that
torch's pipelining IR prints (torch:distributed/pipelining/schedules.py:88-113), so
golden vectors and CSV dumps are interchangeable.  ``ALLREDUCE`` (data-parallel gradient all-reduce) and
``OPTIM`` (fused optimizer step) are additions of this framework; the reference has neither
(/root/reference/LLMsDistributedTrainingHelper.py:113-131).
"""
from __future__ import annotations

import enum
import re
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional, Sequence


class ActionType(enum.Enum):
    F = "F"                      # forward of one micro-batch on one stage
    B = "B"                      # full backward (input grad + weight grad)
    I = "I"                      # input-grad only   (zero-bubble style split; executable, unused by the 3 schedules)
    W = "W"                      # weight-grad only
    SEND_F = "SEND_F"
    RECV_F = "RECV_F"
    SEND_B = "SEND_B"
    RECV_B = "RECV_B"
    REDUCE_GRAD = "REDUCE_GRAD"  # grads / m   (folded into the optimizer kernel on CUDA)
    ALLREDUCE = "ALLREDUCE"      # data-parallel all-reduce of the stage's gradient arena
    OPTIM = "OPTIM"              # fused AdamW step of the stage
    UNSHARD = "UNSHARD"          # FSDP parameter prefetch slot (placeholder: no FSDP in this framework)
    RESHARD = "RESHARD"          # FSDP release slot
    OVERLAP_F_B = "OVERLAP_F_B"  # a forward and a backward of different micro-batches issued as one unit (sub-actions)

    @property
    def is_compute(self) -> bool:
        return self in (ActionType.F, ActionType.B, ActionType.I, ActionType.W)

    @property
    def is_comm(self) -> bool:
        return self in (ActionType.SEND_F, ActionType.RECV_F, ActionType.SEND_B, ActionType.RECV_B)


_NO_MB = (ActionType.REDUCE_GRAD, ActionType.ALLREDUCE, ActionType.OPTIM, ActionType.UNSHARD, ActionType.RESHARD)
# longest names first so that "SEND_F" is not parsed as "S..."
_TYPE_RE = "|".join(sorted((t.value for t in ActionType), key=len, reverse=True))
_ACTION_RE = re.compile(rf"^(\d+)({_TYPE_RE})(\d*)$")


@dataclass(frozen=True)
class Action:
    stage: int
    type: ActionType
    mb: Optional[int] = None
    sub: Optional[tuple] = None     # OVERLAP_F_B only: the (F, B) actions it groups (torch _Action.sub_actions)

    def __str__(self) -> str:
        if self.sub is not None:    # torch prints "(2F1;2B0)OVERLAP_F_B" (schedules.py:88-113)
            return "(" + ";".join(str(a) for a in self.sub) + ")" + self.type.value
        return f"{self.stage}{self.type.value}{'' if self.mb is None else self.mb}"

    __repr__ = __str__

    @staticmethod
    def from_str(s: str) -> "Action":
        s = s.strip()
        if s.startswith("(") and s.endswith(")" + ActionType.OVERLAP_F_B.value):
            subs = tuple(Action.from_str(t) for t in s[1:-len(ActionType.OVERLAP_F_B.value) - 1].split(";") if t.strip())
            if len(subs) < 2:
                raise ValueError(f"OVERLAP_F_B needs at least two sub-actions: {s!r}")
            return Action(subs[0].stage, ActionType.OVERLAP_F_B, subs[0].mb, subs)
        m = _ACTION_RE.match(s)
        if not m:
            raise ValueError(f"cannot parse schedule action {s!r}")
        stage, typ, mb = int(m.group(1)), ActionType(m.group(2)), m.group(3)
        if typ in _NO_MB:
            if mb != "":
                raise ValueError(f"{typ.value} takes no micro-batch index: {s!r}")
            return Action(stage, typ, None)
        if mb == "":
            raise ValueError(f"{typ.value} needs a micro-batch index: {s!r}")
        return Action(stage, typ, int(mb))


def F(stage: int, mb: int) -> Action:
    return Action(stage, ActionType.F, mb)


def B(stage: int, mb: int) -> Action:
    return Action(stage, ActionType.B, mb)


def format_schedule(per_rank: Dict[int, Sequence[Action]]) -> str:
    """One line per rank: ``rank0: 0F0 0SEND_F0 ...``."""
    return "\n".join(f"rank{r}: " + " ".join(str(a) for a in per_rank[r]) for r in sorted(per_rank))


def parse_schedule(text: str) -> Dict[int, List[Action]]:
    out: Dict[int, List[Action]] = {}
    cur: Optional[int] = None
    for line in text.strip().splitlines():
        line = line.strip()
        if not line:
            continue
        m = re.match(r"^rank(\d+):\s*(.*)$", line)
        if m:
            cur = int(m.group(1))
            out[cur] = []
            line = m.group(2)
        if cur is None:
            raise ValueError("schedule text must start with 'rankN:'")
        out[cur].extend(Action.from_str(tok) for tok in line.split())
    return out


def to_csv(per_rank: Dict[int, Sequence[Action]]) -> str:
    """One row per rank, comma separated (same shape as torch's _dump_csv, schedules.py:1454-1837)."""
    return "\n".join(",".join(str(a) for a in per_rank[r]) for r in sorted(per_rank)) + "\n"


def from_csv(text: str) -> Dict[int, List[Action]]:
    return {r: [Action.from_str(t) for t in row.split(",") if t.strip()]
            for r, row in enumerate(l for l in text.strip().splitlines() if l.strip())}


def grid(per_rank: Dict[int, Sequence[Action]], error_at: Optional[tuple] = None, compute_only: bool = False) -> str:
    """Pretty-print the schedule as a rank x step grid; ``error_at=(rank, index)`` marks the failing action
    (torch prints the same '<-- ERROR HERE' marker, schedules.py:236-240)."""
    rows = []
    width = max((len(str(a)) for acts in per_rank.values() for a in acts), default=4) + 1
    for r in sorted(per_rank):
        cells = []
        for i, a in enumerate(per_rank[r]):
            if compute_only and not a.type.is_compute:
                continue
            s = str(a)
            if error_at == (r, i):
                s += " <-- ERROR HERE"
            cells.append(s.ljust(width))
        rows.append(f"rank{r:<3}| " + "".join(cells))
    return "\n".join(rows)
