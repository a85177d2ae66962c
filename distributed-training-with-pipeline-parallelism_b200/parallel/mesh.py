"""(dp, pp, tp) device mesh on top of torch.distributed process groups.

Rank layout: ``rank = (dp_idx * pp_size + pp_idx) * tp_size + tp_idx`` -- tensor-parallel peers are adjacent
(they exchange the most data per step), pipeline neighbours next, data-parallel replicas furthest apart.  On one
NVSwitch box every pair has the same bandwidth, so the layout only matters for readability of rank numbers.
The reference creates one unused group (helper:178); there is no mesh in it."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch.distributed as dist


@dataclass
class Mesh:
    dp_size: int
    pp_size: int
    tp_size: int
    rank: int
    dp_idx: int
    pp_idx: int
    tp_idx: int
    pp_group: Optional[object] = None
    pp_group_bwd: Optional[object] = None   # second communicator over the same ranks (gradient direction)
    dp_group: Optional[object] = None
    tp_group: Optional[object] = None

    @property
    def world_size(self) -> int:
        return self.dp_size * self.pp_size * self.tp_size

    def describe(self) -> str:
        parts = []
        if self.dp_size > 1:
            parts.append(f"dp{self.dp_size}")
        parts.append(f"pp{self.pp_size}")
        if self.tp_size > 1:
            parts.append(f"tp{self.tp_size}")
        return "x".join(parts)


def configure_nccl_for_pipeline() -> None:
    """Call BEFORE ``init_process_group('nccl')``.  Stage-to-stage transfers are 8-16 MiB NCCL send/recv pairs that
    overlap with compute.  Two opposing needs, both measured on 4x B200 (1F1B, flagship model, 32 micro-batches):

    * a posted receive spins on its SMs until the sender arrives, and every SM NCCL holds pushes one CTA of a
      persistent tcgen05 GEMM into a second wave -- so the kernels leave ``PPB_RESERVED_SMS`` SMs free;
    * once the stages are balanced (partition on measured unit times) the steady state of 1F1B has NO slack, so the
      latency of every transfer lands on the critical path once per micro-batch.  With 2 channels a message takes
      ~0.2-0.3 ms and the step was 237.3 ms (critical-rank bubble 0.144 against a bound of 0.086); 8 channels / 8 SMs:
      227.5 ms (0.102); 16 / 16: 225.8 ms (0.094) -- profiles/nccl_p2p_channels_n4.md.
    User-set values win."""
    import os
    os.environ.setdefault("NCCL_MAX_P2P_NCHANNELS", "16")
    os.environ.setdefault("NCCL_MIN_P2P_NCHANNELS", "1")
    os.environ.setdefault("PPB_RESERVED_SMS", "16")


def rank_of(dp: int, pp: int, tp: int, pp_size: int, tp_size: int) -> int:
    return (dp * pp_size + pp) * tp_size + tp


def coords_of(rank: int, pp_size: int, tp_size: int):
    tp = rank % tp_size
    pp = (rank // tp_size) % pp_size
    dp = rank // (tp_size * pp_size)
    return dp, pp, tp


def group_ranks(dp_size: int, pp_size: int, tp_size: int):
    """All (pp, dp, tp) groups as lists of global ranks -- pure function, used by the tests."""
    pp_groups = [[rank_of(d, p, t, pp_size, tp_size) for p in range(pp_size)] for d in range(dp_size) for t in range(tp_size)]
    dp_groups = [[rank_of(d, p, t, pp_size, tp_size) for d in range(dp_size)] for p in range(pp_size) for t in range(tp_size)]
    tp_groups = [[rank_of(d, p, t, pp_size, tp_size) for t in range(tp_size)] for d in range(dp_size) for p in range(pp_size)]
    return pp_groups, dp_groups, tp_groups


def init_mesh(dp_size: int = 1, pp_size: Optional[int] = None, tp_size: int = 1) -> Mesh:
    """Create the process groups (every rank must call this; new_group is collective)."""
    world = dist.get_world_size()
    rank = dist.get_rank()
    if pp_size is None:
        pp_size = world // (dp_size * tp_size)
    if dp_size * pp_size * tp_size != world:
        raise ValueError(f"mesh {dp_size}x{pp_size}x{tp_size} does not match world size {world}")
    dp, pp, tp = coords_of(rank, pp_size, tp_size)
    mesh = Mesh(dp_size, pp_size, tp_size, rank, dp, pp, tp)
    pp_groups, dp_groups, tp_groups = group_ranks(dp_size, pp_size, tp_size)
    if dp_size * tp_size == 1:
        mesh.pp_group = None  # the default group
    else:
        for ranks in pp_groups:
            g = dist.new_group(ranks)
            if rank in ranks:
                mesh.pp_group = g
    if pp_size > 1:
        for ranks in pp_groups:
            g = dist.new_group(ranks)
            if rank in ranks:
                mesh.pp_group_bwd = g
    if dp_size > 1:
        for ranks in dp_groups:
            g = dist.new_group(ranks)
            if rank in ranks:
                mesh.dp_group = g
    if tp_size > 1:
        for ranks in tp_groups:
            g = dist.new_group(ranks)
            if rank in ranks:
                mesh.tp_group = g
    return mesh
