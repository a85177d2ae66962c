"""Pipeline schedule IR, generators (GPipe / 1F1B / Interleaved-1F1B), comm lowering, validation, simulator."""
from .ir import Action, ActionType, format_schedule, parse_schedule  # noqa: F401
from .generate import (SCHEDULES, EXTRA_SCHEDULES, zero_bubble_order, zero_bubble_interleaved_order, split_bw, merge_bw, overlap_f_b, expand_overlap, build_schedule, gpipe_order, one_f_one_b_order, interleaved_order,  # noqa: F401
                       add_reduce_grad, add_send_recv, add_optimizer, add_unshard_reshard, strip_fsdp_slots, stage_to_rank_loop, validate_schedule,
                       theoretical_bubble)
from .simulate import simulate, SimResult  # noqa: F401
