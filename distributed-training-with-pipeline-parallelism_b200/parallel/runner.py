"""Schedule runner: interprets a rank's lowered action list on CUDA streams (or on the CPU with gloo).

Behavioural counterpart of torch's ``PipelineScheduleSingle/Multi.step`` + ``_PipelineScheduleRuntime``
(schedules.py:506-630, :1855-2279), which the reference drives from
/root/reference/LLMsDistributedTrainingHelper.py:113-131 (``schedule.step(x)`` on the first rank,
``schedule.step(target=y, losses=losses)`` on the last, ``schedule.step()`` in between).

Design:
  * ONE interpreter for all three schedules: GPipe, 1F1B and Interleaved-1F1B differ only in the action list;
  * no host-blocking waits on CUDA: receives are posted on dedicated anchor streams into a persistent, pre-sized
    ring (one buffer per stage x direction x micro-batch) and consumers wait on CUDA events; sends are issued
    right behind the producing compute and complete asynchronously on NCCL's per-peer streams, so the transfer
    of micro-batch i+1 overlaps the compute of micro-batch i and the critical path carries no exposed copy;
  * every compute action is bracketed by CUDA events -> measured timeline -> measured bubble fraction;
  * gradient averaging over micro-batches and data-parallel replicas is folded into the fused optimizer kernel;
    the optimizer step is part of the schedule (``{s}OPTIM``), i.e. inside the timed region.
"""
from __future__ import annotations

import time
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist

from .schedule import Action, ActionType as T, build_schedule, theoretical_bubble
from .schedule import ir as sched_ir
from .stage import PipelineStage


class PipelineSchedule:
    schedule_name = "1F1B"

    def __init__(self, stages: Union[PipelineStage, Sequence[PipelineStage]], n_microbatches: int, loss_fn=None, *,
                 schedule: Optional[str] = None, pp_group=None, pp_group_bwd=None, dp_group=None,
                 optimizer: Optional[dict] = None,
                 scale_grads: bool = True, record_timeline: bool = True, fused_loss: Optional[bool] = None,
                 cuda_graphs: Optional[bool] = None, graph_warmup_steps: int = 2, actions=None, recompute: bool = False,
                 extra_warmup: Optional[int] = None):
        self.stages: List[PipelineStage] = [stages] if isinstance(stages, PipelineStage) else list(stages)
        if not self.stages:
            raise ValueError("need at least one stage")
        self.name = schedule or self.schedule_name
        self.m = int(n_microbatches)
        self.loss_fn = loss_fn
        # a python loss_fn is honoured through one autograd hop; loss_fn=None or fused_loss=True uses the fused
        # softmax-cross-entropy kernel (mean token CE, identical to helper:197-201)
        self.fused_loss = (loss_fn is None) if fused_loss is None else fused_loss
        self.pp_group = pp_group
        # gradients travel on their own communicator so that a receive posted early in one direction can never
        # sit in front of a send of the other direction on the same NCCL stream
        self.pp_group_bwd = pp_group_bwd if pp_group_bwd is not None else pp_group
        self.dp_group = dp_group
        self.optimizer = optimizer
        self.scale_grads = scale_grads
        self.record_timeline = record_timeline
        if dist.is_available() and dist.is_initialized():
            self.pp_size = dist.get_world_size(pp_group)
            self.pp_rank = dist.get_rank(pp_group)
            self.dp_size = dist.get_world_size(dp_group) if dp_group is not None else 1
        else:
            self.pp_size, self.pp_rank, self.dp_size = 1, 0, 1
        self.n_virtual = len(self.stages)
        self.tp_size = self.stages[0].tp_size
        if any(st.tp_size != self.tp_size for st in self.stages):
            raise ValueError("all stages of a rank must use the same tensor-parallel degree")
        self.num_stages = self.pp_size * self.n_virtual
        for st in self.stages:
            if st.num_stages != self.num_stages:
                raise ValueError(f"stage {st.stage_index} was built for {st.num_stages} stages, schedule has {self.num_stages}")
            if st.stage_index % self.pp_size != self.pp_rank:
                raise ValueError(f"stage {st.stage_index} does not belong to pipeline rank {self.pp_rank} (loop placement)")
        self.local: Dict[int, PipelineStage] = {st.stage_index: st for st in self.stages}
        self.device = self.stages[0].device
        self.cuda = self.device.type == "cuda"
        self.extra_warmup = 0
        if actions is not None:
            order = self._load_actions(actions, optimizer is not None)
            self.name = "custom" if schedule is None else schedule
        else:
            import os
            if extra_warmup is None:   # 1F1B only: forwards kept in flight beyond torch's warm-up count (see one_f_one_b_order)
                extra_warmup = int(os.environ.get("PPB_1F1B_EXTRA_WARMUP", "0"))
            self.extra_warmup = int(extra_warmup)
            order = build_schedule(self.name, self.pp_size, self.m, n_virtual=self.n_virtual,
                                   dp_allreduce=self.dp_size > 1, with_optimizer=optimizer is not None,
                                   extra_warmup=self.extra_warmup)
        self.full_order = order
        self.order: List[Action] = order[self.pp_rank]
        self._retire_in_order = self._retires_in_forward_order(self.order)
        self.theory_bubble = theoretical_bubble(self.pp_size, self.m, self.n_virtual)
        self._shape: Optional[Tuple[int, int]] = None
        self._ring: Dict[Tuple[int, str, int], torch.Tensor] = {}
        self._anchors: Dict[str, Any] = {}
        self.timeline: List[Tuple[Action, Any, Any]] = []
        self._step_span: Optional[Tuple[Any, Any]] = None
        self.steps_done = 0
        for st in self.stages:
            st.materialize()
            if st.dp_group is None:
                st.dp_group = dp_group
        self.fused_dp = False
        if self.dp_size > 1 and optimizer is not None and self.cuda:
            self.fused_dp = all([st.enable_fused_dp(self.m) for st in self.stages if st.n_total > 0])
        self._b_seen: Dict[int, int] = {}
        # CUDA graphs: one forward + one backward graph per (stage, in-flight slot); eager for the first steps
        import os
        want = (os.environ.get("PPB_CUDA_GRAPHS", "1") != "0") if cuda_graphs is None else cuda_graphs
        self.recompute = bool(recompute)    # trade one extra forward per micro-batch for O(1) stashed activations per slot
        for st in self.stages:
            st.recompute = self.recompute
        # slot = mb % nslots is only safe when every stage retires micro-batches in the order it forwarded them
        # (true for all built-in schedules; a user-supplied action list that does not is run without graphs)
        self.use_graphs = bool(want and self.cuda and self.fused_loss and optimizer is not None and self.tp_size == 1
                               and not self.recompute and self._retire_in_order)
        self.graph_warmup_steps = graph_warmup_steps
        self.nvtx = os.environ.get("PPB_NVTX", "0") == "1"   # NVTX range per action, named with its IR string
        if self.use_graphs:
            for st in self.stages:
                st.enable_graphs(self._max_inflight(st.stage_index))

    # ---- helpers ------------------------------------------------------------------------------------------------
    def _load_actions(self, actions, with_optimizer: bool):
        """User-supplied schedule: {rank: [Action | str]}, the text of ``format_schedule`` / ``to_csv``, or a path to
        such a file (torch _PipelineScheduleRuntime._load_actions / _load_csv, schedules.py:1976-2027).  Compute-only
        lists (no SEND/RECV) are lowered here: reduce-grad, comms, optimizer; everything is validated before use."""
        import os
        from .schedule import generate as G
        if isinstance(actions, str):
            text = open(actions).read() if os.path.exists(actions) else actions
            order = sched_ir.parse_schedule(text) if ":" in text else sched_ir.from_csv(text)
        else:
            order = {int(r): [a if isinstance(a, Action) else Action.from_str(a) for a in acts] for r, acts in actions.items()}
        order = G.expand_overlap(order)       # OVERLAP_F_B groups run as their sub-actions, in order
        for r in range(self.pp_size):
            order.setdefault(r, [])
        comm = (T.SEND_F, T.SEND_B, T.RECV_F, T.RECV_B)
        s2r = G.stage_to_rank_loop(self.pp_size)
        if self.pp_size > 1 and not any(a.type in comm for acts in order.values() for a in acts):
            if not any(a.type == T.REDUCE_GRAD for acts in order.values() for a in acts):
                order = G.add_reduce_grad(order, self.m, also_allreduce=self.dp_size > 1)
            order = G.add_send_recv(order, s2r, self.num_stages)
        elif not any(a.type == T.REDUCE_GRAD for acts in order.values() for a in acts):
            order = G.add_reduce_grad(order, self.m, also_allreduce=self.dp_size > 1)
        if with_optimizer and not any(a.type == T.OPTIM for acts in order.values() for a in acts):
            order = G.add_optimizer(order)
        G.validate_schedule(order, self.num_stages, self.m, s2r)
        from .schedule import simulate
        simulate(order, self.num_stages)     # raises DeadlockError on an order that cannot complete
        return order

    @staticmethod
    def _retires_in_forward_order(order: List[Action]) -> bool:
        """True when, per stage, backwards (B or I, and W) come in the same micro-batch order as the forwards."""
        seq: Dict[Tuple[int, str], List[int]] = {}
        for a in order:
            if a.type in (T.F, T.B, T.I, T.W):
                kind = "F" if a.type == T.F else ("W" if a.type == T.W else "B")
                seq.setdefault((a.stage, kind), []).append(a.mb)
        for (stage, kind), mbs in seq.items():
            if kind != "F" and mbs != [m for m in seq.get((stage, "F"), []) if m in set(mbs)]:
                return False
        return True

    def _global_rank(self, pp_rank: int) -> int:
        if self.pp_group is None:
            return pp_rank
        return dist.get_global_rank(self.pp_group, pp_rank)

    def _rank_of_stage(self, stage: int) -> int:
        return stage % self.pp_size

    def _max_inflight(self, stage: int) -> int:
        live = peak = 0
        for a in self.order:
            if a.stage != stage:
                continue
            if a.type == T.F:
                live += 1
                peak = max(peak, live)
            elif a.type in (T.B, T.W):
                live -= 1
        return max(1, peak)

    def _sync_shape(self, x: Optional[torch.Tensor]) -> Tuple[int, int]:
        """One-time (B, S) agreement: the rank that owns stage 0 knows it from its input.  Replaces torch's
        pickle-based per-step shape inference (stage.py:1421-1533)."""
        if self._shape is not None:
            return self._shape
        if self.pp_size == 1:
            if x is None:
                raise ValueError("the single pipeline rank needs the input batch")
            self._shape = (int(x.shape[0]), int(x.shape[1]))
            return self._shape
        dev = self.device if self.cuda else torch.device("cpu")
        t = torch.zeros(2, dtype=torch.int64, device=dev)
        if 0 in self.local:
            if x is None:
                raise ValueError("the rank that owns stage 0 must be given the input batch")
            t[0], t[1] = int(x.shape[0]), int(x.shape[1])
        dist.broadcast(t, src=self._global_rank(self._rank_of_stage(0)), group=self.pp_group)
        self._shape = (int(t[0].item()), int(t[1].item()))
        return self._shape

    def _mb_sizes(self, B: int) -> List[int]:
        """torch.tensor_split semantics (uneven batches allowed; microbatch.py:315-427)."""
        base, rem = divmod(B, self.m)
        return [base + (1 if i < rem else 0) for i in range(self.m)]

    def _ring_buf(self, stage: int, kind: str, mb: int, b: int, S: int) -> torch.Tensor:
        key = (stage, kind, mb)
        buf = self._ring.get(key)
        st = self.local[stage]
        # a boundary behind a self-attention block carries the packed [x1 ; x0] tensor (and its gradient): 2x the rows
        b = b * (st.in_rows if kind == "F" else st.out_rows)
        if buf is None or buf.shape[0] != (b if self.tp_size == 1 else b * S // self.tp_size):
            shape = (b, S, st.args.dim) if self.tp_size == 1 else (b * S // self.tp_size, st.args.dim)
            buf = torch.empty(shape, dtype=st.dtype, device=self.device)
            self._ring[key] = buf
        return buf

    # ---- the step -------------------------------------------------------------------------------------------------
    def step(self, *args, target: Optional[torch.Tensor] = None, losses: Optional[list] = None,
             return_outputs: bool = True):
        """Run one full training step (all micro-batches forward + backward, gradient reduction, optimizer)."""
        x = args[0] if args else None
        B, S = self._sync_shape(x)
        sizes = self._mb_sizes(B)
        x_chunks = list(torch.tensor_split(x, self.m, dim=0)) if (x is not None and 0 in self.local) else None
        last_stage = self.num_stages - 1
        y_chunks = list(torch.tensor_split(target, self.m, dim=0)) if (target is not None and last_stage in self.local) else None
        for st in self.stages:
            st.begin_step()
            st.seq_len = S
        tp = self.tp_size

        def act_view(t, b):   # stage boundary tensors: [b, S, d] ([2b, S, d] when packed), or this tp rank's [b*S/tp, d] rows
            return t.view(-1, S, t.shape[-1]) if tp == 1 else t.view(b * S // tp, -1)

        cuda = self.cuda
        compute = torch.cuda.current_stream(self.device) if cuda else None
        if cuda and not self._anchors:
            self._anchors = {"F": torch.cuda.Stream(self.device), "B": torch.cuda.Stream(self.device)}
        if cuda:
            ev0 = compute.record_event()
            for a in self._anchors.values():
                a.wait_event(ev0)  # ring buffers of the previous step are no longer in use

        recv_state: Dict[Tuple[int, str, int], Any] = {}
        fwd_out: Dict[Tuple[int, int], torch.Tensor] = {}
        bwd_out: Dict[Tuple[int, int], torch.Tensor] = {}
        local_in: Dict[Tuple[int, str, int], torch.Tensor] = {}
        send_works: List[Any] = []
        keep_alive: List[torch.Tensor] = []
        allreduce_works: Dict[int, Any] = {}
        slot_sends: Dict[Tuple[int, str, int], Any] = {}
        # graph slots are captured for ONE micro-batch shape: uneven batches (tensor_split semantics) run eagerly
        graphs_on = (self.use_graphs and self.steps_done >= self.graph_warmup_steps and B % self.m == 0
                     and all(st.training for st in self.stages))
        b_count: Dict[int, int] = {}
        timeline: List[Tuple[Action, Any, Any]] = []
        rec = self.record_timeline

        def tick():
            if not rec:
                return None
            if cuda:
                e = torch.cuda.Event(enable_timing=True)
                e.record(compute)
                return e
            return time.perf_counter()

        t_begin = tick()
        host_t0 = time.perf_counter()
        host_by: Dict[str, float] = {}
        _h0, _ht = host_t0, "begin"
        idx = -1
        clip = None
        nvtx = cuda and self.nvtx
        try:
            for idx, a in enumerate(self.order):
                typ = a.type
                if idx:     # host time spent issuing the previous action, by action type (where does the host block?)
                    _hn = time.perf_counter()
                    host_by[_ht] = host_by.get(_ht, 0.0) + (_hn - _h0)
                    _h0 = _hn
                _ht = typ.value
                if nvtx:   # ranges carry the IR string (2F0, 3SEND_B1, ...) so ncu / nsys timelines line up with the schedule
                    if idx:
                        torch.cuda.nvtx.range_pop()
                    torch.cuda.nvtx.range_push(str(a))
                if typ == T.RECV_F or typ == T.RECV_B:
                    kind = "F" if typ == T.RECV_F else "B"
                    src_stage = a.stage - 1 if kind == "F" else a.stage + 1
                    buf = self._ring_buf(a.stage, kind, a.mb, sizes[a.mb], S)
                    src = self._global_rank(self._rank_of_stage(src_stage))
                    grp = self.pp_group if kind == "F" else self.pp_group_bwd
                    if cuda:
                        anchor = self._anchors[kind]
                        with torch.cuda.stream(anchor):
                            w = dist.irecv(buf, src=src, group=grp)
                            w.wait()  # stream-level: the anchor stream waits, the host does not
                            ev = anchor.record_event()
                        recv_state[(a.stage, kind, a.mb)] = (buf, ev)
                    else:
                        recv_state[(a.stage, kind, a.mb)] = (buf, dist.irecv(buf, src=src, group=grp))
                elif typ == T.SEND_F or typ == T.SEND_B:
                    if typ == T.SEND_F:
                        t = fwd_out.pop((a.stage, a.mb))
                        dst_stage = a.stage + 1
                        grp = self.pp_group
                    else:
                        t = bwd_out.pop((a.stage, a.mb))
                        dst_stage = a.stage - 1
                        grp = self.pp_group_bwd
                    t = act_view(t, sizes[a.mb])
                    dst = self._global_rank(self._rank_of_stage(dst_stage))
                    w = dist.isend(t, dst=dst, group=grp)
                    send_works.append(w)
                    keep_alive.append(t)
                    if graphs_on:
                        slot_sends[(a.stage, "F" if typ == T.SEND_F else "B", a.mb % self.local[a.stage].nslots)] = w
                elif typ == T.F:
                    st = self.local[a.stage]
                    if a.stage == 0:
                        inp = x_chunks[a.mb]
                    else:
                        inp = self._take_input(recv_state, local_in, a.stage, "F", a.mb, compute)
                    tgt = y_chunks[a.mb] if (a.stage == last_stage and y_chunks is not None) else None
                    t0 = tick()
                    if graphs_on:
                        prev = slot_sends.pop((a.stage, "F", a.mb % st.nslots), None)
                        if prev is not None:
                            prev.wait()   # the slot's static output buffer is free again
                        out = st.forward_graphed(a.mb, inp, tgt)
                    else:
                        out = st.forward_one_chunk(a.mb, inp, tgt, 1.0, None if self.fused_loss else self.loss_fn)
                    timeline.append((a, t0, tick()))
                    if a.stage != last_stage:
                        if self._rank_of_stage(a.stage + 1) == self.pp_rank:
                            local_in[(a.stage + 1, "F", a.mb)] = act_view(out, sizes[a.mb])
                        else:
                            fwd_out[(a.stage, a.mb)] = out
                elif typ == T.B:
                    st = self.local[a.stage]
                    dout = None
                    if a.stage != last_stage:
                        dout = self._take_input(recv_state, local_in, a.stage, "B", a.mb, compute)
                    t0 = tick()
                    b_count[a.stage] = b_count.get(a.stage, 0) + 1
                    is_last_b = b_count[a.stage] == self.m
                    if graphs_on:
                        prev = slot_sends.pop((a.stage, "B", a.mb % st.nslots), None)
                        if prev is not None:
                            prev.wait()
                        if is_last_b and st.fused_ar is not None:
                            din = st.backward_eager_on_slot(a.mb, dout, last=True)
                        else:
                            din = st.backward_graphed(a.mb, dout)
                    else:
                        din = st.backward_one_chunk(a.mb, dout, last=is_last_b)
                    timeline.append((a, t0, tick()))
                    if a.stage != 0:
                        if self._rank_of_stage(a.stage - 1) == self.pp_rank:
                            local_in[(a.stage - 1, "B", a.mb)] = act_view(din, sizes[a.mb])
                        else:
                            bwd_out[(a.stage, a.mb)] = din
                elif typ == T.I:
                    st = self.local[a.stage]
                    dout = None
                    if a.stage != last_stage:
                        dout = self._take_input(recv_state, local_in, a.stage, "B", a.mb, compute)
                    t0 = tick()
                    if graphs_on:
                        prev = slot_sends.pop((a.stage, "B", a.mb % st.nslots), None)
                        if prev is not None:
                            prev.wait()
                        din = st.backward_input_graphed(a.mb, dout)
                    else:
                        din = st.backward_input_one_chunk(a.mb, dout)
                    timeline.append((a, t0, tick()))
                    if a.stage != 0:
                        if self._rank_of_stage(a.stage - 1) == self.pp_rank:
                            local_in[(a.stage - 1, "B", a.mb)] = act_view(din, sizes[a.mb])
                        else:
                            bwd_out[(a.stage, a.mb)] = din
                elif typ == T.W:
                    t0 = tick()
                    b_count[a.stage] = b_count.get(a.stage, 0) + 1
                    if graphs_on:
                        self.local[a.stage].backward_weight_graphed(a.mb, last=b_count[a.stage] == self.m)
                    else:
                        self.local[a.stage].backward_weight_one_chunk(a.mb, last=b_count[a.stage] == self.m)
                    timeline.append((a, t0, tick()))
                elif typ == T.REDUCE_GRAD:
                    self.local[a.stage].reduce_tp_grads()
                    # micro-batch averaging: folded into the optimizer kernel; done explicitly when there is none
                    if self.optimizer is None and self.scale_grads and self.local[a.stage].n_total:
                        self.local[a.stage].grad.mul_(1.0 / self.m)
                elif typ == T.ALLREDUCE:
                    allreduce_works[a.stage] = self.local[a.stage].all_reduce_grads(async_op=True)
                elif typ == T.OPTIM:
                    w = allreduce_works.pop(a.stage, None)
                    if w is not None:
                        w.wait()
                    t0 = tick()
                    opt = dict(self.optimizer)
                    gs = (1.0 / self.m if self.scale_grads else 1.0) / self.dp_size
                    max_norm = opt.pop("max_grad_norm", None)
                    if max_norm:
                        if clip is None:   # once per step: global L2 norm over every stage of every pipeline rank
                            for w2 in allreduce_works.values():
                                if w2 is not None:
                                    w2.wait()
                            allreduce_works.clear()
                            clip = self._global_clip_coef(gs, float(max_norm))
                        opt["clip_coef"] = clip
                    self.local[a.stage].optimizer_step(grad_scale=gs, **opt)
                    timeline.append((a, t0, tick()))
                elif typ in (T.UNSHARD, T.RESHARD):
                    pass
                else:
                    raise NotImplementedError(f"action {a} is not executable by this runner")
            if nvtx and idx >= 0:
                torch.cuda.nvtx.range_pop()
        except Exception as e:  # same diagnostics contract as torch's runtime (schedules.py:2257-2269)
            grid = sched_ir.grid({self.pp_rank: self.order}, error_at=(self.pp_rank, idx))
            raise RuntimeError(f"[rank {self.pp_rank}] schedule '{self.name}' failed at action {idx} "
                               f"({self.order[idx] if idx >= 0 else '?'}): {e}\n{grid}") from e

        for w in send_works:
            w.wait()
        for w in allreduce_works.values():
            if w is not None:
                w.wait()
        t_end = tick()
        self._host_ms = (time.perf_counter() - host_t0) * 1e3
        self._host_by_type = {k: v * 1e3 for k, v in host_by.items()}
        self.timeline = timeline
        self._step_span = (t_begin, t_end)
        self._keep_alive = keep_alive
        self.steps_done += 1

        out = None
        if last_stage in self.local:
            st = self.local[last_stage]
            if losses is not None:
                losses.extend(st.losses[i] for i in sorted(st.losses))
            if return_outputs and st.outputs:
                out = torch.cat([st.outputs[i] for i in sorted(st.outputs)], dim=0)
        return out

    def _take_input(self, recv_state, local_in, stage: int, kind: str, mb: int, compute):
        key = (stage, kind, mb)
        if key in local_in:
            return local_in.pop(key)
        if key not in recv_state:
            raise RuntimeError(f"compute for stage {stage} micro-batch {mb} ({kind}) started before its receive was posted")
        buf, handle = recv_state.pop(key)
        if self.cuda:
            compute.wait_event(handle)
        else:
            handle.wait()
        return buf

    def _global_clip_coef(self, grad_scale: float, max_norm: float) -> torch.Tensor:
        """min(1, max_norm / ||g||) as a device scalar, ||g|| = L2 norm of the averaged gradient of the WHOLE model:
        per-stage sums of squares (device), one all-reduce over the pipeline group, coefficient computed on the device
        and consumed by the AdamW kernel -- no host synchronisation.  (The reference has no optimizer)"""
        tot = None
        for st in self.stages:
            s = st.grad_sumsq()
            tot = s.clone() if tot is None else tot.add_(s)
        if self.pp_size > 1:
            dist.all_reduce(tot, group=self.pp_group)
        coef = self.stages[0].clip_coef
        from ..ops import functional as F
        F.clip_coef(tot, grad_scale, max_norm, coef)
        self.last_grad_norm_sq = tot    # (unscaled) for logging: norm = sqrt(tot) * grad_scale
        return coef

    def eval_step(self, *args, target=None, losses=None):
        """Forward-only pass (torch's schedule.eval(), schedules.py:248-457): runs the F actions of the GPipe
        order with dropout off and no gradient work."""
        saved = (self.order, [st.training for st in self.stages])
        try:
            for st in self.stages:
                st.training = False
            self.order = [a for a in self.order if a.type in (T.F, T.SEND_F, T.RECV_F)]
            return self.step(*args, target=target, losses=losses)
        finally:
            self.order = saved[0]
            for st, tr in zip(self.stages, saved[1]):
                st.training = tr
                st.stash.clear()

    # ---- measurement ----------------------------------------------------------------------------------------------
    def measured(self) -> Dict[str, float]:
        """Resolve the last step's timeline: step span, busy time, measured bubble fraction (this rank)."""
        if not self.timeline or self._step_span is None or self._step_span[0] is None:
            return {}
        if self.cuda:
            torch.cuda.synchronize(self.device)
            t0 = self._step_span[0]
            span = t0.elapsed_time(self._step_span[1])
            busy = sum(s.elapsed_time(e) for a, s, e in self.timeline if a.type.is_compute or a.type == T.OPTIM)
        else:
            span = (self._step_span[1] - self._step_span[0]) * 1e3
            busy = sum((e - s) * 1e3 for a, s, e in self.timeline if a.type.is_compute or a.type == T.OPTIM)
        return {"step_ms": span, "busy_ms": busy, "bubble_fraction": max(0.0, 1.0 - busy / span) if span > 0 else 0.0,
                "bubble_theory": self.theory_bubble}

    def summary(self) -> Dict[str, float]:
        """Per-action-type mean durations of the last step (ms) + host time spent issuing the step."""
        if not self.timeline or self._step_span is None or self._step_span[0] is None:
            return {}
        if self.cuda:
            torch.cuda.synchronize(self.device)
        acc: Dict[str, List[float]] = {}
        for a, s, e in self.timeline:
            d = s.elapsed_time(e) if self.cuda else (e - s) * 1e3
            acc.setdefault(a.type.value, []).append(d)
        out = {f"{k}_ms_mean": sum(v) / len(v) for k, v in acc.items()}
        out.update({f"{k}_ms_max": max(v) for k, v in acc.items()})
        out["host_issue_ms"] = getattr(self, "_host_ms", float("nan"))
        # host milliseconds by action type: a rank whose GPU waits for its neighbours runs ahead until the launch
        # queue (or an NCCL enqueue) back-pressures it, so on non-critical ranks host time ~ step time by design
        for k, v in sorted(getattr(self, "_host_by_type", {}).items(), key=lambda kv: -kv[1])[:4]:
            out[f"host_ms_in_{k}"] = v
        out.update(self.measured())
        return out

    def trace_events(self) -> List[dict]:
        """Chrome-trace events (ph='X') of the last step's MEASURED timeline, names = IR strings."""
        if not self.timeline or self._step_span is None or self._step_span[0] is None:
            return []
        evs = []
        if self.cuda:
            torch.cuda.synchronize(self.device)
            t0 = self._step_span[0]
            for a, s, e in self.timeline:
                evs.append({"name": str(a), "ph": "X", "pid": self.pp_rank, "tid": a.stage,
                            "ts": t0.elapsed_time(s) * 1e3, "dur": s.elapsed_time(e) * 1e3, "cat": a.type.value})
        else:
            t0 = self._step_span[0]
            for a, s, e in self.timeline:
                evs.append({"name": str(a), "ph": "X", "pid": self.pp_rank, "tid": a.stage, "ts": (s - t0) * 1e6,
                            "dur": (e - s) * 1e6, "cat": a.type.value})
        return evs


class ScheduleGPipe(PipelineSchedule):
    """All forwards then all backwards (torch ScheduleGPipe, used at helper:216)."""
    schedule_name = "GPipe"


class Schedule1F1B(PipelineSchedule):
    """One-forward-one-backward steady state (torch Schedule1F1B, used at helper:218)."""
    schedule_name = "1F1B"


class ScheduleZeroBubble1F1B(PipelineSchedule):
    """1F1B with backward split into input-gradient (I) and weight-gradient (W) actions, the W's moved into idle
    slots (ZB-H1 style).  An extension beyond the reference's three schedules; exercises the I/W half of the IR."""
    schedule_name = "ZeroBubble1F1B"


class ScheduleZeroBubbleInterleaved(PipelineSchedule):
    """Interleaved (virtual-stage) visiting order with split backward: bubble ~ (p-1)(F + I - W) / v."""
    schedule_name = "ZeroBubbleInterleaved"


class ScheduleInterleaved1F1B(PipelineSchedule):
    """Virtual stages, depth-first interleaved 1F1B (torch ScheduleInterleaved1F1B, used at helper:220)."""
    schedule_name = "Interleaved1F1B"


def make_schedule(name: str, stages, n_microbatches: int, loss_fn=None, **kw) -> PipelineSchedule:
    cls = {"GPipe": ScheduleGPipe, "1F1B": Schedule1F1B, "Interleaved1F1B": ScheduleInterleaved1F1B,
           "ZeroBubble1F1B": ScheduleZeroBubble1F1B, "ZeroBubbleInterleaved": ScheduleZeroBubbleInterleaved}.get(name)
    if cls is None:
        raise ValueError(f"unknown schedule {name!r}")
    if name not in ("Interleaved1F1B", "ZeroBubbleInterleaved") and not isinstance(stages, PipelineStage):
        stages = list(stages)
        if len(stages) != 1:
            raise ValueError(f"{name} runs one stage per rank")
        stages = stages[0]
    return cls(stages, n_microbatches, loss_fn, **kw)
