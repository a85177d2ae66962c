"""Stage runtime: one pipeline stage = a pruned ``Transformer`` + flat parameter/gradient/optimizer arenas +
per-micro-batch activation stashes, with explicit ``forward_one_chunk`` / ``backward_one_chunk`` entry points.

Behavioural counterpart of torch's ``PipelineStage`` + ``_PipelineStageBase`` (stage.py:116-1604; SURVEY.md
section 2.2 T11-T13), which the reference instantiates at /root/reference/LLMsDistributedTrainingHelper.py:93.
Deliberate differences (SURVEY.md section 3.3 "things the new runner must do differently"):
  * shapes are static, so there is no pickle-based run-time shape inference and no neighbour handshake;
  * backward is the explicit fused program of ``models.transformer`` (no autograd graph walk);
  * parameters live in ONE bf16 arena with fp32 master/moment/gradient arenas (memory laid out for the fused
    AdamW kernel and for a single gradient all-reduce), ``state_dict()`` keys/shapes are untouched;
  * micro-batch gradient averaging (p.grad.div_(m), torch stage.py:589-591) is folded into the optimizer kernel.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from ..models import Ctx, GradSink, Transformer, stage_backward, stage_forward, stage_loss
from ..ops import _lib, functional as F


_SEED_OFFSETS: Dict[int, torch.Tensor] = {}


class PipelineStage:
    """A stage of the pipeline.  ``submod`` is a (possibly pruned) :class:`Transformer`."""

    def __init__(self, submodule: Transformer, stage_index: int, num_stages: int, device: torch.device,
                 group=None, dp_group=None, tp_group=None):
        if not (0 <= stage_index < num_stages):
            raise ValueError(f"stage_index {stage_index} out of range for {num_stages} stages")
        self.submod = submodule
        self.stage_index = stage_index
        self.num_stages = num_stages
        self.device = torch.device(device)
        self.group = group
        self.dp_group = dp_group
        # intra-stage tensor parallelism (models/tp.py): ``submodule`` must already hold this rank's shards
        self.tp_group = tp_group
        self.tp_size = int(getattr(submodule, "tp_size", 1))
        self.tp = None            # ops.collective.TPComm, created on the first forward (needs the micro-batch geometry)
        self.seq_len = None       # set by the runner: activations between tensor-parallel stages are [b*S/tp, d] rows
        if self.tp_size > 1 and tp_group is None:
            raise ValueError("a tensor-parallel sharded model needs its tp_group")
        self.is_first = stage_index == 0
        self.is_last = stage_index == num_stages - 1
        self.args = submodule.args
        self._materialized = False
        self.stash: Dict[int, dict] = {}
        self.ctxs: Dict[int, Ctx] = {}
        self.outputs: Dict[int, torch.Tensor] = {}
        self.losses: Dict[int, torch.Tensor] = {}
        self.base_seed = 0
        self.step_count = 0
        self.training = True
        self.fused_ar = None
        self.deferred_wgrads: Dict[int, list] = {}   # micro-batch -> weight-gradient work recorded by an I action
        self.nslots = 0
        self.graphs = None
        self.recompute = False    # activation recomputation: forward keeps its input only, backward re-runs the forward
        self.has_params = any(True for _ in submodule.parameters())
        # a stage boundary behind a self-attention block carries the packed [x1 ; x0] tensor (2x the rows)
        from ..models.transformer import part_io_rows
        parts = [getattr(l, "part", "full") for l in submodule.layers.values()]
        self.in_rows, self.out_rows = part_io_rows(parts[0], parts[-1], self.args.cross_attn and not self.args.norm_first) if parts else (1, 1)

    # ---- memory layout --------------------------------------------------------------------------------------
    def materialize(self, dtype: Optional[torch.dtype] = None) -> "PipelineStage":
        """Move the stage to its device and flatten parameters into arenas (decay group first, then no-decay)."""
        if self._materialized:
            return self
        dev = self.device
        if dtype is None:
            dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
        self.dtype = dtype
        params = [(n, p) for n, p in self.submod.named_parameters()]
        decay = [(n, p) for n, p in params if p.dim() >= 2]
        no_decay = [(n, p) for n, p in params if p.dim() < 2]
        ordered = decay + no_decay
        # 16-byte aligned offsets (in elements; 8 covers bf16 and fp32 vector paths)
        offs, total = [], 0
        for _, p in ordered:
            offs.append(total)
            total += (p.numel() + 7) // 8 * 8
        self.n_decay = sum((p.numel() + 7) // 8 * 8 for _, p in decay)
        self.n_total = total
        self.param_arena = torch.zeros(max(total, 8), dtype=dtype, device=dev)
        mixed = dtype != torch.float32
        self.master = torch.zeros(max(total, 8), dtype=torch.float32, device=dev) if mixed else self.param_arena
        self.grad = torch.zeros(max(total, 8), dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(max(total, 8), dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(max(total, 8), dtype=torch.float32, device=dev)
        self.sink = GradSink()
        if dev.type == "cuda" and self.tp_size == 1:
            self.sink.side_stream = F.side_stream(dev)   # weight-gradient work forked off the input-gradient chain
        self.param_names = []
        self.tp_replicated: List[tuple] = []   # (offset, numel) of parameters every tp rank holds in full
        with torch.no_grad():
            for (n, p), o in zip(ordered, offs):
                view = self.param_arena[o:o + p.numel()].view(p.shape)
                src = p.detach()
                if src.device.type == "meta":
                    raise RuntimeError(f"parameter {n} is still on the meta device")
                view.copy_(src.to(device=dev, dtype=dtype))
                if mixed:
                    self.master[o:o + p.numel()].view(p.shape).copy_(src.to(device=dev, dtype=torch.float32))
                p.data = view
                self.sink.bind(p, self.grad[o:o + p.numel()].view(p.shape))
                self.param_names.append((n, o, tuple(p.shape)))
                if self.tp_size > 1 and getattr(p, "_tp_replicated", True):
                    self.tp_replicated.append((o, p.numel()))
        self.clip_sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.clip_coef = torch.ones(1, dtype=torch.float32, device=dev)
        if dev.type == "cuda":
            # step-varying part of every dropout seed lives in device memory (CUDA-graph friendly)
            key = dev.index if dev.index is not None else torch.cuda.current_device()
            if key not in _SEED_OFFSETS:  # one per device, shared by all (virtual) stages living on it
                _SEED_OFFSETS[key] = torch.zeros(1, dtype=torch.int64, device=dev)
                lib = _lib.load(required=True)
                with torch.cuda.device(dev):
                    _lib.check(lib.ppb_set_dropout_offset_ptr(_SEED_OFFSETS[key].data_ptr()), "ppb_set_dropout_offset_ptr")
            self.seed_offset = _SEED_OFFSETS[key]
        self._materialized = True
        return self

    def enable_fused_dp(self, n_microbatches: int) -> bool:
        """Switch the data-parallel gradient reduction of this stage to the fused in-kernel path (K10).  Returns
        False (and keeps the NCCL all-reduce) when the stage is not on CUDA, has no replicas, or owns no parameters.
        Collective over the dp group: every replica must call it."""
        import os
        import torch.distributed as dist
        if (self.device.type != "cuda" or self.dp_group is None or dist.get_world_size(self.dp_group) == 1
                or self.n_total == 0 or self.tp_size > 1 or os.environ.get("PPB_FUSED_DP", "1") == "0"):
            return False
        from ..ops.collective import FusedGradAllReduce
        dp = dist.get_world_size(self.dp_group)
        self.fused_ar = FusedGradAllReduce(self.grad, self.n_total, self.dp_group, 1.0 / (n_microbatches * dp))
        return True

    def grad_views(self) -> Dict[str, torch.Tensor]:
        return {n: self.grad[o:o + int(torch.Size(s).numel())].view(s) for n, o, s in self.param_names}

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Checkpoint shard of this stage: the reference key layout with GLOBAL layer indices.
        In mixed precision the fp32 master copy is what gets saved."""
        if not self._materialized or self.master is self.param_arena:
            return {k: v.detach().clone() for k, v in self.submod.state_dict().items()}
        return {n: self.master[o:o + int(torch.Size(s).numel())].view(s).detach().clone() for n, o, s in self.param_names}

    def load_state_dict(self, sd: Dict[str, torch.Tensor], strict: bool = True) -> None:
        own = {n for n, _, _ in self.param_names} if self._materialized else set(self.submod.state_dict())
        missing = own - set(sd)
        if strict and missing:
            raise KeyError(f"stage {self.stage_index}: checkpoint lacks {sorted(missing)[:5]}...")
        with torch.no_grad():
            if not self._materialized:
                self.submod.load_state_dict({k: v for k, v in sd.items() if k in own}, strict=strict)
                return
            for n, o, s in self.param_names:
                if n in sd:
                    numel = int(torch.Size(s).numel())
                    src = sd[n].to(self.device)
                    self.param_arena[o:o + numel].view(s).copy_(src.to(self.dtype))
                    if self.master is not self.param_arena:
                        self.master[o:o + numel].view(s).copy_(src.to(torch.float32))

    def optimizer_state_dict(self) -> Dict[str, Dict[str, torch.Tensor]]:
        out = {}
        for n, o, s in self.param_names:
            numel = int(torch.Size(s).numel())
            out[n] = {"exp_avg": self.exp_avg[o:o + numel].view(s).clone(), "exp_avg_sq": self.exp_avg_sq[o:o + numel].view(s).clone()}
        return {"state": out, "step": self.step_count}

    def load_optimizer_state_dict(self, osd) -> None:
        for n, o, s in self.param_names:
            if n in osd["state"]:
                numel = int(torch.Size(s).numel())
                self.exp_avg[o:o + numel].view(s).copy_(osd["state"][n]["exp_avg"].to(self.device))
                self.exp_avg_sq[o:o + numel].view(s).copy_(osd["state"][n]["exp_avg_sq"].to(self.device))
        self.step_count = int(osd.get("step", 0))

    # ---- per-step / per-micro-batch ----------------------------------------------------------------------------
    def begin_step(self) -> None:
        """Counterpart of torch's stage.clear_runtime_states() (stage.py:512-530)."""
        self.stash.clear()
        self.ctxs.clear()
        self.outputs.clear()
        self.losses.clear()
        self.deferred_wgrads.clear()
        if self.device.type == "cuda" and self._materialized:
            self.seed_offset.fill_(self.step_count * 0x632BE5AB)

    def _mb_offset(self, mb: int) -> int:
        """Step- and micro-batch-varying part of every dropout seed (one scheme for CPU, eager CUDA and graph replays)."""
        return (self.step_count * 0x632BE5AB + (mb + 1) * 0x9E3779B1) & 0x7FFFFFFFFFFFFFFF

    def _ctx(self, mb: int, batch: int, seq: int) -> Ctx:
        # on CUDA the offset lives in device memory (set right before the launches of a micro-batch), so the seeds
        # baked into captured kernels are micro-batch independent; the CPU program adds the same number on the host
        off = 0 if self.device.type == "cuda" else self._mb_offset(mb)
        ctx = Ctx(self.args, batch, seq, self.base_seed, 0, self.training, off)
        if self.tp_size > 1:
            rows = batch * seq // self.tp_size
            if batch * seq % self.tp_size or (self.tp is not None and self.tp.rows != rows):
                raise ValueError(f"tensor parallelism needs equal micro-batches with b*S divisible by tp (got {batch}x{seq})")
            if self.tp is None:
                from ..ops.collective import TPComm
                self.tp = TPComm(self.tp_group, rows, self.args.dim, self.device)
            ctx.tp = self.tp
        return ctx

    def forward_one_chunk(self, mb: int, inp: torch.Tensor, target: Optional[torch.Tensor] = None,
                          loss_scale: float = 1.0, loss_fn=None) -> torch.Tensor:
        """Run the stage forward for micro-batch ``mb``.  ``inp`` is tokens [b, S] on the first stage, activations
        [b, S, d] otherwise.  On the last stage with a ``target`` the fused loss runs too (its gradient replaces the
        logits in place).  Returns the stage output ([b*S, d] activations or logits)."""
        if not self._materialized:
            self.materialize()
        if self.tp_size > 1 and not self.is_first:   # [b*S/tp, d] rows from the previous tensor-parallel stage
            S = self.seq_len if self.seq_len is not None else (target.shape[1] if target is not None else None)
            if S is None:
                raise ValueError("a tensor-parallel stage needs seq_len (set by the runner) to interpret its input rows")
            b = inp.reshape(-1, self.args.dim).shape[0] * self.tp_size // S
        else:
            b, S = inp.shape[0] // (1 if self.is_first else self.in_rows), inp.shape[1]
        ctx = self._ctx(mb, b, S)
        st: dict = {}
        if self.device.type == "cuda":
            self._set_mb_seed(mb)
        with torch.no_grad():
            out = stage_forward(self.submod, inp, ctx, st)
            if self.is_last and target is not None and self.submod.output is not None:
                if loss_fn is None:
                    self.losses[mb] = stage_loss(out, target, loss_scale, st, ctx.tp)
                    if ctx.tp is not None:   # every rank holds its rows' share of the mean; report the mean
                        ctx.tp.all_reduce(self.losses[mb])
                elif ctx.tp is not None:
                    raise NotImplementedError("python loss_fn with tensor parallelism: use the fused loss (loss_fn=None)")
                else:  # arbitrary user loss: one autograd hop over the logits only
                    with torch.enable_grad():
                        lg = out.detach().float().view(b, S, -1).requires_grad_(True)
                        loss = loss_fn(lg, target)
                        (g,) = torch.autograd.grad(loss * loss_scale, lg)
                    out.copy_(g.view(out.shape).to(out.dtype))
                    st["dlogits"] = out
                    self.losses[mb] = loss.detach().reshape(1).float()
        if self.is_last and "dlogits" not in st:
            self.outputs[mb] = out.view(b, S, -1) if ctx.tp is None else out
        if self.recompute and self.training and loss_fn is None:
            # activation recomputation: keep only what is needed to run this forward again right before its backward
            # (the counter-based dropout streams regenerate the same masks from the same seed + offset)
            st = {"_recompute": (inp, target, loss_scale)}
        self.stash[mb] = st
        self.ctxs[mb] = ctx
        return out

    def _rebuild_stash(self, mb: int, st: dict, ctx: Ctx) -> dict:
        inp, target, loss_scale = st["_recompute"]
        st = {}
        with torch.no_grad():
            out = stage_forward(self.submod, inp, ctx, st)
            if self.is_last and target is not None and self.submod.output is not None:
                stage_loss(out, target, loss_scale, st, ctx.tp)
        return st

    def backward_one_chunk(self, mb: int, dout: Optional[torch.Tensor] = None, last: bool = False) -> Optional[torch.Tensor]:
        """Backward for micro-batch ``mb``; ``dout`` is the gradient received from the next stage (None on the last
        stage: the fused loss already left dlogits in the stash).  Returns the input gradient [b*S, d] or None."""
        st = self.stash.pop(mb)
        ctx = self.ctxs.pop(mb)
        if self.device.type == "cuda":
            self._set_mb_seed(mb)
        if "_recompute" in st:
            st = self._rebuild_stash(mb, st, ctx)
        if dout is None:
            if "dlogits" not in st:
                raise RuntimeError(f"stage {self.stage_index}: backward of micro-batch {mb} without a gradient or a loss")
            douts = [st["dlogits"]]
        else:
            douts = [dout]
        fused = self.fused_ar if (last and self.fused_ar is not None) else None
        if fused is not None:
            fused.begin_last_backward()   # this backward's weight-gradient GEMMs carry the DP all-reduce
            self.sink.fused_ar = fused
        try:
            with torch.no_grad():
                return stage_backward(self.submod, douts, ctx, st, self.sink, need_input_grad=not self.is_first)
        finally:
            self.sink.fused_ar = None

    def backward_input_one_chunk(self, mb: int, dout: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """IR action ``I``: the input-gradient chain only (dgrad GEMMs, attention backward, norm backward -- whose
        gamma/beta gradients are by-products of the same kernel and land now).  Every weight-gradient GEMM and the
        embedding scatter are recorded, with their operands kept alive, for the matching ``W`` action.  Counterpart
        of torch's stage_backward_input (_backward.py:165, driven from stage.py:873); explicit here because the
        stage program is explicit."""
        if mb in self.deferred_wgrads:
            raise RuntimeError(f"stage {self.stage_index}: I{mb} issued twice without W{mb}")
        self.sink.deferred = []
        try:
            din = self.backward_one_chunk(mb, dout, last=False)
        finally:
            items, self.sink.deferred = self.sink.deferred, None
        self.deferred_wgrads[mb] = items
        return din

    def backward_weight_one_chunk(self, mb: int, last: bool = False) -> None:
        """IR action ``W``: run the weight-gradient work recorded by ``I`` of the same micro-batch (torch
        stage_backward_weight, _backward.py:283).  ``last`` = this is the stage's final W of the step, whose GEMMs
        carry the fused data-parallel all-reduce when that is enabled."""
        items = self.deferred_wgrads.pop(mb, None)
        if items is None:
            raise RuntimeError(f"stage {self.stage_index}: W{mb} without a preceding I{mb}")
        fused = self.fused_ar if (last and self.fused_ar is not None) else None
        if fused is not None:
            fused.begin_last_backward()
            self.sink.fused_ar = fused
        try:
            with torch.no_grad():
                self.sink.run_deferred(items)
        finally:
            self.sink.fused_ar = None

    # ---- CUDA graphs ----------------------------------------------------------------------------------------------
    def enable_graphs(self, nslots: int) -> None:
        """Capture forward / backward of one micro-batch per in-flight slot into CUDA graphs (launch-bound inner
        loops: ~45 / ~100 launches per 3-layer stage).  Micro-batch ``mb`` uses slot ``mb % nslots``; the
        activation stash of a slot lives in the graphs' private memory pool."""
        self.nslots = int(nslots)
        self.graphs: Dict[int, dict] = {}

    def _set_mb_seed(self, mb: int) -> None:
        # micro-batch-varying part of the dropout seed lives in device memory, so graph replays get fresh masks
        self.seed_offset.fill_(self._mb_offset(mb))

    def forward_graphed(self, mb: int, inp: torch.Tensor, target: Optional[torch.Tensor] = None) -> torch.Tensor:
        from ..ops import _lib as L
        slot = mb % self.nslots
        g = self.graphs.get(slot)
        self._set_mb_seed(mb)
        if g is None:
            g = {"in": inp.clone(), "tgt": None if target is None else target.clone(), "F": torch.cuda.CUDAGraph(),
                 "B": None, "dout": None, "din": None, "din_B": None, "din_I": None, "sends": {},
                 # one private pool per slot: within a slot the replay order F -> B (or F -> I -> W) is the capture
                 # order, so its graphs may share scratch; across slots live ranges interleave differently from the
                 # capture order (slot 1's stash is live while slot 0's backward replays), so nothing is shared
                 "pool": torch.cuda.graph_pool_handle()}
            b, S = inp.shape[0] // (1 if self.is_first else self.in_rows), inp.shape[1]
            ctx = Ctx(self.args, b, S, self.base_seed, 0, self.training, 0)   # micro-batch variation comes from the device-side offset only
            st: dict = {}
            n0 = L.launches()
            with torch.cuda.graph(g["F"], pool=g["pool"]):
                with torch.no_grad():
                    out = stage_forward(self.submod, g["in"], ctx, st)
                    loss = None
                    if self.is_last and g["tgt"] is not None and self.submod.output is not None:
                        loss = stage_loss(out, g["tgt"], 1.0, st)
            g.update(out=out, loss=loss, stash=st, ctx=ctx, nF=L.launches() - n0, shape=(b, S))
            self.graphs[slot] = g
        else:
            if tuple(inp.shape) != tuple(g["in"].shape):   # never rely on copy_ broadcasting (uneven micro-batches)
                raise RuntimeError(f"stage {self.stage_index}: graph slot {slot} was captured for input "
                                   f"{tuple(g['in'].shape)}, got {tuple(inp.shape)}")
            if inp.data_ptr() != g["in"].data_ptr():       # a receive that landed in the static input needs no copy
                g["in"].copy_(inp)
            if target is not None:
                g["tgt"].copy_(target)
            L.count_launch(g["nF"])
        g["F"].replay()
        if g["loss"] is not None:
            self.losses[mb] = g["loss"].clone()
        self.stash[mb] = g["stash"]
        self.ctxs[mb] = g["ctx"]
        return g["out"]

    @staticmethod
    def _slot_dout(g: dict, dout: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        """The slot's static output-gradient buffer.  Allocated ONCE per slot and never rebound: a slot can see both a
        B graph and an I graph (ZeroBubble orders merge only some micro-batches back into B), and every captured graph
        of the slot must keep reading the buffer the replays copy into."""
        if dout is None:
            return None
        if g.get("dout") is None:
            g["dout"] = dout.clone()
        elif dout.data_ptr() != g["dout"].data_ptr():
            if dout.numel() != g["dout"].numel():
                raise RuntimeError(f"graph slot captured for an output gradient of {tuple(g['dout'].shape)}, got {tuple(dout.shape)}")
            g["dout"].copy_(dout.reshape(g["dout"].shape))
        return g["dout"]

    def _slot_din(self, g: dict) -> Optional[torch.Tensor]:
        """The slot's static input-gradient buffer, allocated OUTSIDE the graph pool.  A tensor allocated while the
        backward graph is captured may reuse pool memory that the slot's forward graph uses for temporaries; the
        consumer of the gradient (an NCCL send still in flight, or a local virtual stage that runs later) would then
        race with the slot's next forward replay.  The captured backward ends with one copy into this buffer."""
        if self.is_first:
            return None
        if g.get("din") is None:
            b, S = g["shape"]
            g["din"] = torch.empty((self.in_rows * b * S, self.args.dim), dtype=self.dtype, device=self.device)
        return g["din"]

    def backward_graphed(self, mb: int, dout: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        from ..ops import _lib as L
        slot = mb % self.nslots
        g = self.graphs[slot]
        self._set_mb_seed(mb)
        self.stash.pop(mb, None)
        self.ctxs.pop(mb, None)
        sd = self._slot_dout(g, dout)
        if g["B"] is None:
            g["B"] = torch.cuda.CUDAGraph()
            st, ctx = g["stash"], g["ctx"]
            din_static = self._slot_din(g)
            n0 = L.launches()
            with torch.cuda.graph(g["B"], pool=g["pool"]):
                with torch.no_grad():
                    douts = [st["dlogits"]] if sd is None else [sd]
                    din = stage_backward(self.submod, douts, ctx, st, self.sink, need_input_grad=not self.is_first)
                    g["din_B"] = None if din is None else din_static.copy_(din.reshape(din_static.shape))
            g["nB"] = L.launches() - n0
        else:
            L.count_launch(g["nB"])
        g["B"].replay()
        return g["din_B"]

    def backward_input_graphed(self, mb: int, dout: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
        """Graph replay of IR action ``I`` on the slot of ``mb``: the deferred weight-gradient operands are static
        tensors of the graph pool, so the matching ``W`` is a second graph over the same addresses."""
        from ..ops import _lib as L
        slot = mb % self.nslots
        g = self.graphs[slot]
        self._set_mb_seed(mb)
        self.stash.pop(mb, None)
        self.ctxs.pop(mb, None)
        sd = self._slot_dout(g, dout)
        if g.get("I") is None:
            g["I"] = torch.cuda.CUDAGraph()
            st, ctx = g["stash"], g["ctx"]
            din_static = self._slot_din(g)
            n0 = L.launches()
            self.sink.deferred = []
            try:
                with torch.cuda.graph(g["I"], pool=g["pool"]):
                    with torch.no_grad():
                        douts = [st["dlogits"]] if sd is None else [sd]
                        din = stage_backward(self.submod, douts, ctx, st, self.sink, need_input_grad=not self.is_first)
                        g["din_I"] = None if din is None else din_static.copy_(din.reshape(din_static.shape))
            finally:
                g["wlist"], self.sink.deferred = self.sink.deferred, None
            g["nI"] = L.launches() - n0
            g["W"] = None
        else:
            L.count_launch(g["nI"])
        g["I"].replay()
        return g["din_I"]

    def backward_weight_graphed(self, mb: int, last: bool = False) -> None:
        from ..ops import _lib as L
        g = self.graphs[mb % self.nslots]
        if last and self.fused_ar is not None:   # fused DP all-reduce epilogues: a different launch sequence, eager
            self.fused_ar.begin_last_backward()
            self.sink.fused_ar = self.fused_ar
            try:
                with torch.no_grad():
                    self.sink.run_deferred(g["wlist"])
            finally:
                self.sink.fused_ar = None
            return
        if g.get("W") is None:
            g["W"] = torch.cuda.CUDAGraph()
            n0 = L.launches()
            with torch.cuda.graph(g["W"], pool=g["pool"]):
                with torch.no_grad():
                    self.sink.run_deferred(g["wlist"])
            g["nW"] = L.launches() - n0
        else:
            L.count_launch(g["nW"])
        g["W"].replay()

    def backward_eager_on_slot(self, mb: int, dout: Optional[torch.Tensor], last: bool) -> Optional[torch.Tensor]:
        """Eager backward over a graph slot's stash (used for the step's last micro-batch when the fused DP
        all-reduce rides on its weight-gradient GEMMs, which is a different launch sequence)."""
        slot = mb % self.nslots
        g = self.graphs[slot]
        self._set_mb_seed(mb)
        self.stash[mb] = g["stash"]
        self.ctxs[mb] = g["ctx"]
        return self.backward_one_chunk(mb, dout, last=last)

    # ---- gradient reduction + optimizer ---------------------------------------------------------------------------
    def reduce_tp_grads(self) -> None:
        """Sum, over the tensor-parallel group, the gradients of the parameters every tp rank holds in full (norms,
        row-parallel biases, embedding, head): each rank computed them on its own token rows only.  One packed
        all-reduce per step."""
        if self.tp_size == 1 or not self.tp_replicated:
            return
        import torch.distributed as dist
        flat = torch.cat([self.grad[o:o + n] for o, n in self.tp_replicated])
        dist.all_reduce(flat, group=self.tp_group)
        i = 0
        for o, n in self.tp_replicated:
            self.grad[o:o + n].copy_(flat[i:i + n])
            i += n


    def all_reduce_grads(self, async_op: bool = False):
        """Data-parallel all-reduce of the whole gradient arena in one collective (baseline path: NCCL/gloo;
        the fused in-kernel path lives in ops.collective)."""
        import torch.distributed as dist
        if self.dp_group is None or dist.get_world_size(self.dp_group) == 1 or self.n_total == 0:
            return None
        if self.fused_ar is not None:
            self.fused_ar.finish()   # pushes the non-GEMM gradients + device-side barrier; nothing left to wait for
            return None
        return dist.all_reduce(self.grad, group=self.dp_group, async_op=async_op)

    def optimizer_step(self, lr: float = 1e-4, betas=(0.9, 0.95), eps: float = 1e-8, weight_decay: float = 0.1,
                       grad_scale: float = 1.0, max_grad_norm: Optional[float] = None, clip_coef=None) -> None:
        """Fused AdamW over the arenas: two launches (decay group / no-decay group)."""
        if not self._materialized or self.n_total == 0:
            return
        self.step_count += 1
        mixed = self.master is not self.param_arena
        nd = self.n_decay
        coef = clip_coef
        grad = self.grad
        if self.fused_ar is not None:   # already summed over the replicas by the fused GEMM epilogues
            grad = self.fused_ar.final
        for lo, hi, wd in ((0, nd, weight_decay), (nd, self.n_total, 0.0)):
            if hi <= lo:
                continue
            F.adamw_step(self.param_arena[lo:hi] if mixed else None, self.master[lo:hi], self.exp_avg[lo:hi],
                         self.exp_avg_sq[lo:hi], grad[lo:hi], lr, betas[0], betas[1], eps, wd, self.step_count,
                         grad_scale, coef, zero_grad=True)
        if self.fused_ar is not None:
            self.grad.zero_()
            self.fused_ar.after_optimizer()

    def zero_grad(self) -> None:
        if self._materialized:
            self.grad.zero_()

    def grad_sumsq(self) -> torch.Tensor:
        """Sum of squares of this stage's (already reduced) gradients, each parameter counted once across the
        tensor-parallel group: slices add up over the tp ranks, replicated parameters are counted by one rank."""
        grad = self.fused_ar.final if self.fused_ar is not None else self.grad
        self.clip_sumsq.zero_()
        if self.n_total:
            F.sumsq_accum(grad[:self.n_total], self.clip_sumsq)
        if self.tp_size > 1:
            import torch.distributed as dist
            rep = torch.zeros_like(self.clip_sumsq)
            if self.tp_replicated:
                F.sumsq_accum(torch.cat([grad[o:o + n] for o, n in self.tp_replicated]), rep)
            self.clip_sumsq.sub_(rep)
            dist.all_reduce(self.clip_sumsq, group=self.tp_group)
            self.clip_sumsq.add_(rep)
        return self.clip_sumsq
