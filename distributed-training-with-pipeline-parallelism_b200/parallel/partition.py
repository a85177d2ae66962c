"""Stage partitioner: which layers (and which of embedding / final norm / LM head) a pipeline stage owns.

``reference_ranges`` reproduces the helper's rule exactly (/root/reference/LLMsDistributedTrainingHelper.py:70-75):
``layers_per_stage = n_layers // num_stages``, stage i keeps ``[i*lps, (i+1)*lps)`` and the LAST stage takes the
remainder.  ``balanced_ranges`` is the cost-aware alternative: the LM head costs
about as much as a decoder layer and the embedding is nearly free, so the last stage gets fewer layers.

``prune_to_stage`` turns a full ``Transformer`` into the stage's sub-model in place by deleting ``ModuleDict``
entries (global layer indices are preserved, so the union of the stages' ``state_dict()`` is the full checkpoint
with no key translation -- SURVEY.md section 2.3).  ``build_stage_model`` constructs a stage directly without
ever materialising the full model (reference quirk 5: every rank builds the whole model first, helper:207).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..models import ModelArgs, Transformer
from ..models.transformer import DecoderLayer


def reference_ranges(n_layers: int, num_stages: int) -> List[Tuple[int, int]]:
    lps = n_layers // num_stages
    out = []
    for s in range(num_stages):
        start = s * lps
        end = start + lps if s < num_stages - 1 else n_layers
        out.append((start, end))
    return out


# ----------------------------------------------------------------------------------------------------------------
# cost model: measured per-unit times when available, an analytic estimate otherwise
# ----------------------------------------------------------------------------------------------------------------
# Units of a layer (the places a stage boundary may fall, rule 'fine'):
#   post-LN reference block with cross-attention: 'self' | 'cross' | 'ffn'   (a cut behind 'self' ships TWO tensors:
#       x1 and the layer input x0, which is the cross-attention memory -- packed as [2T, d] rows)
#   every other block (pre-LN GPT / Llama, post-LN without cross-attention): 'attn' | 'ffn'
_MEASURED: dict = {}


def layer_units(args: ModelArgs) -> Tuple[str, ...]:
    if not args.norm_first and args.cross_attn:
        return ("self", "cross", "ffn")
    return ("attn", "ffn")


def _cost_key(args: ModelArgs, seq_len: int, tokens: int):
    return (args.dim, args.n_heads, args.ffn_dim, args.vocab_size, args.cross_attn, args.norm_first, args.activation,
            args.causal, round(args.dropout, 4), str(args.dtype), seq_len, tokens)


def set_measured_costs(args: ModelArgs, seq_len: int, costs: dict, tokens: int = 0) -> None:
    """Register measured forward+backward times (any unit, e.g. ms) of one micro-batch for the units of ``args``:
    keys = the names in ``layer_units(args)`` plus 'head' (final norm + LM head + loss) and 'embed'.  From then on every
    partition rule balances these numbers instead of the analytic estimate (``measure_unit_costs`` produces them)."""
    need = set(layer_units(args)) | {"head", "embed"}
    if not need <= set(costs):
        raise ValueError(f"measured costs need the keys {sorted(need)}, got {sorted(costs)}")
    _MEASURED[_cost_key(args, seq_len, 0)] = dict(costs)
    if tokens:
        _MEASURED[_cost_key(args, seq_len, tokens)] = dict(costs)


def clear_measured_costs() -> None:
    _MEASURED.clear()


def unit_costs(args: Optional[ModelArgs], seq_len: int = 1024) -> dict:
    """{'<unit>': cost, ..., 'head': cost, 'embed': cost, 'source': 'measured' | 'model'} per micro-batch."""
    if args is None:
        return {"attn": 0.5, "ffn": 0.5, "self": 0.4, "cross": 0.4, "head": 1.0, "embed": 0.0, "source": "uniform"}
    m = _MEASURED.get(_cost_key(args, seq_len, 0))
    if m is not None:
        return {**m, "source": "measured"}
    # Analytic fallback (no GPU measurement available, e.g. planning on a CPU box): FLOPs per token weighted by the
    # efficiency each part reached on B200 in round 1 (layer GEMMs ~0.40 of the tensor-core peak, attention ~0.20, the
    # vocabulary-sized head GEMMs ~0.9, the loss pass HBM-bound).  Only RATIOS matter.
    d, f, V = args.dim, args.ffn_dim, args.vocab_size
    attn_core = 4 * seq_len * d * (0.5 if args.causal else 1.0) / 0.20
    proj = 8 * d * d / 0.40
    ffn = (6 if args.activation == "swiglu" else 4) * d * f / 0.40
    out = {"ffn": ffn, "head": 2 * d * V / 0.9 + 650.0 * V, "embed": 0.0, "source": "model"}
    if "self" in layer_units(args):
        out.update(self=proj + attn_core, cross=proj + attn_core)
    else:
        out.update(attn=proj + attn_core)
    return out


def _layer_cost(args: Optional[ModelArgs], seq_len: int) -> Tuple[float, float, float]:
    c = unit_costs(args, seq_len)
    units = layer_units(args) if args is not None else ("attn", "ffn")
    return sum(c[u] for u in units), c["head"], c["embed"]


def balanced_ranges(n_layers: int, num_stages: int, args: Optional[ModelArgs] = None, seq_len: int = 1024) -> List[Tuple[int, int]]:
    """Contiguous split at WHOLE-layer granularity minimising the most expensive stage, with the LM head (and loss)
    charged to the last stage and the embedding to the first."""
    layer_cost, head_cost, embed_cost = _layer_cost(args, seq_len)
    if args is None:
        layer_cost, head_cost, embed_cost = 1.0, 1.0, 0.0
    if n_layers < num_stages:   # more stages than layers: every stage gets at most one layer (reference quirk 8)
        return [(min(s, n_layers), min(s + 1, n_layers)) for s in range(num_stages)]
    units = _min_max_partition([layer_cost] * n_layers, num_stages, head_cost, embed_cost)
    return units


def _min_max_partition(costs: List[float], num_stages: int, tail_cost: float, head_cost: float = 0.0) -> List[Tuple[int, int]]:
    """Contiguous split of ``costs`` into ``num_stages`` non-empty groups minimising (largest group, sum of squares);
    ``tail_cost`` is added to the last group, ``head_cost`` to the first."""
    n, S = len(costs), num_stages
    pre = [0.0]
    for c in costs:
        pre.append(pre[-1] + c)
    INF = (float("inf"), float("inf"))
    best = [[INF] * (n + 1) for _ in range(S + 1)]
    cut = [[n] * (n + 1) for _ in range(S + 1)]
    for i in range(n):
        c = pre[n] - pre[i] + tail_cost + (head_cost if S == 1 else 0.0)
        best[S - 1][i] = (c, c * c)
    for s in range(S - 2, -1, -1):
        for i in range(n):
            if s > 0 and i == 0:
                continue
            for j in range(i + 1, n):
                nxt = best[s + 1][j]
                if nxt[0] == float("inf"):
                    continue
                c = pre[j] - pre[i] + (head_cost if s == 0 else 0.0)
                w = (max(c, nxt[0]), c * c + nxt[1])
                if w < best[s][i]:
                    best[s][i], cut[s][i] = w, j
    out, i = [], 0
    for s in range(S):
        j = cut[s][i] if s < S - 1 else n
        out.append((i, j))
        i = j
    return out


def fine_unit_ranges(n_layers: int, num_stages: int, args: ModelArgs, seq_len: int = 1024) -> List[Tuple[int, int]]:
    """Stage boundaries at SUB-layer granularity.  With U = len(layer_units(args)) units per layer, unit U*i + k is
    block k of layer i; returns [u0, u1) unit ranges.  The reference block has three units (self-attention |
    cross-attention | feed-forward, roughly 0.4 / 0.4 / 0.2 of a layer at the flagship shape), so the most expensive
    of 8 stages of the 24-layer flagship drops from 4.0 layer-equivalents (whole layers) to about 3.4."""
    units = layer_units(args)
    c = unit_costs(args, seq_len)
    costs = [c[units[u % len(units)]] for u in range(len(units) * n_layers)]
    if len(costs) < num_stages:
        raise ValueError(f"{n_layers} layers ({len(costs)} units) cannot be cut into {num_stages} stages")
    return _min_max_partition(costs, num_stages, c["head"], c["embed"])


def _stage_costs(n_layers: int, num_stages: int, rule: str, args: ModelArgs, seq_len: int) -> List[float]:
    c = unit_costs(args, seq_len)
    units = layer_units(args)
    out = []
    for s in range(num_stages):
        w = sum(sum(c[u] for u in part_units(part, units)) for _, part in stage_layer_parts(n_layers, num_stages, s, rule, args, seq_len))
        out.append(w + (c["head"] if s == num_stages - 1 else 0.0) + (c["embed"] if s == 0 else 0.0))
    return out


def resolve_rule(rule: str, n_layers: int, num_stages: int, args: Optional[ModelArgs], seq_len: int = 1024,
                 tp: int = 1) -> str:
    """'auto' -> 'fine' when sub-layer cuts lower the most expensive stage by more than 2 % (and the configuration
    supports them), else 'balanced'.  Other rule names pass through."""
    if rule != "auto":
        return rule
    if args is None or num_stages == 1 or tp > 1 or len(layer_units(args)) * n_layers < num_stages:
        return "balanced"
    try:
        fine = max(_stage_costs(n_layers, num_stages, "fine", args, seq_len))
        bal = max(_stage_costs(n_layers, num_stages, "balanced", args, seq_len))
    except ValueError:
        return "balanced"
    return "fine" if fine < 0.98 * bal else "balanced"


def stage_ranges(n_layers: int, num_stages: int, rule: str = "reference", args: Optional[ModelArgs] = None,
                 seq_len: int = 1024) -> List[Tuple[int, int]]:
    """[start, end) LAYER ranges per stage.  For rule 'fine' a boundary layer appears in both neighbouring stages
    (see ``stage_layer_parts`` for which units each of them owns)."""
    rule = resolve_rule(rule, n_layers, num_stages, args, seq_len)
    if rule == "reference":
        return reference_ranges(n_layers, num_stages)
    if rule == "balanced":
        return balanced_ranges(n_layers, num_stages, args, seq_len)
    if rule == "fine":
        U = len(layer_units(args)) if args is not None else 2
        return [(u0 // U, (u1 + U - 1) // U) for u0, u1 in _fine_units(n_layers, num_stages, args, seq_len)]
    raise ValueError(f"unknown partition rule {rule!r}")


def _fine_units(n_layers: int, num_stages: int, args: Optional[ModelArgs], seq_len: int):
    if args is None:
        return [(2 * a, 2 * b) for a, b in balanced_ranges(n_layers, num_stages, args, seq_len)]
    return fine_unit_ranges(n_layers, num_stages, args, seq_len)


def part_name(k0: int, k1: int, units: Tuple[str, ...]) -> str:
    """Name of the contiguous unit range [k0, k1) of one layer: 'full', or the owned unit names joined by '+'
    ('self+cross' is spelled 'attn', as it was before the three-way split existed)."""
    if k0 == 0 and k1 == len(units):
        return "full"
    name = "+".join(units[k0:k1])
    return "attn" if name == "self+cross" else name


def part_units(part: str, units: Tuple[str, ...]) -> Tuple[str, ...]:
    if part == "full":
        return tuple(units)
    if part == "attn" and "attn" not in units:
        return ("self", "cross")
    return tuple(part.split("+"))


def stage_layer_parts(n_layers: int, num_stages: int, stage_index: int, rule: str, args: Optional[ModelArgs] = None,
                      seq_len: int = 1024) -> List[Tuple[int, str]]:
    """[(layer index, part), ...] owned by ``stage_index``; part = 'full' or the owned units ('attn', 'ffn', and for
    the reference block also 'self', 'cross', 'cross+ffn')."""
    rule = resolve_rule(rule, n_layers, num_stages, args, seq_len)
    if rule != "fine":
        start, end = stage_ranges(n_layers, num_stages, rule, args, seq_len)[stage_index]
        return [(i, "full") for i in range(start, end)]
    units = layer_units(args) if args is not None else ("attn", "ffn")
    U = len(units)
    u0, u1 = _fine_units(n_layers, num_stages, args, seq_len)[stage_index]
    out = []
    for i in range(u0 // U, (u1 + U - 1) // U):
        k0, k1 = max(u0, U * i) - U * i, min(u1, U * (i + 1)) - U * i
        out.append((i, part_name(k0, k1, units)))
    return out


_UNIT_MODULES = {"self": ("self_attn", "norm1"), "cross": ("multihead_attn", "norm2"),
                 "attn": ("self_attn", "norm1"), "ffn": ("linear1", "linear2", "norm3")}


def _apply_part(layer: DecoderLayer, part: str) -> DecoderLayer:
    """Keep only the parameters of the units in ``part`` (checkpoint keys of what remains are unchanged)."""
    layer.part = part
    if part == "full":
        return layer
    has_cross = hasattr(layer, "multihead_attn") or hasattr(layer, "norm2")
    units = ("self", "cross", "ffn") if has_cross else ("attn", "ffn")
    keep = {m for u in part_units(part, units) for m in _UNIT_MODULES[u]}
    for u in units:
        for name in _UNIT_MODULES[u]:
            if name not in keep and hasattr(layer, name):
                delattr(layer, name)
    return layer


def prune_to_stage(model: Transformer, stage_index: int, num_stages: int, rule: str = "reference",
                   seq_len: int = 1024) -> Transformer:
    """In-place pruning, same observable result as helper:60-91."""
    n_layers = len(model.layers)
    parts = dict(stage_layer_parts(n_layers, num_stages, stage_index, rule, model.args, seq_len))
    for i in [k for k in model.layers.keys() if int(k) not in parts]:
        del model.layers[i]
    for k, layer in model.layers.items():
        _apply_part(layer, parts[int(k)])
    if stage_index != 0:
        model.tok_embeddings = None
    if stage_index != num_stages - 1:
        model.norm = None
        model.output = None
    return model


def build_stage_model(args: ModelArgs, stage_index: int, num_stages: int, rule: str = "reference", seq_len: int = 1024,
                      device: Optional[torch.device] = None, seed: Optional[int] = None) -> Transformer:
    """Build ONLY the modules of one stage (meta-device skeleton, then materialise the slice)."""
    parts = stage_layer_parts(args.n_layers, num_stages, stage_index, rule, args, seq_len)
    with torch.device("meta"):
        skel = Transformer(ModelArgs(**{**args.__dict__, "n_layers": 0}))
    model = skel
    model.args = args
    gen_dev = device if device is not None else torch.device("cpu")
    if seed is not None:
        torch.manual_seed(seed + 7919 * stage_index)
    with torch.device(gen_dev):
        for i, part in parts:
            model.layers[str(i)] = _apply_part(DecoderLayer(args), part)
        full_tail = Transformer(ModelArgs(**{**args.__dict__, "n_layers": 0}))
    model.tok_embeddings = full_tail.tok_embeddings if stage_index == 0 else None
    model.norm = full_tail.norm if stage_index == num_stages - 1 else None
    model.output = full_tail.output if stage_index == num_stages - 1 else None
    return model


# ----------------------------------------------------------------------------------------------------------------
# measured unit costs (what 'balanced' / 'fine' / 'auto' balance on a GPU)
# ----------------------------------------------------------------------------------------------------------------
def measure_unit_costs(args: ModelArgs, micro_batch: int, seq_len: int, device: torch.device, iters: int = 4) -> dict:
    """Forward+backward time (ms, CUDA events) of ONE micro-batch for every unit a stage boundary can separate, for the
    final norm + LM head + loss ('head') and for the embedding ('embed'): each is built as a one-unit stage and timed
    as the runner executes it (CUDA-graph replays of the stage program), so the numbers track the kernels as they
    are today instead of constants calibrated once.  Takes about a second."""
    from .stage import PipelineStage
    dev = torch.device(device)
    d, T = args.dim, micro_batch * seq_len
    dt = args.dtype if args.dtype != torch.float32 or dev.type != "cuda" else torch.bfloat16
    margs = ModelArgs(**{**args.__dict__, "n_layers": 1, "dtype": dt})
    units = layer_units(margs)
    out: dict = {}

    def run(stage: "PipelineStage", inp, tgt, dout) -> float:
        stage.materialize()
        stage.training = True
        stage.base_seed = 1
        stage.begin_step()
        stage.forward_one_chunk(0, inp, tgt)        # eager once: module loading / attribute setting outside any capture
        stage.backward_one_chunk(0, dout)
        stage.enable_graphs(1)
        best = float("inf")
        for it in range(iters + 2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            stage.forward_graphed(0, inp, tgt)
            stage.backward_graphed(0, dout)
            e1.record()
            e1.synchronize()
            if it >= 2:
                best = min(best, e0.elapsed_time(e1))
        return best

    def one_unit_model(part: Optional[str], embed: bool, head: bool, n_layers: int = 1) -> Transformer:
        with torch.device(dev):
            m = Transformer(ModelArgs(**{**margs.__dict__, "n_layers": n_layers}))
        if part is None:
            for k in list(m.layers.keys()):
                del m.layers[k]
        else:
            for layer in m.layers.values():
                _apply_part(layer, part)
        if not embed:
            m.tok_embeddings = None
        if not head:
            m.norm = None
            m.output = None
        return m

    gen = torch.Generator(device="cpu").manual_seed(7)
    tokens = torch.randint(0, args.vocab_size, (micro_batch, seq_len), generator=gen).to(dev)
    act = lambda rows: torch.randn(rows, seq_len, d, device=dev).to(dt) * 0.5          # noqa: E731
    grad = lambda rows: torch.randn(rows * seq_len, d, device=dev).to(dt) * 0.01        # noqa: E731
    for u in units:
        st = PipelineStage(one_unit_model(u, False, False), 1, 3, dev)
        out[u] = run(st, act(st.in_rows * micro_batch), None, grad(st.out_rows * micro_batch))
        del st
    st = PipelineStage(one_unit_model(None, False, True), 2, 3, dev)
    out["head"] = run(st, act(micro_batch), tokens, None)
    del st
    st = PipelineStage(one_unit_model(None, True, False), 0, 3, dev)
    out["embed"] = run(st, tokens, None, grad(micro_batch))
    del st
    # Every one-unit measurement carries a fixed per-replay cost (two graph launches, the input / gradient copies into
    # the slot's static buffers) that a real stage pays once, not once per unit: a one-layer and a two-layer stage give
    # the true marginal cost of a layer and that fixed part, which is taken out of every unit before they are used.
    t1 = run(PipelineStage(one_unit_model("full", False, False, 1), 1, 3, dev), act(micro_batch), None, grad(micro_batch))
    t2 = run(PipelineStage(one_unit_model("full", False, False, 2), 1, 3, dev), act(micro_batch), None, grad(micro_batch))
    layer = max(t2 - t1, 0.25 * t1)
    fixed = max(0.0, t1 - layer)
    out["_raw"] = {**{k: v for k, v in out.items()}, "layer1": t1, "layer2": t2, "fixed": fixed}
    for k in list(units) + ["head", "embed"]:
        out[k] = max(out[k] - fixed, 0.1 * out[k])
    scale = layer / max(sum(out[u] for u in units), 1e-9)      # the units of a layer add up to the marginal layer cost
    for u in units:
        out[u] *= scale
    torch.cuda.synchronize(dev)
    torch.cuda.empty_cache()
    return out


def calibrate(args: ModelArgs, micro_batch: int, seq_len: int, device: torch.device, group=None, src: int = 0) -> dict:
    """Measure the unit costs on this GPU and make every rank of ``group`` use rank ``src``'s numbers (the partition
    must be identical on all ranks).  No-op on CPU devices (the analytic model stays in charge)."""
    import torch.distributed as dist
    dev = torch.device(device)
    if dev.type != "cuda":
        return unit_costs(args, seq_len)
    costs = measure_unit_costs(args, micro_batch, seq_len, dev)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        box = [costs]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, src) if group is not None else src, group=group)
        costs = box[0]
    set_measured_costs(args, seq_len, costs)
    return {**costs, "source": "measured"}
