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


def balanced_ranges(n_layers: int, num_stages: int, args: Optional[ModelArgs] = None, seq_len: int = 1024) -> List[Tuple[int, int]]:
    """Contiguous split minimising the most expensive stage, with the LM head (and loss) charged to the last one."""
    if args is None:
        layer_cost, head_cost = 1.0, 1.0
    else:
        # FLOPs per token weighted by the efficiency each part reaches on B200, calibrated on measured per-action
        # times (profiles/bench_r1_n4_1f1b_v2.json: forward+backward of a decoder layer 1.00 ms, of embedding + final
        # norm + head + loss 1.46 ms per 4096-token micro-batch): the vocabulary-sized head GEMMs run near the
        # tensor-core roofline (~0.9), the d x d / d x f GEMMs of a layer at ~0.40, attention (exp-bound softmax) at
        # ~0.20; the cross-entropy pass is HBM-bound (reads and rewrites the [tokens, V] logits).
        d, f, V = args.dim, args.ffn_dim, args.vocab_size
        blocks = 2 if args.cross_attn else 1
        layer_cost = (blocks * 8 * d * d + (6 if args.activation == 'swiglu' else 4) * d * f) / 0.40 + blocks * 4 * seq_len * d * (0.5 if args.causal else 1.0) / 0.20
        head_cost = 2 * d * V / 0.9 + 650.0 * V
    # DP over (stage, first layer of the stage): minimise (most expensive stage, then sum of squared costs);
    # every stage gets at least one layer when there are enough layers.
    INF = (float("inf"), float("inf"))
    S, L = num_stages, n_layers
    min_layers = 1 if L >= S else 0
    best = [[INF] * (L + 1) for _ in range(S + 1)]   # best[s][i]: stages s.. cover layers i..L-1
    cut = [[L] * (L + 1) for _ in range(S + 1)]
    for i in range(L + 1):
        if L - i >= min_layers:
            c = (L - i) * layer_cost + head_cost
            best[S - 1][i] = (c, c * c)
    for s in range(S - 2, -1, -1):
        for i in range(L + 1):
            for j in range(i + min_layers, L + 1):
                nxt = best[s + 1][j]
                if nxt[0] == float("inf"):
                    continue
                c = (j - i) * layer_cost
                w = (max(c, nxt[0]), c * c + nxt[1])
                if w < best[s][i]:
                    best[s][i], cut[s][i] = w, j
    out, i = [], 0
    for s in range(S):
        j = cut[s][i] if s < S - 1 else L
        out.append((i, j))
        i = j
    return out


def _min_max_partition(costs: List[float], num_stages: int, tail_cost: float) -> List[Tuple[int, int]]:
    """Contiguous split of ``costs`` into ``num_stages`` non-empty groups minimising (largest group, sum of squares);
    ``tail_cost`` is added to the last group."""
    n, S = len(costs), num_stages
    pre = [0.0]
    for c in costs:
        pre.append(pre[-1] + c)
    INF = (float("inf"), float("inf"))
    best = [[INF] * (n + 1) for _ in range(S + 1)]
    cut = [[n] * (n + 1) for _ in range(S + 1)]
    for i in range(n):
        c = pre[n] - pre[i] + tail_cost
        best[S - 1][i] = (c, c * c)
    for s in range(S - 2, -1, -1):
        for i in range(n):
            for j in range(i + 1, n):
                nxt = best[s + 1][j]
                if nxt[0] == float("inf"):
                    continue
                c = pre[j] - pre[i]
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
    """Stage boundaries at HALF-layer granularity: unit 2i = the attention sub-blocks of layer i (post-LN reference block:
    self + cross attention, ends at the input of the feed-forward block; pre-LN block: attention + its residual add),
    unit 2i+1 = its feed-forward block.
    Returns [u0, u1) unit ranges.  Only one [tokens, d] tensor crosses such a cut (the cross-attention memory is the
    layer input, so a cut between the two attention blocks would have to ship two).  At 8 stages the flagship model
    goes from a most expensive stage of 4.0 layer-equivalents (whole layers) to 3.7."""
    d, f, V = args.dim, args.ffn_dim, args.vocab_size
    blocks = 2 if args.cross_attn else 1
    attn_cost = blocks * 8 * d * d / 0.40 + blocks * 4 * seq_len * d * (0.5 if args.causal else 1.0) / 0.20
    ffn_cost = (6 if args.activation == "swiglu" else 4) * d * f / 0.40
    head_cost = 2 * d * V / 0.9 + 650.0 * V
    costs = [attn_cost if u % 2 == 0 else ffn_cost for u in range(2 * n_layers)]
    if 2 * n_layers < num_stages:
        raise ValueError(f"{n_layers} layers cannot be cut into {num_stages} stages")
    return _min_max_partition(costs, num_stages, head_cost)


def stage_ranges(n_layers: int, num_stages: int, rule: str = "reference", args: Optional[ModelArgs] = None,
                 seq_len: int = 1024) -> List[Tuple[int, int]]:
    """[start, end) LAYER ranges per stage.  For rule 'fine' a boundary layer appears in both neighbouring stages
    (see ``stage_layer_parts`` for which half each of them owns)."""
    if rule == "reference":
        return reference_ranges(n_layers, num_stages)
    if rule == "balanced":
        return balanced_ranges(n_layers, num_stages, args, seq_len)
    if rule == "fine":
        return [(u0 // 2, (u1 + 1) // 2) for u0, u1 in _fine_units(n_layers, num_stages, args, seq_len)]
    raise ValueError(f"unknown partition rule {rule!r}")


def _fine_units(n_layers: int, num_stages: int, args: Optional[ModelArgs], seq_len: int):
    if args is None:
        return [(2 * a, 2 * b) for a, b in balanced_ranges(n_layers, num_stages, args, seq_len)]
    return fine_unit_ranges(n_layers, num_stages, args, seq_len)


def stage_layer_parts(n_layers: int, num_stages: int, stage_index: int, rule: str, args: Optional[ModelArgs] = None,
                      seq_len: int = 1024) -> List[Tuple[int, str]]:
    """[(layer index, 'full' | 'attn' | 'ffn'), ...] owned by ``stage_index``."""
    if rule != "fine":
        start, end = stage_ranges(n_layers, num_stages, rule, args, seq_len)[stage_index]
        return [(i, "full") for i in range(start, end)]
    u0, u1 = _fine_units(n_layers, num_stages, args, seq_len)[stage_index]
    out = []
    for i in range(u0 // 2, (u1 + 1) // 2):
        has_attn, has_ffn = u0 <= 2 * i < u1, u0 <= 2 * i + 1 < u1
        out.append((i, "full" if (has_attn and has_ffn) else ("attn" if has_attn else "ffn")))
    return out


_ATTN_MODULES = ("self_attn", "multihead_attn", "norm1", "norm2")
_FFN_MODULES = ("linear1", "linear2", "norm3")


def _apply_part(layer: DecoderLayer, part: str) -> DecoderLayer:
    layer.part = part
    for name in (_FFN_MODULES if part == "attn" else _ATTN_MODULES if part == "ffn" else ()):
        if hasattr(layer, name):
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
