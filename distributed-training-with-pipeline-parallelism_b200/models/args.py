"""Model hyper-parameters.

``ModelArgs()`` with no arguments is the reference's model exactly (/root/reference/LLMsDistributedTrainingHelper.py:23-28:
dim 768, 8 layers, 8 heads, vocab 10000) including everything the reference inherits silently from
``nn.TransformerDecoderLayer`` defaults: FFN width fixed at 2048, dropout 0.1, ReLU,
post-LN, bias, eps 1e-5, bidirectional self-attention plus a cross-attention block over the layer's own input.
The extra fields make the north-star configs (GPT-2 medium / XL, Llama-style) expressible.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Dict

import torch


@dataclass
class ModelArgs:
    dim: int = 768
    n_layers: int = 8
    n_heads: int = 8
    vocab_size: int = 10000
    # ---- extensions; the defaults reproduce the reference ----
    ffn_dim: int = 2048          # torch default dim_feedforward (NOT 4*dim)
    dropout: float = 0.1         # live in the reference (modules stay in train mode)
    causal: bool = False         # the reference passes no tgt_mask (helper:52)
    cross_attn: bool = True      # layer(h, h): cross-attention whose memory is the layer input
    activation: str = "relu"     # 'relu' | 'gelu' | 'swiglu' (pre-LN block only; linear1 packs gate|up = [2*ffn_dim, dim])
    norm_first: bool = False     # post-LN like the reference; True = GPT-style pre-LN
    norm: str = "layernorm"      # 'layernorm' | 'rmsnorm'
    layer_norm_eps: float = 1e-5
    rope: bool = False           # rotary position embedding on q, k (pre-LN block only); the reference has no positions
    rope_theta: float = 10000.0
    dtype: torch.dtype = torch.float32   # parameter/activation dtype: fp32 on CPU, bf16 on B200

    def __post_init__(self):
        if self.dim % self.n_heads != 0:
            raise ValueError(f"dim ({self.dim}) must be divisible by n_heads ({self.n_heads})")
        if self.activation not in ("relu", "gelu", "swiglu"):
            raise ValueError(f"unknown activation {self.activation!r}")
        if self.norm not in ("layernorm", "rmsnorm"):
            raise ValueError(f"unknown norm {self.norm!r}")
        if self.norm_first and self.cross_attn:
            raise ValueError("the pre-LN (GPT) block has no cross-attention; set cross_attn=False")
        if (self.activation == "swiglu" or self.rope) and not self.norm_first:
            raise ValueError("swiglu / rope are implemented for the pre-LN block (norm_first=True)")
        if self.rope and (self.dim // self.n_heads) % 16:
            raise ValueError("rope needs head_dim % 16 == 0")

    @property
    def head_dim(self) -> int:
        return self.dim // self.n_heads

    def params_per_layer(self) -> int:
        d, f = self.dim, self.ffn_dim
        attn = 4 * d * d + 4 * d
        up = 2 if self.activation == "swiglu" else 1
        n = attn * (2 if self.cross_attn else 1) + (1 + up) * d * f + up * f + d
        n += (3 if self.cross_attn else 2) * 2 * d
        return n

    def num_params(self) -> int:
        d, V = self.dim, self.vocab_size
        return self.n_layers * self.params_per_layer() + V * d + 2 * d + V * d + V

    def flops_per_token(self, seq_len: int) -> float:
        """fwd+bwd model FLOPs per token (3 x forward; attention counted un-masked unless causal)."""
        d, f, V = self.dim, self.ffn_dim, self.vocab_size
        attn_blocks = 2 if self.cross_attn else 1
        gemm = 2 * (attn_blocks * 4 * d * d + (3 if self.activation == "swiglu" else 2) * d * f)
        attn = attn_blocks * 4 * seq_len * d * (0.5 if self.causal else 1.0)
        return 3.0 * (self.n_layers * (gemm + attn) + 2 * d * V)


PRESETS: Dict[str, ModelArgs] = {
    # the reference's sweep model (helper:23-28, nb:346-349); L/H are overridden per experiment
    "ref-small": ModelArgs(),
    # "same LM scaled to GPT-2 medium" (BASELINE.json config 2): the helper's block at GPT-2-medium width/depth/vocab.
    # The helper cannot change the FFN width, so it stays 2048 -- this is what the reference arm can run too.
    "ref-gpt2-medium": ModelArgs(dim=1024, n_layers=24, n_heads=16, vocab_size=50257),
    "ref-gpt2-xl": ModelArgs(dim=1600, n_layers=48, n_heads=25, vocab_size=50257),
    # GPT-style blocks (pre-LN, causal, GELU, 4*dim FFN, no cross-attention)
    "gpt2-medium": ModelArgs(dim=1024, n_layers=24, n_heads=16, vocab_size=50257, ffn_dim=4096, causal=True,
                             cross_attn=False, activation="gelu", norm_first=True),
    "gpt2-xl": ModelArgs(dim=1600, n_layers=48, n_heads=25, vocab_size=50257, ffn_dim=6400, causal=True,
                         cross_attn=False, activation="gelu", norm_first=True),
    # Llama width/depth with the GPT (GELU, learned-position-free) block: the pre-SwiGLU/RoPE comparison point
    # Llama-style block: pre-LN RMSNorm, causal attention with rotary embeddings, SwiGLU FFN, no dropout
    "llama-7b": ModelArgs(dim=4096, n_layers=32, n_heads=32, vocab_size=32000, ffn_dim=11008, causal=True, cross_attn=False,
                          activation="swiglu", norm_first=True, norm="rmsnorm", dropout=0.0, rope=True),
    "llama-7b-shape": ModelArgs(dim=4096, n_layers=32, n_heads=32, vocab_size=32000, ffn_dim=11008, causal=True,
                                cross_attn=False, activation="gelu", norm_first=True, norm="rmsnorm", dropout=0.0),
}


def preset(name: str, **overrides) -> ModelArgs:
    if name not in PRESETS:
        raise KeyError(f"unknown model preset {name!r}; have {sorted(PRESETS)}")
    return replace(PRESETS[name], **overrides)
