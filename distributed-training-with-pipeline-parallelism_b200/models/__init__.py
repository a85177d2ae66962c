"""Model family: the helper-compatible Transformer and its explicit stage program."""
from .args import ModelArgs, PRESETS, preset  # noqa: F401
from .transformer import (Transformer, DecoderLayer, Ctx, GradSink, layer_forward, layer_backward, stage_forward,  # noqa: F401
                          stage_backward, stage_loss, site_seed)
