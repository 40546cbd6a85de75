"""llm_fp8_b200 — B200-native (sm_100a) FP8 / MXFP8 linear-layer path behind the Transformer-Engine plugin surface
that xuanvinh1997/llm-fp8's decoder swap uses (te_llama.py, te_llama_hybrid.py, te_llama_mxfp8.py).

Python here is host plumbing only (parameters, autograd wiring, recipe state); all arithmetic is in
libllmfp8_b200.so (llm_fp8_b200/csrc, C ABI in include/llmfp8_b200.h).  There is no CPU or library fallback.
"""
from . import _lib, ops
from .fp8 import (flush_all, fp8_autocast, get_fp8_recipe, is_fp8_available, is_fp8_enabled, is_mxfp8_available, reset_fp8_state)
from .modules import (LayerNormLinear, LayerNormMLP, Linear, MultiheadAttention, RotaryPositionEmbedding, TransformerLayer,
                      apply_rotary_pos_emb)
from .lm_head import LMHeadLoss, bf16_gemm, lm_head_cross_entropy
from .optim import FusedAdamW
from .recipe import DelayedScaling, Format, MXFP8BlockScaling

__all__ = ["Linear", "LayerNormLinear", "LayerNormMLP", "MultiheadAttention", "RotaryPositionEmbedding", "TransformerLayer",
           "fp8_autocast", "is_fp8_available", "is_mxfp8_available", "DelayedScaling", "MXFP8BlockScaling", "Format", "ops", "LMHeadLoss", "lm_head_cross_entropy", "bf16_gemm", "FusedAdamW"]
