"""`te.pytorch.*` names the reference touches: MultiheadAttention, LayerNormMLP, fp8_autocast, TransformerLayer (te_llama.py:44-79),
is_fp8_available (te_qwen.py:268), plus Linear / LayerNormLinear."""
from llm_fp8_b200.fp8 import fp8_autocast, is_fp8_available, is_mxfp8_available  # noqa: F401
from llm_fp8_b200.modules import (LayerNormLinear, LayerNormMLP, Linear, MultiheadAttention, RotaryPositionEmbedding,  # noqa: F401
                                  TransformerLayer)
from . import attention  # noqa: F401
