"""`from transformer_engine.pytorch.attention import RotaryPositionEmbedding` (te_llama.py:13)."""
from llm_fp8_b200.modules import MultiheadAttention, RotaryPositionEmbedding, apply_rotary_pos_emb  # noqa: F401
