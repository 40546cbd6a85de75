"""`from transformer_engine.common.recipe import Format, DelayedScaling, MXFP8BlockScaling` (te_llama.py:25, te_llama_mxfp8.py:25)."""
from llm_fp8_b200.recipe import DelayedScaling, Format, MXFP8BlockScaling, Recipe  # noqa: F401
