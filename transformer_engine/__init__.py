"""Compatibility shim: lets the reference's te_llama*.py / te_qwen.py import unchanged
(`import transformer_engine as te`, te_llama.py:12) and land on llm_fp8_b200."""
from . import common, pytorch  # noqa: F401

__version__ = "2.5.0+llm_fp8_b200"
