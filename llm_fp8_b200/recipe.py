"""FP8 recipes — the configuration surface of the path (mirrors transformer_engine.common.recipe as the reference
uses it: te_llama.py:25,39-40; te_llama_hybrid.py:39; te_llama_mxfp8.py:25,28-29; train_fp8.py:136-165)."""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum
from typing import NamedTuple


class _FormatHelper(NamedTuple):
    max_fwd: float
    max_bwd: float


class Format(Enum):
    """E4M3: 448 both ways; E5M2: 57344 both ways; HYBRID: E4M3 forward tensors, E5M2 gradients."""
    E4M3 = _FormatHelper(max_fwd=448.0, max_bwd=448.0)
    E5M2 = _FormatHelper(max_fwd=57344.0, max_bwd=57344.0)
    HYBRID = _FormatHelper(max_fwd=448.0, max_bwd=57344.0)


FMT_E4M3, FMT_E5M2 = 0, 1


def fwd_fmt(fmt: Format) -> int:
    return FMT_E5M2 if fmt is Format.E5M2 else FMT_E4M3


def bwd_fmt(fmt: Format) -> int:
    return FMT_E4M3 if fmt is Format.E4M3 else FMT_E5M2


class Recipe:
    def delayed(self) -> bool:
        return isinstance(self, DelayedScaling)

    def mxfp8(self) -> bool:
        return isinstance(self, MXFP8BlockScaling)


@dataclass
class DelayedScaling(Recipe):
    """Per-tensor scaling from an amax history (defaults follow TE: margin 0, HYBRID, history 1024, "max")."""
    margin: int = 0
    fp8_format: Format = Format.HYBRID
    amax_history_len: int = 1024
    amax_compute_algo: str = "max"
    reduce_amax: bool = True
    fp8_dpa: bool = False
    fp8_mha: bool = False

    def __post_init__(self):
        if self.amax_compute_algo != "max":
            raise NotImplementedError("only amax_compute_algo='max' is built (the reference uses 'max': te_llama.py:39-40)")
        if self.amax_history_len < 1:
            raise ValueError("amax_history_len must be >= 1")
        if self.fp8_dpa or self.fp8_mha:
            raise NotImplementedError("FP8 attention is out of scope (reference runs bf16 flash-attention)")

    def key(self):
        return ("delayed", self.fp8_format.name, self.margin, self.amax_history_len, self.amax_compute_algo, self.reduce_amax)


@dataclass
class MXFP8BlockScaling(Recipe):
    """E8M0 scale per 32 contiguous elements along the contraction dimension; E4M3 data."""
    margin: int = 0
    fp8_format: Format = Format.E4M3
    fp8_dpa: bool = False
    fp8_mha: bool = False

    def __post_init__(self):
        if self.fp8_format is not Format.E4M3:
            raise NotImplementedError("MXFP8BlockScaling is built for Format.E4M3 (te_llama_mxfp8.py:28)")

    def key(self):
        return ("mxfp8", self.fp8_format.name)


def get_default_fp8_recipe() -> Recipe:
    return DelayedScaling()
