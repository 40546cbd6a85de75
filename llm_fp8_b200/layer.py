"""The linear-layer path of one decoder layer, as the reference's TELlamaDecoderLayer drives it (te_llama.py:41-82):

    with fp8_autocast(attn_recipe):  h += proj( attention_core( LayerNormLinear_qkv(h) ) )
    with fp8_autocast(mlp_recipe):   h += LayerNormMLP(h)

`DecoderLinearPath` keeps exactly the FP8 part of that structure — the four linears, two fused RMSNorms, SwiGLU and the
two residual adds — with the bf16 attention core (flash-attention in the reference, README.md:27; not on the FP8 path)
replaced by the identity on the query slice, so every FLOP it executes is a linear-layer FLOP.  bench.py, smoke() and the
data-parallel tests run stacks of it; the full block with a real attention core is modules.MultiheadAttention.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn

from .fp8 import fp8_autocast
from .modules import LayerNormLinear, LayerNormMLP, Linear
from .recipe import DelayedScaling, Format, MXFP8BlockScaling, Recipe

# H, I, heads, kv_heads, head_dim, layers   (public HF configs)
MODEL_SHAPES = {
    "llama-3.2-1b": (2048, 8192, 32, 8, 64, 16),
    "llama-3.2-3b": (3072, 8192, 24, 8, 128, 28),
    "llama-3.1-8b": (4096, 14336, 32, 8, 128, 32),
    "qwen2.5-14b": (5120, 13824, 40, 8, 128, 48),
}


def recipes_for(scenario: str):
    """(attn_recipe, mlp_recipe) of the reference's three decoder swaps: 'default' = te_llama.py:39-40,
    'hybrid' = te_llama_hybrid.py:39, 'mxfp8' = te_llama_mxfp8.py:28-29."""
    if scenario == "default":
        return (DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max"),
                DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16, amax_compute_algo="max"))
    if scenario == "hybrid":
        r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
        return r, r
    if scenario == "mxfp8":
        r = MXFP8BlockScaling(fp8_format=Format.E4M3)
        return r, r
    raise ValueError(f"unknown fp8 scenario {scenario!r} (default | hybrid | mxfp8)")


def linear_params_per_layer(H: int, I: int, heads: int, kv: int, hd: int) -> int:
    """P of SURVEY.md §8(d): weight elements of qkv + proj + fc1 + fc2 (fwd+bwd linear FLOPs of a layer = 6*M*P)."""
    return (heads * hd + 2 * kv * hd) * H + H * heads * hd + 2 * I * H + H * I


class _QuerySlice(torch.autograd.Function):
    """The attention-core stand-in: context = the query columns of the fused qkv output, handed on as a VIEW (the proj quantiser
    reads it through its row pitch, no copy).  Backward writes dq next to a zero dk|dv block with one vectorised concat instead
    of autograd's zero-fill + strided copy — a real attention backward would emit the whole [M, q+2kv] gradient itself."""

    @staticmethod
    def forward(ctx, qkv, q_out):
        ctx.rest = qkv.shape[-1] - q_out
        return qkv[..., :q_out]

    @staticmethod
    def backward(ctx, dq):
        if ctx.rest == 0:
            return dq, None
        # the Linear that consumed the slice hands dq over as columns [0, q) of a zero-padded [M, q + rest] buffer its dgrad GEMM wrote
        # in place (functional._padded_dx): that buffer IS the gradient of qkv
        b = dq._base
        W = dq.shape[-1] + ctx.rest
        if b is not None and getattr(b, "_llmfp8_pad", None) == (0, dq.shape[-1]) and b.shape[-1] == W and b.is_contiguous() \
                and b.numel() == dq.numel() // dq.shape[-1] * W:
            return b.view(*dq.shape[:-1], W), None
        return torch.cat((dq, dq.new_zeros(*dq.shape[:-1], ctx.rest)), dim=-1), None


class DecoderLinearPath(nn.Module):
    """`fuse_residual` (default on; LLMFP8_FUSE_RESIDUAL=0 turns it off): the two residual adds ride in the proj / fc2 GEMM epilogues
    and their gradients in the RMSNorm backward kernels — four [M, H] elementwise passes per layer fewer than the plain
    `h = h + f(h)` form, which stays available as the reference-shaped path (what te_llama.py:78, 81 spell out)."""

    def __init__(self, H: int, I: int, heads: int, kv: int, hd: int, attn_recipe: Recipe, mlp_recipe: Recipe, rms_eps: float = 1e-5,
                 params_dtype: torch.dtype = torch.bfloat16, device=None, fuse_residual: Optional[bool] = None):
        super().__init__()
        q_out, kv_out = heads * hd, kv * hd
        self.q_out = q_out
        self.fuse_residual = (os.environ.get("LLMFP8_FUSE_RESIDUAL", "1") != "0") if fuse_residual is None else fuse_residual
        self.attn_recipe, self.mlp_recipe = attn_recipe, mlp_recipe
        # same constructor arguments the reference passes (te_llama.py:45-63): attention without bias and with the model's eps,
        # LayerNormMLP with the TE defaults (bias=True, eps=1e-5)
        self.layernorm_qkv = LayerNormLinear(H, q_out + 2 * kv_out, eps=rms_eps, bias=False, normalization="RMSNorm",
                                             parameters_split={"query": q_out, "key": kv_out, "value": kv_out},
                                             params_dtype=params_dtype, device=device)
        self.proj = Linear(q_out, H, bias=False, params_dtype=params_dtype, device=device)
        self.layernorm_mlp = LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=params_dtype, device=device)

    def forward(self, h: torch.Tensor, fp8_group=None) -> torch.Tensor:
        if self.fuse_residual:
            with fp8_autocast(enabled=True, fp8_recipe=self.attn_recipe, fp8_group=fp8_group):
                qkv, h_res = self.layernorm_qkv(h, return_input=True)
                h = self.proj(_QuerySlice.apply(qkv, self.q_out), residual=h_res)
            with fp8_autocast(enabled=True, fp8_recipe=self.mlp_recipe, fp8_group=fp8_group):
                return self.layernorm_mlp(h, add_residual=True)
        with fp8_autocast(enabled=True, fp8_recipe=self.attn_recipe, fp8_group=fp8_group):
            qkv = self.layernorm_qkv(h)
            a = self.proj(_QuerySlice.apply(qkv, self.q_out))
        h = h + a
        with fp8_autocast(enabled=True, fp8_recipe=self.mlp_recipe, fp8_group=fp8_group):
            f = self.layernorm_mlp(h)
        return h + f


class LinearPathStack(nn.Module):
    """`layers` DecoderLinearPath blocks back to back (what a data-parallel rank runs per step)."""

    def __init__(self, model: str, scenario: str, layers: Optional[int] = None, device=None, seed: int = 1234, fuse_residual: Optional[bool] = None):
        super().__init__()
        H, I, heads, kv, hd, L = MODEL_SHAPES[model]
        self.H, self.I = H, I
        self.n_layers = layers or L
        self.params_per_layer = linear_params_per_layer(H, I, heads, kv, hd)
        ar, mr = recipes_for(scenario)
        torch.manual_seed(seed)
        self.layers = nn.ModuleList(DecoderLinearPath(H, I, heads, kv, hd, ar, mr, device=device, fuse_residual=fuse_residual)
                                    for _ in range(self.n_layers))

    def forward(self, h: torch.Tensor) -> torch.Tensor:
        for layer in self.layers:
            h = layer(h)
        return h

    def linear_flops(self, tokens: int) -> float:
        return 6.0 * tokens * self.params_per_layer * self.n_layers
