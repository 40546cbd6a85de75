"""Linear / LayerNormLinear / LayerNormMLP / MultiheadAttention — the plugin surface the reference's decoder swap builds
(te_llama.py:45-66; te_llama_hybrid.py:44-65; te_llama_mxfp8.py:53-76; te_qwen.py:76-104) with the same constructor
arguments, parameter names (te_llama.py:181-239) and defaults (bias=True, eps=1e-5, params in the default dtype).

Inside `fp8_autocast(enabled=True)` every GEMM runs in FP8 through the C ABI; outside it (or with enabled=False) the
modules raise: this package is the FP8 path only and carries no bf16 fallback implementation of its own.
"""
from __future__ import annotations

import math
import os
from typing import Optional, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import functional as fn
from .fp8 import get_fp8_recipe, is_fp8_enabled
from .recipe import DelayedScaling, MXFP8BlockScaling


def _init_weight(shape, dtype, device, std=0.023):
    w = torch.empty(shape, dtype=dtype, device=device)
    nn.init.normal_(w, mean=0.0, std=std)
    return nn.Parameter(w)


def _check_active(name: str):
    if not is_fp8_enabled():
        raise RuntimeError(f"{name}: call inside fp8_autocast(enabled=True); llm_fp8_b200 implements the FP8 path only")


class _Fp8Module(nn.Module):
    """Carries the delayed-scaling slot bookkeeping and round-trips it through state_dict as `_extra_state`."""
    _N_FWD, _N_BWD = 3, 2

    def __init__(self):
        super().__init__()
        self._meta = fn.Fp8Meta(self._N_FWD, self._N_BWD)
        # quantised weight copies kept between the micro-batches of one optimizer step (forward(..., is_first_microbatch=True / False);
        # None = cast on every call, what the reference's training loop gets: it never passes the argument — train_fp8.py:276-291)
        self._wcache = {}

    def _apply(self, fn_, *args, **kwargs):
        self._wcache.clear()                 # .cuda() / .to() replace the parameters the copies were cast from
        return super()._apply(fn_, *args, **kwargs)

    def get_extra_state(self):
        """This module's delayed-scaling slots, exactly as they stand — including amaxes recorded but not yet folded into the scales
        (`pending`), so a restore resumes mid-window without a collective (state_dict() is often called on one rank only)."""
        out = {}
        for (rkey, dev), s in self._meta._slots.items():
            pf, pb = s["pf"], s["pb"]
            out[str((rkey, "fwd"))] = {k: v[..., s["f0"]:s["f0"] + self._N_FWD].detach().cpu() for k, v in
                                       (("amax_history", pf.amax_history), ("scale", pf.scale), ("scale_inv", pf.scale_inv))}
            out[str((rkey, "fwd"))]["pending"] = bool(pf.pending)
            out[str((rkey, "bwd"))] = {k: v[..., s["b0"]:s["b0"] + self._N_BWD].detach().cpu() for k, v in
                                       (("amax_history", pb.amax_history), ("scale", pb.scale), ("scale_inv", pb.scale_inv))}
            out[str((rkey, "bwd"))]["pending"] = bool(pb.pending)
        return out

    def set_extra_state(self, state):
        """Only this package's own dict format is understood.  A checkpoint written with Transformer Engine modules carries a pickled
        uint8 tensor / BytesIO here: it is ignored with a warning (the delayed-scaling windows restart empty; weights load normally)."""
        if state is None or (isinstance(state, dict) and not state):
            self._pending_extra_state = None
            return
        if not isinstance(state, dict):
            if not (isinstance(state, torch.Tensor) and state.numel() == 0):
                import warnings
                warnings.warn(f"{type(self).__name__}: ignoring an _extra_state of type {type(state).__name__} (not written by llm_fp8_b200); "
                              "delayed-scaling state starts fresh")
            self._pending_extra_state = None
            return
        self._pending_extra_state = state      # applied when the slots are claimed for that recipe

    def _apply_extra_state(self, recipe, device):
        st = getattr(self, "_pending_extra_state", None)
        if not st:
            return
        s = self._meta.get(recipe, device)
        for direction, pool, off, n in (("fwd", s["pf"], s["f0"], self._N_FWD), ("bwd", s["pb"], s["b0"], self._N_BWD)):
            blob = st.get(str((recipe.key(), direction)))
            if blob is None:
                continue
            pool.amax_history[:, off:off + n].copy_(blob["amax_history"])
            pool.scale[off:off + n].copy_(blob["scale"])
            pool.scale_inv[off:off + n].copy_(blob["scale_inv"])
            pool.pending = pool.pending or bool(blob.get("pending", False))
        self._pending_extra_state = None


class Linear(_Fp8Module):
    """y = x W^T + b in FP8.  Parameters: `weight` [out, in], `bias` [out] (te_llama.py:213-216 copies `proj.weight`)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, params_dtype: Optional[torch.dtype] = None,
                 device: Union[torch.device, str, None] = None, init_method=None, **_unused):
        super().__init__()
        dtype = params_dtype or torch.get_default_dtype()
        self.in_features, self.out_features = in_features, out_features
        self.weight = _init_weight((out_features, in_features), dtype, device)
        if init_method is not None:
            init_method(self.weight)
        self.bias = nn.Parameter(torch.zeros(out_features, dtype=dtype, device=device)) if bias else None

    def forward(self, inp: torch.Tensor, is_first_microbatch=None, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        """`residual` (an extension over the TE surface): returns residual + Linear(inp), the add fused into the GEMM epilogue —
        the `h + proj(...)` of TELlamaDecoderLayer.forward (te_llama.py:78) without the separate elementwise pass."""
        _check_active("Linear")
        recipe = get_fp8_recipe()
        g = torch.is_grad_enabled()
        first = is_first_microbatch
        if isinstance(recipe, MXFP8BlockScaling):
            out = fn.MXLinearFn.apply(inp, self.weight, self.bias, None, 0.0, g, residual, False, self._wcache, first)
        else:
            self._apply_extra_state(recipe, inp.device)
            out = fn.LinearFn.apply(inp, self.weight, self.bias, self._meta, recipe, None, 0.0, g, residual, False, self._wcache, first)
        return out

_JOIN_SPLIT_WEIGHTS = os.environ.get("LLMFP8_JOIN_SPLIT_WEIGHTS", "1") != "0"


def _joined_rows(params, joined: Optional[torch.Tensor]) -> torch.Tensor:
    """-> the [sum of rows, in] tensor whose consecutive row slices ARE the storage of `params`.  Parameters are separate tensors after
    construction and after anything that goes through Module._apply (.cuda(), .to(dtype)) or re-assigns them; whenever the aliasing
    does not hold (checked by address) the rows are gathered once into a new buffer and each Parameter's .data is re-pointed at its
    slice.  In-place updates (optimizer steps, load_state_dict, replace_params' .data.copy_: te_llama.py:196-238) keep it."""
    p0 = params[0]
    cols, item = p0.shape[1], p0.element_size()
    ok = joined is not None and joined.device == p0.device and joined.dtype == p0.dtype and joined.shape[1] == cols
    if ok:
        addr = joined.data_ptr()
        for p in params:
            ok = ok and p.dtype == p0.dtype and p.is_contiguous() and p.data_ptr() == addr
            addr += p.shape[0] * cols * item
        ok = ok and addr == joined.data_ptr() + joined.numel() * item
    if ok:
        return joined
    with torch.no_grad():
        joined = torch.cat([p.detach() for p in params], dim=0)
        off = 0
        for p in params:
            p.data = joined[off:off + p.shape[0]]
            off += p.shape[0]
    return joined


class _JoinRows(torch.autograd.Function):
    """joined buffer in, the same memory out — with the split Parameters as autograd inputs, so that the weight gradient [out, in] a
    linear Function returns is routed to them as row slices (views: no split copies either)."""

    @staticmethod
    def forward(ctx, joined, *params):
        ctx.sizes = [p.shape[0] for p in params]
        return joined.detach()

    @staticmethod
    def backward(ctx, dw):
        return (None,) + tuple(dw.split(ctx.sizes, dim=0))


class LayerNormLinear(_Fp8Module):
    """RMSNorm fused into the FP8 input cast, then one GEMM.  With `parameters_split` the weight is exposed as separate
    `<name>_weight` Parameters (query/key/value: te_llama.py:198-211) that feed a single GEMM."""

    def __init__(self, in_features: int, out_features: int, eps: float = 1e-5, bias: bool = True, normalization: str = "LayerNorm",
                 parameters_split: Union[None, Tuple[str, ...], dict] = None, params_dtype: Optional[torch.dtype] = None,
                 device=None, zero_centered_gamma: bool = False, return_layernorm_output: bool = False, init_method=None, **_unused):
        super().__init__()
        if normalization != "RMSNorm":
            raise NotImplementedError("only normalization='RMSNorm' is built (the reference uses RMSNorm: te_llama.py:51,61)")
        if zero_centered_gamma or return_layernorm_output:
            raise NotImplementedError("zero_centered_gamma / return_layernorm_output are not used by the reference path")
        dtype = params_dtype or torch.get_default_dtype()
        self.in_features, self.out_features, self.eps = in_features, out_features, eps
        self.layer_norm_weight = nn.Parameter(torch.ones(in_features, dtype=dtype, device=device))
        if parameters_split is None:
            self._split = None
            self.weight = _init_weight((out_features, in_features), dtype, device)
            if init_method is not None:
                init_method(self.weight)
            self.bias = nn.Parameter(torch.zeros(out_features, dtype=dtype, device=device)) if bias else None
        else:
            if isinstance(parameters_split, dict):
                names, sizes = list(parameters_split.keys()), list(parameters_split.values())
            else:
                names = list(parameters_split)
                sizes = [out_features // len(names)] * len(names)
            if sum(sizes) != out_features:
                raise ValueError("parameters_split sizes must sum to out_features")
            self._split = [n if n.endswith("_") else n + "_" for n in names]
            for n, sz in zip(self._split, sizes):
                w = _init_weight((sz, in_features), dtype, device)
                if init_method is not None:
                    init_method(w)
                self.register_parameter(n + "weight", w)
                self.register_parameter(n + "bias", nn.Parameter(torch.zeros(sz, dtype=dtype, device=device)) if bias else None)
            self.use_bias = bias

    def _weight_bias(self):
        if self._split is None:
            return self.weight, self.bias
        ws = [getattr(self, n + "weight") for n in self._split]
        if _JOIN_SPLIT_WEIGHTS:
            # the split Parameters live as row slices of ONE [out, in] buffer: the GEMM path reads it in place (no torch.cat of the
            # weight per forward: 2 x 50 MB of traffic for the qkv weight at 8B) and the wgrad output is handed back as row slices
            self._joined = _joined_rows(ws, getattr(self, "_joined", None))
            w = _JoinRows.apply(self._joined, *ws)
        else:
            w = torch.cat(ws, dim=0)
        b = torch.cat([getattr(self, n + "bias") for n in self._split], dim=0) if self.use_bias else None
        return w, b

    def forward(self, inp: torch.Tensor, is_first_microbatch=None, return_input: bool = False):
        """`return_input` (an extension over the TE surface): also returns an alias of `inp` to take the residual branch from; its
        gradient then arrives in this module's backward and is folded into the RMSNorm backward kernel."""
        _check_active("LayerNormLinear")
        recipe = get_fp8_recipe()
        w, b = self._weight_bias()
        g = torch.is_grad_enabled()
        first = is_first_microbatch
        if isinstance(recipe, MXFP8BlockScaling):
            out = fn.MXLinearFn.apply(inp, w, b, self.layer_norm_weight, self.eps, g, None, return_input, self._wcache, first)
        else:
            self._apply_extra_state(recipe, inp.device)
            out = fn.LinearFn.apply(inp, w, b, self._meta, recipe, self.layer_norm_weight, self.eps, g, None, return_input, self._wcache, first)
        return out

class LayerNormMLP(_Fp8Module):
    """RMSNorm -> fc1 (+bias) -> SwiGLU -> fc2 (+bias).  Parameters: layer_norm_weight [H], fc1_weight [2I, H] (gate rows first:
    te_llama.py:225-233), fc1_bias [2I], fc2_weight [H, I], fc2_bias [H].  bias=True and eps=1e-5 by default: the reference
    constructs it without either argument (te_llama.py:58-63)."""
    _N_FWD, _N_BWD = 6, 4

    def __init__(self, hidden_size: int, ffn_hidden_size: int, eps: float = 1e-5, bias: bool = True, normalization: str = "LayerNorm",
                 activation: str = "gelu", params_dtype: Optional[torch.dtype] = None, device=None, zero_centered_gamma: bool = False,
                 return_layernorm_output: bool = False, init_method=None, output_layer_init_method=None, **_unused):
        super().__init__()
        if normalization != "RMSNorm" or activation != "swiglu":
            raise NotImplementedError("only normalization='RMSNorm', activation='swiglu' are built (te_llama.py:61-62)")
        if zero_centered_gamma or return_layernorm_output:
            raise NotImplementedError("zero_centered_gamma / return_layernorm_output are not used by the reference path")
        dtype = params_dtype or torch.get_default_dtype()
        self.hidden_size, self.ffn_hidden_size, self.eps = hidden_size, ffn_hidden_size, eps
        self.layer_norm_weight = nn.Parameter(torch.ones(hidden_size, dtype=dtype, device=device))
        self.fc1_weight = _init_weight((2 * ffn_hidden_size, hidden_size), dtype, device)
        self.fc2_weight = _init_weight((hidden_size, ffn_hidden_size), dtype, device)
        if init_method is not None:
            init_method(self.fc1_weight)
        if output_layer_init_method is not None:
            output_layer_init_method(self.fc2_weight)
        self.fc1_bias = nn.Parameter(torch.zeros(2 * ffn_hidden_size, dtype=dtype, device=device)) if bias else None
        self.fc2_bias = nn.Parameter(torch.zeros(hidden_size, dtype=dtype, device=device)) if bias else None

    def forward(self, inp: torch.Tensor, is_first_microbatch=None, add_residual: bool = False) -> torch.Tensor:
        """`add_residual` (an extension over the TE surface): returns inp + MLP(inp) — the `h + layernorm_mlp(h)` of
        TELlamaDecoderLayer.forward (te_llama.py:81) with the add in the fc2 GEMM epilogue and its gradient in the RMSNorm backward."""
        _check_active("LayerNormMLP")
        recipe = get_fp8_recipe()
        g = torch.is_grad_enabled()
        first = is_first_microbatch
        if isinstance(recipe, MXFP8BlockScaling):
            out = fn.MXLayerNormMLPFn.apply(inp, self.layer_norm_weight, self.fc1_weight, self.fc1_bias, self.fc2_weight, self.fc2_bias, self.eps, g,
                                            add_residual, self._wcache, first)
        else:
            self._apply_extra_state(recipe, inp.device)
            out = fn.LayerNormMLPFn.apply(inp, self.layer_norm_weight, self.fc1_weight, self.fc1_bias, self.fc2_weight, self.fc2_bias, self._meta,
                                          recipe, self.eps, g, add_residual, self._wcache, first)
        return out

# ------------------------------------------------------------------------------------------ attention shell
class RotaryPositionEmbedding(nn.Module):
    """RoPE angle table; call with max_seq_len -> freqs [S, 1, 1, dim] (te_llama.py:65-66)."""

    def __init__(self, dim: int, rotary_percent: float = 1.0, seq_len_interpolation_factor=None, pretrained_max_position_embeddings=None,
                 rotary_base: float = 10000.0, interleaved: bool = False):
        super().__init__()
        if rotary_percent < 1.0:
            dim = int(dim * rotary_percent)
        self.dim = dim
        inv_freq = 1.0 / (rotary_base ** (torch.arange(0, dim, 2, dtype=torch.float32) / dim))
        self.register_buffer("inv_freq", inv_freq, persistent=False)

    def forward(self, max_seq_len: int, offset: int = 0) -> torch.Tensor:
        seq = torch.arange(max_seq_len, device=self.inv_freq.device, dtype=torch.float32) + offset
        freqs = torch.outer(seq, self.inv_freq)
        emb = torch.cat((freqs, freqs), dim=-1)
        return emb.reshape(emb.size(0), 1, 1, emb.size(1))


def apply_rotary_pos_emb(t: torch.Tensor, freqs: torch.Tensor) -> torch.Tensor:
    """t [B, S, h, d] (bshd), freqs [S_max, 1, 1, d]: non-interleaved rotate-half."""
    S = t.shape[1]
    f = freqs[:S, 0, 0, :].to(torch.float32)
    cos, sin = f.cos()[None, :, None, :], f.sin()[None, :, None, :]
    tf = t.float()
    d = tf.shape[-1]
    rot = torch.cat((-tf[..., d // 2:], tf[..., : d // 2]), dim=-1)
    return (tf * cos + rot * sin).to(t.dtype)


class MultiheadAttention(nn.Module):
    """Attention block shell: our LayerNormLinear (qkv) and Linear (proj) around a bf16 attention core
    (torch SDPA — the reference's core is bf16 flash-attention, NVTE_FLASH_ATTN=1 README.md:27; out of the FP8 path)."""

    def __init__(self, hidden_size: int, num_attention_heads: int, kv_channels: Optional[int] = None, attention_dropout: float = 0.1,
                 layernorm_epsilon: float = 1e-5, init_method=None, output_layer_init_method=None, layer_number=None,
                 attn_mask_type: str = "causal", num_gqa_groups: Optional[int] = None, return_layernorm_output: bool = False,
                 input_layernorm: bool = False, attention_type: str = "self", fuse_qkv_params: bool = False, bias: bool = True,
                 normalization: str = "LayerNorm", qkv_format: str = "sbhd", params_dtype=None, device=None, **_unused):
        super().__init__()
        if attention_type != "self" or not input_layernorm:
            raise NotImplementedError("built for self-attention with input_layernorm=True (te_llama.py:45-56)")
        self.hidden_size, self.num_heads = hidden_size, num_attention_heads
        self.num_gqa_groups = num_gqa_groups or num_attention_heads
        self.head_dim = kv_channels or hidden_size // num_attention_heads
        self.qkv_format, self.attn_mask_type, self.dropout = qkv_format, attn_mask_type, attention_dropout
        q_out = self.num_heads * self.head_dim
        kv_out = self.num_gqa_groups * self.head_dim
        split = {"query": q_out, "key": kv_out, "value": kv_out} if not fuse_qkv_params else None
        self.layernorm_qkv = LayerNormLinear(hidden_size, q_out + 2 * kv_out, eps=layernorm_epsilon, bias=bias, normalization=normalization,
                                             parameters_split=split, params_dtype=params_dtype, device=device, init_method=init_method)
        self.proj = Linear(q_out, hidden_size, bias=bias, params_dtype=params_dtype, device=device, init_method=output_layer_init_method)

    def forward(self, hidden_states: torch.Tensor, attention_mask: Optional[torch.Tensor] = None, rotary_pos_emb=None,
                is_first_microbatch=None, **_unused):
        if self.qkv_format == "sbhd":
            hidden_states = hidden_states.transpose(0, 1)
        B, S, _ = hidden_states.shape
        qkv = self.layernorm_qkv(hidden_states, is_first_microbatch=is_first_microbatch)
        q_out, kv_out = self.num_heads * self.head_dim, self.num_gqa_groups * self.head_dim
        q, k, v = qkv.split([q_out, kv_out, kv_out], dim=-1)
        q = q.reshape(B, S, self.num_heads, self.head_dim)
        k = k.reshape(B, S, self.num_gqa_groups, self.head_dim)
        v = v.reshape(B, S, self.num_gqa_groups, self.head_dim)
        if rotary_pos_emb is not None:
            if isinstance(rotary_pos_emb, (tuple, list)):
                qf, kf = rotary_pos_emb
            else:
                qf = kf = rotary_pos_emb
            q, k = apply_rotary_pos_emb(q, qf), apply_rotary_pos_emb(k, kf)
        causal = "causal" in self.attn_mask_type
        mask = None
        if attention_mask is not None and not causal:
            mask = ~attention_mask if attention_mask.dtype == torch.bool else attention_mask     # TE: True = masked out
        ctx = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), attn_mask=mask,
                                             dropout_p=self.dropout if self.training else 0.0, is_causal=causal and mask is None,
                                             enable_gqa=self.num_gqa_groups != self.num_heads)
        ctx = ctx.transpose(1, 2).reshape(B, S, q_out)
        out = self.proj(ctx, is_first_microbatch=is_first_microbatch)
        if self.qkv_format == "sbhd":
            out = out.transpose(0, 1)
        return out


class TransformerLayer(nn.Module):
    """Name only: the reference evaluates `te.pytorch.TransformerLayer` as a bare expression (te_llama.py:44) and never builds one."""

    def __init__(self, *a, **k):
        raise NotImplementedError("TransformerLayer is not part of the FP8 linear path; build MultiheadAttention + LayerNormMLP")
