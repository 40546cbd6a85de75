"""PyTorch-CPU oracle for the FP8 / MXFP8 linear-layer hot path.

TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may import this module; the product package (llm_fp8_b200) never does.

PARITY UNPINNED at the Transformer-Engine boundary: the reference repo contains no arithmetic for
this path (it calls `transformer_engine`, third-party, un-vendored: install.sh:4; "2.5.0" per
paper/conference_101719.tex:236) and no tests.  This file restates the rules BASELINE.json's north
star fixes (saturating RNE casts to torch.float8_e4m3fn / e5m2, amax/scale rules, E8M0 round-up,
GEMM = dequantise -> fp32 matmul) on top of the reference's call sites:

  te_llama.py:39-40        attn = DelayedScaling(HYBRID, hist 16, "max"); mlp = DelayedScaling(E4M3, ...)
  te_llama_hybrid.py:39    one DelayedScaling(HYBRID, 16, "max")
  te_llama_mxfp8.py:28-29  MXFP8BlockScaling(fp8_format=Format.E4M3)
  te_llama.py:45-63        MultiheadAttention(RMSNorm, input_layernorm) + LayerNormMLP(RMSNorm, swiglu)
  te_llama.py:225-233      fc1 rows = gate then up

It is pinned on (i) the three valid FP8 encodings of thesis/chapters/c2/c2_quantization.tex:100,102,103,
(ii) agreement between two independent restatements — torch's float8 casts here and the integer
bit-level C code in oracle_c.c — over every bf16 input value (tests/test_oracle.py).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch

E4M3 = torch.float8_e4m3fn
E5M2 = torch.float8_e5m2
FMAX = {E4M3: 448.0, E5M2: 57344.0}
MX_BLOCK = 32


# --------------------------------------------------------------------------- casts
def sat_cast(v: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """Saturating RNE cast: torch's CPU cast is RNE but not saturating."""
    fmax = FMAX[dtype]
    return v.float().clamp(-fmax, fmax).to(dtype)


def amax_of(x: torch.Tensor) -> torch.Tensor:
    """max |x| in fp32, NaN ignored (fmaxf semantics, what the kernels compute)."""
    a = x.float().abs()
    a = torch.where(torch.isnan(a), torch.zeros_like(a), a)
    return a.max() if a.numel() else torch.zeros((), dtype=torch.float32)


def quantize(x: torch.Tensor, scale: float | torch.Tensor, dtype: torch.dtype):
    """K2: q = sat_cast(x * scale) with fp32 multiply; returns (q, amax)."""
    s = torch.as_tensor(scale, dtype=torch.float32)
    return sat_cast(x.float() * s, dtype), amax_of(x)


def dequantize(q: torch.Tensor, scale_inv: float | torch.Tensor) -> torch.Tensor:
    return q.float() * torch.as_tensor(scale_inv, dtype=torch.float32)


# --------------------------------------------------------------------------- delayed scaling (K4)
@dataclass
class DelayedScalingState:
    """amax_history[L, T], scale[T], scale_inv[T]; SURVEY.md Appendix A update rule."""
    num_tensors: int
    history_len: int = 16
    margin: int = 0
    fmax: torch.Tensor = None  # [T]
    amax_history: torch.Tensor = field(init=False)
    scale: torch.Tensor = field(init=False)
    scale_inv: torch.Tensor = field(init=False)

    def __post_init__(self):
        T = self.num_tensors
        self.amax_history = torch.zeros(self.history_len, T, dtype=torch.float32)
        self.scale = torch.ones(T, dtype=torch.float32)
        self.scale_inv = torch.ones(T, dtype=torch.float32)
        if self.fmax is None:
            self.fmax = torch.full((T,), 448.0, dtype=torch.float32)

    def record(self, slot: int, amax: float | torch.Tensor):
        self.amax_history[0, slot] = torch.maximum(self.amax_history[0, slot], torch.as_tensor(amax, dtype=torch.float32))

    def update(self):
        amax_scale_update(self.amax_history, self.scale, self.scale_inv, self.fmax, self.margin)


def amax_scale_update(hist, scale, scale_inv, fmax, margin=0):
    """In-place: amax = max(hist, 0); sf = fmax/amax/2^margin (guards); roll history; zero slot 0."""
    finite = torch.isfinite(hist).all(dim=0)
    h = torch.where(torch.isnan(hist), torch.zeros_like(hist), hist)
    amax = h.max(dim=0).values
    sf = (fmax / amax) / (2.0 ** margin)
    sf = torch.where(amax > 0, sf, scale)
    sf = torch.where(finite, sf, scale)
    sf = torch.where(torch.isinf(sf), torch.full_like(sf, torch.finfo(torch.float32).max), sf)
    scale.copy_(sf)
    scale_inv.copy_(1.0 / sf)
    hist.copy_(torch.roll(hist, -1, 0))
    hist[0].zero_()


# --------------------------------------------------------------------------- E8M0 / MXFP8 (K2m)
def e8m0_roundup(v: torch.Tensor) -> torch.Tensor:
    """E8M0 biased exponent of v (= amax/448) rounded UP to a power of two, satfinite (Appendix C)."""
    bits = v.float().contiguous().view(torch.int32).to(torch.int64) & 0x7FFFFFFF
    exp = bits >> 23
    man = bits & 0x7FFFFF
    bump = (man > 0) & (exp != 0xFE) & ~((exp == 0) & (man <= 0x400000))
    e = torch.where(bump, exp + 1, exp)
    e = torch.where(bits == 0, torch.zeros_like(e), e)
    e = torch.where(bits > 0x7F800000, torch.full_like(e, 0xFF), e)
    return e.to(torch.uint8)


def e8m0_inv_scale(e: torch.Tensor) -> torch.Tensor:
    """2^(127-e) as fp32 (denormal for e >= 254)."""
    e64 = e.to(torch.int64)
    bits = torch.where(e64 <= 253, (254 - e64) << 23,
                       torch.where(e64 == 254, torch.full_like(e64, 0x00400000), torch.full_like(e64, 0x00200000)))
    return bits.to(torch.int32).view(torch.float32)


def e8m0_to_f32(e: torch.Tensor) -> torch.Tensor:
    """2^(e-127) (dequant scale)."""
    return torch.pow(torch.tensor(2.0, dtype=torch.float64), e.to(torch.float64) - 127.0).float()


def mx_quantize_rowwise(x: torch.Tensor):
    """x [R, C] (C % 32 == 0): blocks of 32 along the last dim.  Returns (q e4m3 [R,C], e8m0 u8 [R, C/32])."""
    xf = x.float()
    R, C = xf.shape
    blk = xf.reshape(R, C // MX_BLOCK, MX_BLOCK)
    a = blk.abs()
    a = torch.where(torch.isnan(a), torch.zeros_like(a), a)
    am = a.amax(dim=-1)
    e = e8m0_roundup(am / 448.0)
    inv = e8m0_inv_scale(e)
    q = sat_cast(blk * inv.unsqueeze(-1), E4M3).reshape(R, C)
    return q, e


def mx_quantize_colwise(x: torch.Tensor):
    """Blocks of 32 along dim 0.  Returns (q e4m3 [R,C] same layout as x, e8m0 u8 [C, R/32])."""
    qt, e = mx_quantize_rowwise(x.float().t().contiguous())
    return qt.t().contiguous(), e


def mx_dequantize_rowwise(q: torch.Tensor, e: torch.Tensor) -> torch.Tensor:
    R, C = q.shape
    return (q.float().reshape(R, C // MX_BLOCK, MX_BLOCK) * e8m0_to_f32(e).unsqueeze(-1)).reshape(R, C)


def mx_dequantize_colwise(q: torch.Tensor, e: torch.Tensor) -> torch.Tensor:
    return mx_dequantize_rowwise(q.t().contiguous(), e).t().contiguous()


def sf_swizzled_size(rows: int, kblocks: int) -> int:
    return (math.ceil(rows / 128) * 128) * (math.ceil(kblocks / 4) * 4)


def swizzle_sf(e: torch.Tensor) -> torch.Tensor:
    """[rows, kblocks] u8 -> flat u8 in the 128x4-tile layout tcgen05.cp / cuBLASLt consume:
    tile (r/128, kb/4) row-major, 512 B per tile, in-tile offset (r%32)*16 + ((r%128)//32)*4 + kb%4."""
    rows, kb = e.shape
    rp, kp = math.ceil(rows / 128) * 128, math.ceil(kb / 4) * 4
    pad = torch.zeros(rp, kp, dtype=torch.uint8)
    pad[:rows, :kb] = e
    t = pad.reshape(rp // 128, 4, 32, kp // 4, 4)      # [rt, r/32%4, r%32, kt, kb%4]
    return t.permute(0, 3, 2, 1, 4).contiguous().reshape(-1)


def unswizzle_sf(flat: torch.Tensor, rows: int, kb: int) -> torch.Tensor:
    rp, kp = math.ceil(rows / 128) * 128, math.ceil(kb / 4) * 4
    t = flat.reshape(rp // 128, kp // 4, 32, 4, 4).permute(0, 3, 2, 1, 4).contiguous().reshape(rp, kp)
    return t[:rows, :kb].contiguous()


# --------------------------------------------------------------------------- RMSNorm / SwiGLU (float refs)
def rmsnorm_ref(x: torch.Tensor, gamma: torch.Tensor, eps: float):
    """fp32 reference (reduction order free): used for tolerance checks; the bit-pinned order lives
    in oracle_c.c:orc_rms_rsigma."""
    xf = x.float()
    rs = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return xf * rs * gamma.float(), rs.squeeze(-1)


def swiglu_ref(fc1_out: torch.Tensor) -> torch.Tensor:
    g, u = fc1_out.float().chunk(2, dim=-1)
    return torch.nn.functional.silu(g) * u


# --------------------------------------------------------------------------- GEMM reference
def gemm_ref(aq: torch.Tensor, a_scale_inv, bq: torch.Tensor, b_scale_inv, bias=None) -> torch.Tensor:
    """Y[M,N] = dequant(Aq)[M,K] @ dequant(Bq)[N,K]^T (+bias), fp32 accumulate."""
    y = dequantize(aq, a_scale_inv) @ dequantize(bq, b_scale_inv).t()
    if bias is not None:
        y = y + bias.float()
    return y


# --------------------------------------------------------------------------- fake-quant linear path (CPU baseline)
class _FakeQuant(torch.autograd.Function):
    """Quantise->dequantise with a straight-through gradient that is itself fake-quantised to `bwd`."""

    @staticmethod
    def forward(ctx, x, fwd_dtype, bwd_dtype, mx):
        ctx.bwd_dtype, ctx.mx = bwd_dtype, mx
        return fake_quant(x, fwd_dtype, mx)

    @staticmethod
    def backward(ctx, g):
        return fake_quant(g, ctx.bwd_dtype, ctx.mx), None, None, None


def fake_quant(x: torch.Tensor, dtype: torch.dtype, mx: bool) -> torch.Tensor:
    """Current-scaling fake quantisation used by the CPU baseline (just-in-time amax; the delayed-scaling
    history only changes *which* scale is used, not the cost)."""
    shp = x.shape
    x2 = x.reshape(-1, shp[-1])
    if mx:
        q, e = mx_quantize_rowwise(x2)
        out = mx_dequantize_rowwise(q, e)
    else:
        am = amax_of(x2)
        scale = torch.where(am > 0, FMAX[dtype] / am, torch.ones_like(am))
        q, _ = quantize(x2, scale, dtype)
        out = dequantize(q, 1.0 / scale)
    return out.reshape(shp).to(x.dtype)


def fq_linear(x, w, b, fwd_dtype, bwd_dtype, mx):
    xq = _FakeQuant.apply(x, fwd_dtype, bwd_dtype, mx)
    wq = _FakeQuant.apply(w, fwd_dtype, bwd_dtype, mx)
    y = torch.nn.functional.linear(xq.float(), wq.float(), None if b is None else b.float())
    return y


def fq_layer_linear_path(x, p, eps_attn, eps_mlp, fwd_dtype=E4M3, bwd_dtype=E5M2, mx=False):
    """The linear-layer path of one decoder layer, attention core replaced by identity on V-sized
    slices (the attention core is bf16 flash-attention in the reference and out of scope):
      LayerNormLinear(qkv) -> [take first H columns] -> Linear(proj) ; LayerNormMLP (fc1, swiglu, fc2).
    p: dict with ln1, wqkv, wproj, ln2, w1, b1, w2, b2 (fp32 leaf tensors)."""
    H = x.shape[-1]
    y1, _ = rmsnorm_ref(x, p["ln1"], eps_attn)
    qkv = fq_linear(y1, p["wqkv"], None, fwd_dtype, bwd_dtype, mx)
    attn = qkv[..., :H]
    h = x.float() + fq_linear(attn, p["wproj"], None, fwd_dtype, bwd_dtype, mx)
    y2, _ = rmsnorm_ref(h, p["ln2"], eps_mlp)
    f1 = fq_linear(y2, p["w1"], p["b1"], fwd_dtype, bwd_dtype, mx)
    a = swiglu_ref(f1)
    out = h + fq_linear(a, p["w2"], p["b2"], fwd_dtype, bwd_dtype, mx)
    return out
