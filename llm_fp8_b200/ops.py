"""Tensor-level wrappers over the C ABI (include/llmfp8_b200.h).  PyTorch supplies device memory and the current
stream; every byte of arithmetic happens in libllmfp8_b200.so."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from . import _lib
from .fp8 import ScalePool
from .recipe import FMT_E4M3, FMT_E5M2

TORCH_FP8 = {FMT_E4M3: torch.float8_e4m3fn, FMT_E5M2: torch.float8_e5m2}


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _req(t: torch.Tensor, dtype=None):
    if not t.is_cuda:
        raise RuntimeError("llm_fp8_b200 operates on CUDA tensors only (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("llm_fp8_b200 needs contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected {dtype}, got {t.dtype}")
    return t


def _req_rows(t: torch.Tensor, dtype=None):
    """2D tensor whose rows are contiguous but may be spaced (a column slice of a wider matrix): -> (tensor, row pitch in elements).
    Anything else is made contiguous by the caller."""
    if not t.is_cuda:
        raise RuntimeError("llm_fp8_b200 operates on CUDA tensors only (no CPU fallback)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected {dtype}, got {t.dtype}")
    if t.dim() != 2 or t.stride(1) != 1 or t.stride(0) < t.shape[1] or t.stride(0) % 8 or t.data_ptr() % 16:
        raise RuntimeError("llm_fp8_b200 needs a 2D tensor with contiguous, 16 B aligned rows")
    return t, t.stride(0)


def _req_out(t: torch.Tensor):
    """GEMM output: contiguous, or a column slice of a wider row-major matrix (the TMA store descriptor carries the row pitch)."""
    if t.dim() == 2 and t.is_cuda and not t.is_contiguous():
        return _req_rows(t)[0]
    return _req(t)


def rows_ok(t: torch.Tensor) -> bool:
    """True when `t` ([R, C]) can be read in place by the strided quantisers (else the caller makes it contiguous)."""
    return (t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1] and t.stride(0) % 8 == 0 and t.data_ptr() % 16 == 0
            and t.shape[1] % 16 == 0)


@dataclass
class QTensor:
    """Per-tensor-scaled fp8 matrix [rows, cols] (bytes) + where its dequant scale lives."""
    data: torch.Tensor            # uint8 [R, C]
    fmt: int
    scale_inv: torch.Tensor       # fp32 tensor holding the scale_inv (pool storage or a private 1-element tensor)
    scale_inv_off: int = 0        # element offset into scale_inv

    @property
    def scale_inv_ptr(self) -> int:
        return self.scale_inv.data_ptr() + 4 * self.scale_inv_off

    def dequantize(self) -> torch.Tensor:
        return self.data.view(TORCH_FP8[self.fmt]).float() * self.scale_inv[self.scale_inv_off]


@dataclass
class MXTensor:
    """MXFP8 matrix [rows, cols]: rowwise copy (blocks along cols) and/or colwise copy (blocks along rows)."""
    rows: int
    cols: int
    row: Optional[torch.Tensor] = None      # uint8 [R, C]
    sf_row: Optional[torch.Tensor] = None   # uint8 swizzled, scale matrix (r, c/32)
    col: Optional[torch.Tensor] = None      # uint8 [R, C]
    sf_col: Optional[torch.Tensor] = None   # uint8 swizzled, scale matrix (c, r/32)
    src: Optional[tuple] = None             # (data_ptr, version) of the bf16 tensor these copies were made from, when its producer vouches for it


def mx_scale_bytes(rows: int, k: int) -> int:
    return ((rows + 127) // 128 * 128) * (((k + 31) // 32 + 3) // 4 * 4)


# ------------------------------------------------------------------------------------------ per-tensor quantise
def quantize(x: torch.Tensor, fmt: int, pool: ScalePool, slot: int) -> QTensor:
    """K2 with the pool's scale; amax lands in the pool's history row 0.  A 2D x may be a column slice (spaced rows)."""
    if x.dim() == 2 and not x.is_contiguous():
        x, ldx = _req_rows(x, torch.bfloat16)
        q = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
        rc = _lib.load().llmfp8_quant_cast_amax_2d(x.data_ptr(), ldx, q.data_ptr(), x.shape[0], x.shape[1], fmt, pool.scale_ptr(slot), pool.amax_ptr(slot),
                                                   _stream(x))
        _lib.check(rc, "quant_cast_amax_2d")
        return QTensor(q, fmt, pool.scale_inv, slot)
    x = _req(x, torch.bfloat16)
    q = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    rc = _lib.load().llmfp8_quant_cast_amax(x.data_ptr(), q.data_ptr(), x.numel(), fmt, pool.scale_ptr(slot), pool.amax_ptr(slot), _stream(x))
    _lib.check(rc, "quant_cast_amax")
    return QTensor(q, fmt, pool.scale_inv, slot)


def quantize_raw(x: torch.Tensor, fmt: int, scale: Optional[torch.Tensor], amax: Optional[torch.Tensor]) -> torch.Tensor:
    """K2 with explicit device scalars (tests / benches)."""
    if x.dim() == 2 and x.is_cuda and not x.is_contiguous():
        x, ldx = _req_rows(x, torch.bfloat16)
        q = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
        rc = _lib.load().llmfp8_quant_cast_amax_2d(x.data_ptr(), ldx, q.data_ptr(), x.shape[0], x.shape[1], fmt, scale.data_ptr() if scale is not None else None,
                                                   amax.data_ptr() if amax is not None else None, _stream(x))
        _lib.check(rc, "quant_cast_amax_2d")
        return q
    x = _req(x, torch.bfloat16)
    q = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    rc = _lib.load().llmfp8_quant_cast_amax(x.data_ptr(), q.data_ptr(), x.numel(), fmt, scale.data_ptr() if scale is not None else None,
                                            amax.data_ptr() if amax is not None else None, _stream(x))
    _lib.check(rc, "quant_cast_amax")
    return q


def quantize_bgrad_raw(x: torch.Tensor, fmt: int, scale: Optional[torch.Tensor], amax: Optional[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
    x = _req(x, torch.bfloat16)
    R, C = x.shape
    q = torch.empty((R, C), dtype=torch.uint8, device=x.device)
    ws = torch.empty(C, dtype=torch.float32, device=x.device)
    bgrad = torch.empty(C, dtype=torch.bfloat16, device=x.device)
    rc = _lib.load().llmfp8_quant_cast_amax_bgrad(x.data_ptr(), q.data_ptr(), R, C, fmt, scale.data_ptr() if scale is not None else None,
                                                  amax.data_ptr() if amax is not None else None, ws.data_ptr(), bgrad.data_ptr(), _stream(x))
    _lib.check(rc, "quant_cast_amax_bgrad")
    return q, bgrad


def quantize_bgrad(x: torch.Tensor, fmt: int, pool: ScalePool, slot: int) -> Tuple[QTensor, torch.Tensor]:
    x = _req(x, torch.bfloat16)
    R, C = x.shape
    q = torch.empty((R, C), dtype=torch.uint8, device=x.device)
    ws = torch.empty(C, dtype=torch.float32, device=x.device)
    bgrad = torch.empty(C, dtype=torch.bfloat16, device=x.device)
    rc = _lib.load().llmfp8_quant_cast_amax_bgrad(x.data_ptr(), q.data_ptr(), R, C, fmt, pool.scale_ptr(slot), pool.amax_ptr(slot), ws.data_ptr(),
                                                  bgrad.data_ptr(), _stream(x))
    _lib.check(rc, "quant_cast_amax_bgrad")
    return QTensor(q, fmt, pool.scale_inv, slot), bgrad


def rmsnorm_quant_raw(x, gamma, eps, fmt, scale, amax):
    x, gamma = _req(x, torch.bfloat16), _req(gamma, torch.bfloat16)
    M, H = x.shape
    q = torch.empty((M, H), dtype=torch.uint8, device=x.device)
    rs = torch.empty(M, dtype=torch.float32, device=x.device)
    rc = _lib.load().llmfp8_rmsnorm_quant_fwd(x.data_ptr(), gamma.data_ptr(), q.data_ptr(), rs.data_ptr(), M, H, eps, fmt,
                                              scale.data_ptr() if scale is not None else None, amax.data_ptr() if amax is not None else None, _stream(x))
    _lib.check(rc, "rmsnorm_quant_fwd")
    return q, rs


def rmsnorm_quant(x, gamma, eps, fmt, pool: ScalePool, slot: int) -> Tuple[QTensor, torch.Tensor]:
    x, gamma = _req(x, torch.bfloat16), _req(gamma, torch.bfloat16)
    M, H = x.shape
    q = torch.empty((M, H), dtype=torch.uint8, device=x.device)
    rs = torch.empty(M, dtype=torch.float32, device=x.device)
    rc = _lib.load().llmfp8_rmsnorm_quant_fwd(x.data_ptr(), gamma.data_ptr(), q.data_ptr(), rs.data_ptr(), M, H, eps, fmt, pool.scale_ptr(slot),
                                              pool.amax_ptr(slot), _stream(x))
    _lib.check(rc, "rmsnorm_quant_fwd")
    return QTensor(q, fmt, pool.scale_inv, slot), rs


def rmsnorm_rsigma(x, eps) -> torch.Tensor:
    x = _req(x, torch.bfloat16)
    M, H = x.shape
    rs = torch.empty(M, dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().llmfp8_rmsnorm_rsigma(x.data_ptr(), rs.data_ptr(), M, H, eps, _stream(x)), "rmsnorm_rsigma")
    return rs


def rmsnorm_bwd(dy, x, gamma, rsigma, add: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """K5.  `add` (bf16 [M, H]): the gradient of the residual branch around the norm, folded into dx in the same pass."""
    dy, x, gamma = _req(dy, torch.bfloat16), _req(x, torch.bfloat16), _req(gamma, torch.bfloat16)
    M, H = x.shape
    lib = _lib.load()
    dx = torch.empty_like(x)
    dgamma = torch.empty_like(gamma)
    ws = torch.empty(lib.llmfp8_rmsnorm_bwd_ws_rows(M) * H, dtype=torch.float32, device=x.device)
    if add is not None:
        add = _req(add, torch.bfloat16)
        if add.numel() != x.numel():
            raise RuntimeError("rmsnorm_bwd: residual gradient shape mismatch")
        rc = lib.llmfp8_rmsnorm_bwd_add(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), rsigma.data_ptr(), add.data_ptr(), dx.data_ptr(), dgamma.data_ptr(),
                                        ws.data_ptr(), M, H, _stream(x))
    else:
        rc = lib.llmfp8_rmsnorm_bwd(dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), rsigma.data_ptr(), dx.data_ptr(), dgamma.data_ptr(), ws.data_ptr(), M, H,
                                    _stream(x))
    _lib.check(rc, "rmsnorm_bwd")
    return dx, dgamma


def swiglu_quant_raw(fc1_out, fmt, scale, amax) -> torch.Tensor:
    fc1_out = _req(fc1_out, torch.bfloat16)
    M, I2 = fc1_out.shape
    q = torch.empty((M, I2 // 2), dtype=torch.uint8, device=fc1_out.device)
    rc = _lib.load().llmfp8_swiglu_quant_fwd(fc1_out.data_ptr(), q.data_ptr(), M, I2 // 2, fmt, scale.data_ptr() if scale is not None else None,
                                             amax.data_ptr() if amax is not None else None, _stream(fc1_out))
    _lib.check(rc, "swiglu_quant_fwd")
    return q


def swiglu_quant(fc1_out, fmt, pool: ScalePool, slot: int) -> QTensor:
    fc1_out = _req(fc1_out, torch.bfloat16)
    M, I2 = fc1_out.shape
    q = torch.empty((M, I2 // 2), dtype=torch.uint8, device=fc1_out.device)
    rc = _lib.load().llmfp8_swiglu_quant_fwd(fc1_out.data_ptr(), q.data_ptr(), M, I2 // 2, fmt, pool.scale_ptr(slot), pool.amax_ptr(slot), _stream(fc1_out))
    _lib.check(rc, "swiglu_quant_fwd")
    return QTensor(q, fmt, pool.scale_inv, slot)


def dswiglu_quant_raw(da, fc1_out, fmt, scale, amax, want_dbias=True):
    da, fc1_out = _req(da, torch.bfloat16), _req(fc1_out, torch.bfloat16)
    M, I = da.shape
    q = torch.empty((M, 2 * I), dtype=torch.uint8, device=da.device)
    ws = torch.empty(2 * I, dtype=torch.float32, device=da.device) if want_dbias else None
    dbias = torch.empty(2 * I, dtype=torch.bfloat16, device=da.device) if want_dbias else None
    rc = _lib.load().llmfp8_dswiglu_quant_bwd(da.data_ptr(), fc1_out.data_ptr(), q.data_ptr(), M, I, fmt, scale.data_ptr() if scale is not None else None,
                                              amax.data_ptr() if amax is not None else None, ws.data_ptr() if want_dbias else None,
                                              dbias.data_ptr() if want_dbias else None, _stream(da))
    _lib.check(rc, "dswiglu_quant_bwd")
    return q, dbias


def dswiglu_quant(da, fc1_out, fmt, pool: ScalePool, slot: int, want_dbias=True) -> Tuple[QTensor, Optional[torch.Tensor]]:
    da, fc1_out = _req(da, torch.bfloat16), _req(fc1_out, torch.bfloat16)
    M, I = da.shape
    q = torch.empty((M, 2 * I), dtype=torch.uint8, device=da.device)
    ws = torch.empty(2 * I, dtype=torch.float32, device=da.device) if want_dbias else None
    dbias = torch.empty(2 * I, dtype=torch.bfloat16, device=da.device) if want_dbias else None
    rc = _lib.load().llmfp8_dswiglu_quant_bwd(da.data_ptr(), fc1_out.data_ptr(), q.data_ptr(), M, I, fmt, pool.scale_ptr(slot), pool.amax_ptr(slot),
                                              ws.data_ptr() if want_dbias else None, dbias.data_ptr() if want_dbias else None, _stream(da))
    _lib.check(rc, "dswiglu_quant_bwd")
    return QTensor(q, fmt, pool.scale_inv, slot), dbias


# ------------------------------------------------------------------------------------------ MXFP8 quantise
def mx_quantize(x: torch.Tensor, rowwise: bool = True, colwise: bool = True, rsigma: Optional[torch.Tensor] = None,
                gamma: Optional[torch.Tensor] = None) -> MXTensor:
    if x.dim() == 2 and x.is_cuda and not x.is_contiguous():
        x, ldx = _req_rows(x, torch.bfloat16)          # a column slice: the TMA descriptor carries the row pitch
    else:
        x = _req(x, torch.bfloat16)
        ldx = x.shape[1]
    R, C = x.shape
    dev = x.device
    out = MXTensor(R, C)
    if rowwise:
        out.row = torch.empty((R, C), dtype=torch.uint8, device=dev)
        out.sf_row = torch.empty(mx_scale_bytes(R, C), dtype=torch.uint8, device=dev)
    if colwise:
        out.col = torch.empty((R, C), dtype=torch.uint8, device=dev)
        out.sf_col = torch.empty(mx_scale_bytes(C, R), dtype=torch.uint8, device=dev)
    rc = _lib.load().llmfp8_mx_quantize_ld(x.data_ptr(), ldx, R, C, out.row.data_ptr() if rowwise else None, out.sf_row.data_ptr() if rowwise else None,
                                        out.col.data_ptr() if colwise else None, out.sf_col.data_ptr() if colwise else None,
                                        rsigma.data_ptr() if rsigma is not None else None, gamma.data_ptr() if gamma is not None else None, _stream(x))
    _lib.check(rc, "mx_quantize")
    return out


def mx_quantize_bgrad(x: torch.Tensor, rowwise: bool = True):
    """mx_quantize(x, rowwise, colwise=True) of a gradient dY [R, C] plus its column sums (the bias gradient), in one pass."""
    x = _req(x, torch.bfloat16)
    R, C = x.shape
    out = _mx_out(R, C, rowwise, True, x.device)
    ws = torch.empty(C, dtype=torch.float32, device=x.device)
    bgrad = torch.empty(C, dtype=torch.bfloat16, device=x.device)
    rc = _lib.load().llmfp8_mx_quantize_bgrad(x.data_ptr(), C, R, C, _p(out.row), _p(out.sf_row), out.col.data_ptr(), out.sf_col.data_ptr(), ws.data_ptr(),
                                              bgrad.data_ptr(), _stream(x))
    _lib.check(rc, "mx_quantize_bgrad")
    return out, bgrad


# ------------------------------------------------------------------------------------------ GEMMs
def fp8_gemm(a: QTensor, a_mn: bool, b: QTensor, b_mn: bool, M: int, N: int, K: int, bias: Optional[torch.Tensor] = None,
             out: Optional[torch.Tensor] = None, accumulate: bool = False, out_dtype: torch.dtype = torch.bfloat16, max_ctas: int = 0,
             residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """D[M,N] = A[M,K] B[N,K]^T * (a.scale_inv * b.scale_inv) (+bias) (+D) (+residual[M,N], bf16, fused into the epilogue).
    a_mn / b_mn: operand stored [K, M] / [K, N]."""
    dev = a.data.device
    if out is None:
        if accumulate:
            raise RuntimeError("accumulate=True needs an output tensor")
        out = torch.empty((M, N), dtype=out_dtype, device=dev)
    _req_out(out)
    if out.dtype not in (torch.bfloat16, torch.float32):
        raise RuntimeError("fp8_gemm output must be bf16 or fp32")
    lda, ldb = a.data.shape[-1], b.data.shape[-1]
    if residual is not None:
        residual, ldr = _req_rows(residual, torch.bfloat16)
        if tuple(residual.shape) != (M, N) or out.dtype != torch.bfloat16 or accumulate:
            raise RuntimeError("fp8_gemm: the fused residual needs a bf16 [M, N] residual, a bf16 output and no accumulate")
        rc = _lib.load().llmfp8_fp8_gemm_res(a.data.data_ptr(), lda, a.fmt, int(a_mn), b.data.data_ptr(), ldb, b.fmt, int(b_mn), out.data_ptr(), out.stride(0),
                                             M, N, K, a.scale_inv_ptr, b.scale_inv_ptr, bias.data_ptr() if bias is not None else None,
                                             residual.data_ptr(), ldr, max_ctas, _stream(out))
        _lib.check(rc, "fp8_gemm_res")
        return out
    rc = _lib.load().llmfp8_fp8_gemm(a.data.data_ptr(), lda, a.fmt, int(a_mn), b.data.data_ptr(), ldb, b.fmt, int(b_mn), out.data_ptr(), out.stride(0),
                                     int(out.dtype == torch.float32), M, N, K, a.scale_inv_ptr, b.scale_inv_ptr,
                                     bias.data_ptr() if bias is not None else None, int(accumulate), max_ctas, _stream(out))
    _lib.check(rc, "fp8_gemm")
    return out


def mxfp8_gemm(a: torch.Tensor, sfa: torch.Tensor, a_mn: bool, b: torch.Tensor, sfb: torch.Tensor, b_mn: bool, M: int, N: int, K: int,
               bias: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, accumulate: bool = False,
               out_dtype: torch.dtype = torch.bfloat16, max_ctas: int = 0, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    dev = a.device
    if out is None:
        if accumulate:
            raise RuntimeError("accumulate=True needs an output tensor")
        out = torch.empty((M, N), dtype=out_dtype, device=dev)
    _req_out(out)
    if residual is not None:
        residual, ldr = _req_rows(residual, torch.bfloat16)
        if tuple(residual.shape) != (M, N) or out.dtype != torch.bfloat16 or accumulate:
            raise RuntimeError("mxfp8_gemm: the fused residual needs a bf16 [M, N] residual, a bf16 output and no accumulate")
        rc = _lib.load().llmfp8_mxfp8_gemm_res(a.data_ptr(), a.shape[-1], int(a_mn), sfa.data_ptr(), b.data_ptr(), b.shape[-1], int(b_mn), sfb.data_ptr(),
                                               out.data_ptr(), out.stride(0), M, N, K, bias.data_ptr() if bias is not None else None,
                                               residual.data_ptr(), ldr, max_ctas, _stream(out))
        _lib.check(rc, "mxfp8_gemm_res")
        return out
    rc = _lib.load().llmfp8_mxfp8_gemm(a.data_ptr(), a.shape[-1], int(a_mn), sfa.data_ptr(), b.data_ptr(), b.shape[-1], int(b_mn), sfb.data_ptr(),
                                       out.data_ptr(), out.stride(0), int(out.dtype == torch.float32), M, N, K,
                                       bias.data_ptr() if bias is not None else None, int(accumulate), max_ctas, _stream(out))
    _lib.check(rc, "mxfp8_gemm")
    return out


# ------------------------------------------------------------------------------------------ bf16-output activations (MXFP8 path)
FMT_BF16 = 2


def swiglu_bf16(fc1_out: torch.Tensor) -> torch.Tensor:
    fc1_out = _req(fc1_out, torch.bfloat16)
    M, I2 = fc1_out.shape
    a = torch.empty((M, I2 // 2), dtype=torch.bfloat16, device=fc1_out.device)
    rc = _lib.load().llmfp8_swiglu_quant_fwd(fc1_out.data_ptr(), a.data_ptr(), M, I2 // 2, FMT_BF16, None, None, _stream(fc1_out))
    _lib.check(rc, "swiglu_quant_fwd(bf16)")
    return a


def dswiglu_bf16(da: torch.Tensor, fc1_out: torch.Tensor, want_dbias: bool = True):
    da, fc1_out = _req(da, torch.bfloat16), _req(fc1_out, torch.bfloat16)
    M, I = da.shape
    d = torch.empty((M, 2 * I), dtype=torch.bfloat16, device=da.device)
    ws = torch.empty(2 * I, dtype=torch.float32, device=da.device) if want_dbias else None
    dbias = torch.empty(2 * I, dtype=torch.bfloat16, device=da.device) if want_dbias else None
    rc = _lib.load().llmfp8_dswiglu_quant_bwd(da.data_ptr(), fc1_out.data_ptr(), d.data_ptr(), M, I, FMT_BF16, None, None,
                                              ws.data_ptr() if want_dbias else None, dbias.data_ptr() if want_dbias else None, _stream(da))
    _lib.check(rc, "dswiglu_quant_bwd(bf16)")
    return d, dbias


# ------------------------------------------------------------------------------------------ activation + MXFP8 quantiser in one pass
# (csrc/mx_act_quantize.cu — not yet run on hardware; functional.py selects these only with LLMFP8_MX_FUSED_ACT=1)
def _mx_out(R: int, C: int, rowwise: bool, colwise: bool, dev) -> MXTensor:
    out = MXTensor(R, C)
    if rowwise:
        out.row = torch.empty((R, C), dtype=torch.uint8, device=dev)
        out.sf_row = torch.empty(mx_scale_bytes(R, C), dtype=torch.uint8, device=dev)
    if colwise:
        out.col = torch.empty((R, C), dtype=torch.uint8, device=dev)
        out.sf_col = torch.empty(mx_scale_bytes(C, R), dtype=torch.uint8, device=dev)
    return out


def _p(t: Optional[torch.Tensor]):
    return t.data_ptr() if t is not None else None


def mx_swiglu_quantize(fc1_out: torch.Tensor, rowwise: bool = True, colwise: bool = True) -> MXTensor:
    """== mx_quantize(swiglu_bf16(fc1_out), rowwise, colwise) without the bf16 activation ever reaching HBM."""
    fc1_out = _req(fc1_out, torch.bfloat16)
    M, I2 = fc1_out.shape
    out = _mx_out(M, I2 // 2, rowwise, colwise, fc1_out.device)
    rc = _lib.load().llmfp8_mx_swiglu_quantize(fc1_out.data_ptr(), M, I2 // 2, _p(out.row), _p(out.sf_row), _p(out.col), _p(out.sf_col), _stream(fc1_out))
    _lib.check(rc, "mx_swiglu_quantize")
    return out


def mx_dswiglu_quantize(da: torch.Tensor, fc1_out: torch.Tensor, rowwise: bool = True, colwise: bool = True, want_dbias: bool = True):
    """== (mx_quantize(d, rowwise, colwise), dbias) for d, dbias = dswiglu_bf16(da, fc1_out, want_dbias), without d reaching HBM."""
    da, fc1_out = _req(da, torch.bfloat16), _req(fc1_out, torch.bfloat16)
    M, I = da.shape
    if tuple(fc1_out.shape) != (M, 2 * I):
        raise RuntimeError(f"mx_dswiglu_quantize: fc1_out {tuple(fc1_out.shape)} does not match da {tuple(da.shape)}")
    out = _mx_out(M, 2 * I, rowwise, colwise, da.device)
    ws = torch.empty(2 * I, dtype=torch.float32, device=da.device) if want_dbias else None
    dbias = torch.empty(2 * I, dtype=torch.bfloat16, device=da.device) if want_dbias else None
    rc = _lib.load().llmfp8_mx_dswiglu_quantize(da.data_ptr(), fc1_out.data_ptr(), M, I, _p(out.row), _p(out.sf_row), _p(out.col), _p(out.sf_col),
                                                _p(ws), _p(dbias), _stream(da))
    _lib.check(rc, "mx_dswiglu_quantize")
    return out, dbias
