"""ctypes binding of oracle/liboracle_c.so (oracle_c.c).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle_c.so")
FMT = {torch.float8_e4m3fn: 0, torch.float8_e5m2: 1}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle_c.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.orc_fp8_to_f32.restype = C.c_float
        _lib.orc_f32_to_fp8.restype = C.c_uint8
        _lib.orc_f32_to_fp8.argtypes = [C.c_float, C.c_int]
        _lib.orc_fp8_to_f32.argtypes = [C.c_uint8, C.c_int]
        _lib.orc_e8m0_roundup.restype = C.c_uint8
        _lib.orc_e8m0_roundup.argtypes = [C.c_float]
        _lib.orc_exp_neg.restype = C.c_float
        _lib.orc_exp_neg.argtypes = [C.c_float]
        _lib.orc_sigmoid.restype = C.c_float
        _lib.orc_sigmoid.argtypes = [C.c_float]
        _lib.orc_rms_rsigma.restype = C.c_float
        _lib.orc_sf_offset.restype = C.c_int64
        _lib.orc_sf_offset.argtypes = [C.c_int64] * 3
    return _lib


def _p(t: torch.Tensor):
    assert t.is_contiguous() and t.device.type == "cpu"
    return C.c_void_p(t.data_ptr())


def _bf16(x):
    return x.detach().to(torch.bfloat16).contiguous().cpu()


def cast_f32(x: torch.Tensor, dtype) -> torch.Tensor:
    x = x.detach().float().contiguous()
    q = torch.empty(x.shape, dtype=torch.uint8)
    lib().orc_cast_f32(_p(x), C.c_int64(x.numel()), FMT[dtype], _p(q))
    return q


def decode(q: torch.Tensor, dtype) -> torch.Tensor:
    q = q.contiguous().view(torch.uint8)
    out = torch.empty(q.shape, dtype=torch.float32)
    lib().orc_decode(_p(q), C.c_int64(q.numel()), FMT[dtype], _p(out))
    return out


def quantize_bf16(x: torch.Tensor, scale: float, dtype):
    x = _bf16(x)
    q = torch.empty(x.shape, dtype=torch.uint8)
    amax = C.c_float(0)
    lib().orc_quantize_bf16(_p(x), C.c_int64(x.numel()), C.c_float(scale), FMT[dtype], _p(q), C.byref(amax))
    return q, amax.value


def colsum_bf16(x: torch.Tensor) -> torch.Tensor:
    x = _bf16(x)
    R, Cc = x.shape
    out = torch.empty(Cc, dtype=torch.float32)
    lib().orc_colsum_bf16(_p(x), C.c_int64(R), C.c_int64(Cc), _p(out))
    return out


def amax_scale_update(hist, scale, scale_inv, fmax, margin=0):
    L, T = hist.shape
    lib().orc_amax_scale_update(_p(hist), _p(scale), _p(scale_inv), _p(fmax), L, T, margin)


def rmsnorm_quant(x, gamma, eps, scale, dtype, want_y=False):
    x, gamma = _bf16(x), _bf16(gamma)
    M, H = x.shape
    q = torch.empty(M, H, dtype=torch.uint8)
    rs = torch.empty(M, dtype=torch.float32)
    y = torch.empty(M, H, dtype=torch.float32) if want_y else None
    amax = C.c_float(0)
    lib().orc_rmsnorm_quant(_p(x), _p(gamma), C.c_int64(M), C.c_int64(H), C.c_float(eps), C.c_float(scale),
                            FMT[dtype], _p(q), _p(rs), C.byref(amax), _p(y) if want_y else None)
    return q, rs, amax.value, y


def rmsnorm_y(x, gamma, eps):
    x, gamma = _bf16(x), _bf16(gamma)
    M, H = x.shape
    y = torch.empty(M, H, dtype=torch.float32)
    rs = torch.empty(M, dtype=torch.float32)
    lib().orc_rmsnorm_y(_p(x), _p(gamma), C.c_int64(M), C.c_int64(H), C.c_float(eps), _p(y), _p(rs))
    return y, rs


def swiglu_f32(fc1):
    fc1 = _bf16(fc1)
    M, I2 = fc1.shape
    a = torch.empty(M, I2 // 2, dtype=torch.float32)
    lib().orc_swiglu_f32(_p(fc1), C.c_int64(M), C.c_int64(I2 // 2), _p(a))
    return a


def swiglu_quant(fc1, scale, dtype):
    fc1 = _bf16(fc1)
    M, I2 = fc1.shape
    q = torch.empty(M, I2 // 2, dtype=torch.uint8)
    amax = C.c_float(0)
    lib().orc_swiglu_quant(_p(fc1), C.c_int64(M), C.c_int64(I2 // 2), C.c_float(scale), FMT[dtype], _p(q), C.byref(amax))
    return q, amax.value


def dswiglu_f32(da, fc1):
    da, fc1 = _bf16(da), _bf16(fc1)
    M, I = da.shape
    d = torch.empty(M, 2 * I, dtype=torch.float32)
    lib().orc_dswiglu_f32(_p(da), _p(fc1), C.c_int64(M), C.c_int64(I), _p(d))
    return d


def mx_quantize_f32(x, rowwise=True, colwise=True):
    """Returns dict(q_row, sf_row (swizzled flat), q_col, sf_col (swizzled flat))."""
    x = x.detach().float().contiguous().cpu()
    R, Cc = x.shape
    out = {}

    def sfsize(rows, kb):
        return ((rows + 127) // 128 * 128) * ((kb + 3) // 4 * 4)

    q_row = torch.empty(R, Cc, dtype=torch.uint8) if rowwise else None
    sf_row = torch.zeros(sfsize(R, Cc // 32), dtype=torch.uint8) if rowwise else None
    q_col = torch.empty(R, Cc, dtype=torch.uint8) if colwise else None
    sf_col = torch.zeros(sfsize(Cc, R // 32), dtype=torch.uint8) if colwise else None
    lib().orc_mx_quantize_f32(_p(x), C.c_int64(R), C.c_int64(Cc), int(rowwise), int(colwise),
                              _p(q_row) if rowwise else None, _p(sf_row) if rowwise else None,
                              _p(q_col) if colwise else None, _p(sf_col) if colwise else None)
    return dict(q_row=q_row, sf_row=sf_row, q_col=q_col, sf_col=sf_col)


def num_threads() -> int:
    return lib().orc_num_threads()
