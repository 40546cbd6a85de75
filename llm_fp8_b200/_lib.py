"""ctypes binding of libllmfp8_b200.so (include/llmfp8_b200.h).

The product path has no fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libllmfp8_b200.so")

_lib = None
_lock = threading.Lock()

i64, i32, f32, vp = C.c_int64, C.c_int, C.c_float, C.c_void_p

# name -> argtypes (all return int unless listed in _RESTYPES)
_PROTOTYPES = {
    "llmfp8_abi_version": [],
    "llmfp8_last_error": [],
    "llmfp8_sm_count": [],
    "llmfp8_quant_cast_amax": [vp, vp, i64, i32, vp, vp, vp],
    "llmfp8_quant_cast_amax_bgrad": [vp, vp, i64, i64, i32, vp, vp, vp, vp, vp],
    "llmfp8_rmsnorm_quant_fwd": [vp, vp, vp, vp, i64, i64, f32, i32, vp, vp, vp],
    "llmfp8_rmsnorm_bwd": [vp, vp, vp, vp, vp, vp, vp, i64, i64, vp],
    "llmfp8_rmsnorm_bwd_ws_rows": [i64],
    "llmfp8_swiglu_quant_fwd": [vp, vp, i64, i64, i32, vp, vp, vp],
    "llmfp8_dswiglu_quant_bwd": [vp, vp, vp, i64, i64, i32, vp, vp, vp, vp, vp],
    "llmfp8_amax_scale_update": [vp, vp, vp, vp, i32, i32, i32, vp],
    "llmfp8_mx_quantize": [vp, i64, i64, vp, vp, vp, vp, vp, vp, vp],
    "llmfp8_rmsnorm_rsigma": [vp, vp, i64, i64, f32, vp],
    "llmfp8_mx_scale_bytes": [i64, i64],
    "llmfp8_fp8_gemm": [vp, i64, i32, i32, vp, i64, i32, i32, vp, i64, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp],
    "llmfp8_mxfp8_gemm": [vp, i64, i32, vp, vp, i64, i32, vp, vp, i64, i32, i32, i32, i32, vp, i32, i32, vp],
}
_RESTYPES = {"llmfp8_last_error": C.c_char_p, "llmfp8_mx_scale_bytes": i64}
EXPORTS = tuple(_PROTOTYPES)


def load():
    """Load the shared library (once).  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(or `make -C llm_fp8_b200/csrc`).  llm_fp8_b200 has no CPU or library fallback.")
            lib = C.CDLL(LIB_PATH)
            for name, args in _PROTOTYPES.items():
                fn = getattr(lib, name)
                fn.argtypes = args
                fn.restype = _RESTYPES.get(name, i32)
            if lib.llmfp8_abi_version() != 1:
                raise RuntimeError("libllmfp8_b200.so ABI version mismatch")
            _lib = lib
    return _lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().llmfp8_last_error()
        raise RuntimeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
