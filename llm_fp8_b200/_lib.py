"""ctypes binding of libllmfp8_b200.so (include/llmfp8_b200.h).

The product path has no fallback: if the CUDA library is missing or a call fails, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# LLMFP8_PDL=1 loads the programmatic-dependent-launch build of the same sources (csrc/Makefile, csrc/common.cuh) — same C ABI; that
# variant has not run on hardware yet, so it is never the default.
PDL = os.environ.get("LLMFP8_PDL", "0") == "1"
LIB_PATH = os.path.join(_HERE, "libllmfp8_b200_pdl.so" if PDL else "libllmfp8_b200.so")

_lib = None
_lock = threading.Lock()

i64, i32, f32, vp = C.c_int64, C.c_int, C.c_float, C.c_void_p

# name -> argtypes (all return int unless listed in _RESTYPES)
_PROTOTYPES = {
    "llmfp8_abi_version": [],
    "llmfp8_last_error": [],
    "llmfp8_sm_count": [],
    "llmfp8_launch_count": [],
    "llmfp8_quant_cast_amax": [vp, vp, i64, i32, vp, vp, vp],
    "llmfp8_quant_cast_amax_2d": [vp, i64, vp, i64, i64, i32, vp, vp, vp],
    "llmfp8_quant_cast_amax_bgrad": [vp, vp, i64, i64, i32, vp, vp, vp, vp, vp],
    "llmfp8_rmsnorm_quant_fwd": [vp, vp, vp, vp, i64, i64, f32, i32, vp, vp, vp],
    "llmfp8_rmsnorm_bwd": [vp, vp, vp, vp, vp, vp, vp, i64, i64, vp],
    "llmfp8_rmsnorm_bwd_ws_rows": [i64],
    "llmfp8_rmsnorm_bwd_add": [vp, vp, vp, vp, vp, vp, vp, vp, i64, i64, vp],
    "llmfp8_rmsnorm_bwd_noalloc": [vp, vp, vp, vp, vp, vp, vp, i64, i64, vp],
    "llmfp8_rmsnorm_bwd_noalloc_max_floats": [],
    "llmfp8_swiglu_quant_fwd": [vp, vp, i64, i64, i32, vp, vp, vp],
    "llmfp8_dswiglu_quant_bwd": [vp, vp, vp, i64, i64, i32, vp, vp, vp, vp, vp],
    "llmfp8_amax_scale_update": [vp, vp, vp, vp, i32, i32, i32, vp],
    "llmfp8_mx_quantize": [vp, i64, i64, vp, vp, vp, vp, vp, vp, vp],
    "llmfp8_mx_quantize_ld": [vp, i64, i64, i64, vp, vp, vp, vp, vp, vp, vp],
    "llmfp8_mx_quantize_bgrad": [vp, i64, i64, i64, vp, vp, vp, vp, vp, vp, vp],
    "llmfp8_mx_swiglu_quantize": [vp, i64, i64, vp, vp, vp, vp, vp],
    "llmfp8_mx_dswiglu_quantize": [vp, vp, i64, i64, vp, vp, vp, vp, vp, vp, vp],
    "llmfp8_rmsnorm_rsigma": [vp, vp, i64, i64, f32, vp],
    "llmfp8_mx_scale_bytes": [i64, i64],
    "llmfp8_bf16_gemm": [vp, i64, i32, vp, i64, i32, vp, i64, i32, i32, i32, i32, vp, i32, i32, vp],
    "llmfp8_adamw_mx_step": [vp, vp, vp, vp, i64, i64, f32, f32, f32, f32, f32, f32, f32, vp, vp, vp, vp, vp],
    "llmfp8_lm_head_partials": [i32],
    "llmfp8_lm_head_ce_stats": [vp, i64, vp, i64, i32, i32, i32, vp, vp, vp, vp],
    "llmfp8_ce_reduce_ws_floats": [i32],
    "llmfp8_ce_reduce": [vp, i32, i32, vp, vp, i64, vp, vp, vp, vp],
    "llmfp8_lm_head_dlogits": [vp, i64, vp, i64, i32, i32, i32, vp, i64, vp, vp, vp, i64, vp],
    "llmfp8_gemm_force_bn": [i32],
    "llmfp8_gemm_force_cg": [i32],
    "llmfp8_gemm_debug": [i32],
    "llmfp8_gemm_trace": [vp],
    "llmfp8_fp8_gemm": [vp, i64, i32, i32, vp, i64, i32, i32, vp, i64, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp],
    "llmfp8_mxfp8_gemm": [vp, i64, i32, vp, vp, i64, i32, vp, vp, i64, i32, i32, i32, i32, vp, i32, i32, vp],
    "llmfp8_fp8_gemm_res": [vp, i64, i32, i32, vp, i64, i32, i32, vp, i64, i32, i32, i32, vp, vp, vp, vp, i64, i32, vp],
    "llmfp8_nccl_unique_id": [vp],
    "llmfp8_nccl_init": [vp, i32, i32, C.POINTER(vp)],
    "llmfp8_nccl_allreduce_max_f32": [vp, vp, i64, vp],
    "llmfp8_nccl_allreduce_sum_bf16": [vp, vp, i64, vp],
    "llmfp8_nccl_destroy": [vp],
    "llmfp8_nvls_allreduce_sum_bf16": [vp, i64, vp, i64, i32, i32, i32, i32, vp],
    "llmfp8_nvls_grid": [i64, i32, i32],
    "llmfp8_nvls_amax_scale_update": [vp, vp, vp, vp, vp, i32, i32, i32, vp, i64, i32, i32, i32, vp],
    "llmfp8_nvls_timeouts": [vp],
    "llmfp8_nvls_timeouts_async": [vp, vp],
    "llmfp8_nvls_reset": [],
    "llmfp8_mxfp8_gemm_res": [vp, i64, i32, vp, vp, i64, i32, vp, vp, i64, i32, i32, i32, vp, vp, i64, i32, vp],
}
_RESTYPES = {"llmfp8_last_error": C.c_char_p, "llmfp8_mx_scale_bytes": i64, "llmfp8_rmsnorm_bwd_noalloc_max_floats": i64, "llmfp8_launch_count": C.c_uint64, "llmfp8_gemm_force_bn": None, "llmfp8_gemm_trace": None, "llmfp8_gemm_force_cg": None, "llmfp8_gemm_debug": None}
EXPORTS = tuple(_PROTOTYPES)


def load():
    """Load the shared library (once).  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is not None:
        return _profiled if _profiled is not None else _lib
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
            dbg = os.environ.get("LLMFP8_GEMM_DEBUG")          # experiments only: the bits of llmfp8_gemm_debug (include/llmfp8_b200.h)
            if dbg or os.environ.get("LLMFP8_GEMM_FORCE_CG") or os.environ.get("LLMFP8_GEMM_FORCE_BN"):
                os.environ["LLMFP8_ALLOW_TUNING"] = "1"        # the knobs are inert without it; read by the library at its first setter call
            if dbg:
                lib.llmfp8_gemm_debug(int(dbg, 0))
            for var, setter in (("LLMFP8_GEMM_FORCE_CG", lib.llmfp8_gemm_force_cg), ("LLMFP8_GEMM_FORCE_BN", lib.llmfp8_gemm_force_bn)):
                if os.environ.get(var):                       # tuning probes (tools/gpu_round2_scale.sh); never set in normal use
                    setter(int(os.environ[var]))
            _lib = lib
    return _lib


# ---------------------------------------------------------------------------------------------- per-call device timing
# bench.py's roofline leg: while a profile is active every C-ABI call is bracketed by CUDA events on the current stream and
# recorded with its algorithmic work (FLOPs for GEMMs, HBM bytes for the memory-bound kernels: SURVEY.md §8(d)).
def _work(name: str, a):
    if name == "llmfp8_fp8_gemm" or name == "llmfp8_mxfp8_gemm":
        return "flop", 2.0 * a[11] * a[12] * a[13]
    if name == "llmfp8_fp8_gemm_res" or name == "llmfp8_mxfp8_gemm_res":
        return "flop", 2.0 * a[10] * a[11] * a[12]
    if name == "llmfp8_quant_cast_amax":
        return "byte", 3.0 * a[2]
    if name == "llmfp8_quant_cast_amax_2d":
        return "byte", 3.0 * a[3] * a[4]
    if name == "llmfp8_quant_cast_amax_bgrad":
        return "byte", 3.0 * a[2] * a[3]
    if name == "llmfp8_rmsnorm_quant_fwd":
        return "byte", 3.0 * a[4] * a[5] + 4.0 * a[4]
    if name == "llmfp8_rmsnorm_bwd":
        return "byte", 6.0 * a[7] * a[8]
    if name == "llmfp8_rmsnorm_bwd_add":
        return "byte", 8.0 * a[8] * a[9]
    if name == "llmfp8_rmsnorm_bwd_noalloc":
        return "byte", (8.0 if a[4] is not None else 6.0) * a[7] * a[8]
    if name == "llmfp8_rmsnorm_rsigma":
        return "byte", 2.0 * a[2] * a[3]
    if name == "llmfp8_swiglu_quant_fwd":
        return "byte", (6.0 if a[4] == 2 else 5.0) * a[2] * a[3]
    if name == "llmfp8_dswiglu_quant_bwd":
        return "byte", (10.0 if a[5] == 2 else 8.0) * a[3] * a[4]
    if name == "llmfp8_mx_quantize":
        rc = float(a[1]) * a[2]
        return "byte", 2.0 * rc + (rc + rc / 32) * ((a[3] is not None) + (a[5] is not None))
    if name == "llmfp8_mx_swiglu_quantize":       # read gate | up (4 B per output element), write the copies + scales
        mi = float(a[1]) * a[2]
        return "byte", 4.0 * mi + (mi + mi / 32) * ((a[3] is not None) + (a[5] is not None))
    if name == "llmfp8_mx_dswiglu_quantize":      # read da + gate | up (6 B per [M, I] element), write the copies + scales of [M, 2I]
        mi = float(a[2]) * a[3]
        return "byte", 6.0 * mi + 2.0 * (mi + mi / 32) * ((a[4] is not None) + (a[6] is not None))
    if name == "llmfp8_adamw_mx_step":        # read w, g, m, v; write w, m, v; write the copies + scales
        rc = float(a[4]) * a[5]
        return "byte", 14.0 * rc + (rc + rc / 32) * ((a[13] is not None) + (a[15] is not None))
    if name == "llmfp8_bf16_gemm":
        return "flop", 2.0 * a[9] * a[10] * a[11]
    if name in ("llmfp8_lm_head_ce_stats", "llmfp8_lm_head_dlogits"):
        return "flop", 2.0 * a[4] * a[5] * a[6]
    if name == "llmfp8_mx_quantize_bgrad":
        rc = float(a[2]) * a[3]
        return "byte", 2.0 * rc + (rc + rc / 32) * ((a[4] is not None) + (a[6] is not None))
    if name == "llmfp8_mx_quantize_ld":
        rc = float(a[2]) * a[3]
        return "byte", 2.0 * rc + (rc + rc / 32) * ((a[4] is not None) + (a[6] is not None))
    return "none", 0.0


class _Profiled:
    """Proxy over the ctypes library that times every launch-bearing call with CUDA events."""

    def __init__(self, lib, records):
        self._lib, self._records = lib, records

    def __getattr__(self, name):
        fn = getattr(self._lib, name)
        if name in ("llmfp8_gemm_trace", "llmfp8_gemm_force_bn", "llmfp8_gemm_force_cg", "llmfp8_gemm_debug", "llmfp8_last_error", "llmfp8_abi_version", "llmfp8_sm_count", "llmfp8_launch_count", "llmfp8_mx_scale_bytes",
                    "llmfp8_rmsnorm_bwd_ws_rows", "llmfp8_rmsnorm_bwd_noalloc_max_floats", "llmfp8_lm_head_partials", "llmfp8_ce_reduce_ws_floats") or name.startswith(("llmfp8_nccl_", "llmfp8_nvls_")):
            return fn
        import torch

        def timed(*a):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = fn(*a)
            e1.record()
            unit, work = _work(name, a)
            self._records.append((name, unit, work, e0, e1))
            return rc
        return timed


_profiled = None


def profile_start():
    global _profiled
    _profiled = _Profiled(load(), [])


def profile_stop():
    """-> list of (c_function, unit, algorithmic work, milliseconds); synchronises the device."""
    global _profiled
    import torch
    torch.cuda.synchronize()
    recs = [(n, u, w, e0.elapsed_time(e1)) for n, u, w, e0, e1 in _profiled._records]
    _profiled = None
    return recs


def check(rc: int, what: str):
    if rc != 0:
        msg = load().llmfp8_last_error()
        raise RuntimeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
