"""Go / no-go for the overlapped-accumulator MXFP8 GEMM (the default; debug bit 4 = 16 selects the single-accumulator kernel —
when profiles/r01c_gemm_ovl_check.txt was taken the bit had the opposite sense): bitwise equality with the single-accumulator kernel
(same MMA order, so the outputs must be identical) on the three layouts, bf16 and fp32 outputs, full and ragged shapes, then an
alternating A/B timing at the Llama-3.1-8B fc1 shape.  Writes gpurun_out/gemm_ovl_check.txt.   python tools/gemm_ovl_check.py"""
import os
import statistics
import sys

import torch

os.environ.setdefault("LLMFP8_ALLOW_TUNING", "1")      # this tool drives the GEMM tuning knobs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llm_fp8_b200 import _lib, ops  # noqa: E402

lib = _lib.load()
dev = "cuda"
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
log = open(os.path.join(ROOT, "gpurun_out", "gemm_ovl_check.txt"), "w")


def say(*a):
    print(*a, flush=True)
    print(*a, file=log, flush=True)


def operands(M, N, K):
    g = torch.Generator(device=dev).manual_seed(M + 3 * N + 7 * K)
    r = lambda *s: torch.randint(0, 120, s, device=dev, dtype=torch.uint8, generator=g)
    sf = lambda rows, k: torch.randint(120, 134, (ops.mx_scale_bytes(rows, k),), device=dev, dtype=torch.uint8, generator=g)
    x, w, dy = r(M, K), r(N, K), r(M, N)
    return {"fprop": lambda dt: ops.mxfp8_gemm(x, sf(M, K), False, w, sf(N, K), False, M, N, K, out_dtype=dt),
            "dgrad": lambda dt: ops.mxfp8_gemm(dy, sf(M, N), False, w, sf(K, N), True, M, K, N, out_dtype=dt),
            "wgrad": lambda dt: ops.mxfp8_gemm(dy, sf(N, M), True, x, sf(K, M), True, N, K, M, out_dtype=dt)}


ok = True
for (M, N, K) in [(6144, 28672, 4096), (992, 1344, 992), (512, 768, 4096)]:
    for name in ("fprop", "dgrad", "wgrad"):
        for dt in (torch.bfloat16, torch.float32):
            outs = []
            for mode in (0, 16):
                lib.llmfp8_gemm_debug(mode)
                lib.llmfp8_gemm_force_bn(256)
                fn = operands(M, N, K)[name]          # same seed -> same operands and scales for both modes
                outs.append(fn(dt))
            torch.cuda.synchronize()
            same = torch.equal(outs[0], outs[1])
            ok &= same
            say(f"{name} {M}x{N}x{K} {str(dt)[6:]}: {'identical' if same else 'DIFFERENT'}  finite={bool(torch.isfinite(outs[1].float()).all())}")
lib.llmfp8_gemm_force_bn(0)


def t(fn, iters=6):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


M, N, K = 6144, 28672, 4096
x = torch.randint(0, 120, (M, K), device=dev, dtype=torch.uint8)
w = torch.randint(0, 120, (N, K), device=dev, dtype=torch.uint8)
dy = torch.randint(0, 120, (M, N), device=dev, dtype=torch.uint8)
one = lambda rows, k: torch.full((ops.mx_scale_bytes(rows, k),), 127, device=dev, dtype=torch.uint8)
sx, sw, sdy, swc, sdyc, sxc = one(M, K), one(N, K), one(M, N), one(K, N), one(N, M), one(K, M)
cases = [("G1m fprop fc1", lambda: ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K)),
         ("G2m dgrad fc1", lambda: ops.mxfp8_gemm(dy, sdy, False, w, swc, True, M, K, N)),
         ("G3m wgrad fc1", lambda: ops.mxfp8_gemm(dy, sdyc, True, x, sxc, True, N, K, M))]
flops = 2.0 * M * N * K
for cname, fn in cases:
    res = {0: [], 16: []}
    for m in res:
        lib.llmfp8_gemm_debug(m)
        t(fn, 2)
    for _ in range(4):
        for m in res:
            lib.llmfp8_gemm_debug(m)
            res[m].append(t(fn))
    say(cname, "  ".join(f"mode{m}: {statistics.median(v) * 1e3:6.0f} us {flops / statistics.median(v) / 1e9:6.0f} TF" for m, v in res.items()))
lib.llmfp8_gemm_debug(0)
say("OVL_CHECK", "PASS" if ok else "FAIL")
