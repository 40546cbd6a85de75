"""A few launches of the fc1-shaped fprop GEMM (Llama-3.1-8B: 6144 x 28672 x 4096) per variant, for an ncu capture with source-level
stall sampling:  per-tensor, per-tensor on 5 ring stages (debug 64), MXFP8 (overlapped accumulators), MXFP8 single accumulator (debug 16).
    ncu --set full --import-source on -k regex:fp8_gemm_kernel -o gpurun_out/probe python tools/gemm_probe.py"""
import os
import sys

import torch

os.environ.setdefault("LLMFP8_ALLOW_TUNING", "1")      # this tool drives the GEMM tuning knobs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llm_fp8_b200 import _lib, ops  # noqa: E402

lib = _lib.load()
dev = "cuda"
M, N, K = 6144, 28672, 4096
if len(sys.argv) > 1 and sys.argv[1] == "fc2":
    M, N, K = 6144, 4096, 14336
g = torch.Generator(device=dev).manual_seed(0)
x = (torch.randn(M, K, device=dev, generator=g)).bfloat16()
w = (torch.randn(N, K, device=dev, generator=g) * 0.02).bfloat16()
xm, wm = ops.mx_quantize(x, True, False), ops.mx_quantize(w, True, False)
one = torch.ones(1, device=dev)
X = ops.QTensor(xm.row, 0, one)          # realistic fp8 bit patterns (block-scaled e4m3 of gaussian data) for both kernels
W = ops.QTensor(wm.row, 0, one)
for debug, mx in ((0, False), (64, False), (0, True), (16, True)):
    lib.llmfp8_gemm_debug(debug)
    for _ in range(2):
        if mx:
            ops.mxfp8_gemm(xm.row, xm.sf_row, False, wm.row, wm.sf_row, False, M, N, K)
        else:
            ops.fp8_gemm(X, False, W, False, M, N, K)
    torch.cuda.synchronize()
lib.llmfp8_gemm_debug(0)
print("probe done")
