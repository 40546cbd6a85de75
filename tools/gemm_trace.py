"""Where does the MXFP8 GEMM's tensor pipe idle between tiles?  clock64 stamps of cluster 0 (llmfp8_gemm_trace) for fc1 fprop at Llama-3.1-8B
shapes: per tile, the MMA thread's waits (accumulator free, shared columns free, first stage full) and the epilogue's phases.
    python tools/gemm_trace.py [mx|pt]"""
import os
import sys

import torch

os.environ.setdefault("LLMFP8_ALLOW_TUNING", "1")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llm_fp8_b200 import _lib, ops  # noqa: E402

lib = _lib.load()
dev = "cuda"
M, N, K = 6144, 28672, 4096
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(M, K, device=dev, generator=g).bfloat16()
w = (torch.randn(N, K, device=dev, generator=g) * 0.02).bfloat16()
xm, wm = ops.mx_quantize(x, True, False), ops.mx_quantize(w, True, False)
one = torch.ones(1, device=dev)
X, W = ops.QTensor(xm.row, 0, one), ops.QTensor(wm.row, 0, one)
for kind in (sys.argv[1:] or ["mx", "pt"]):
    run = (lambda: ops.mxfp8_gemm(xm.row, xm.sf_row, False, wm.row, wm.sf_row, False, M, N, K)) if kind == "mx" else (lambda: ops.fp8_gemm(X, False, W, False, M, N, K))
    for _ in range(3):
        run()
    buf = torch.zeros(64 * 16, dtype=torch.int64, device=dev)
    lib.llmfp8_gemm_trace(buf.data_ptr())
    run()
    torch.cuda.synchronize()
    lib.llmfp8_gemm_trace(None)
    t = buf.cpu().view(64, 16)
    n = int((t[:, 0] > 0).sum())
    print(f"== {kind}: {n} tiles on cluster 0 (clock64 cycles)")
    print("tile  period | mma: acc_free  ovl_free  stage_full  issue_span | since last issue: epi_wake  ld1  ld2  epi_done   [warp 2 / warp 6]")
    for i in range(1, n):
        a, p = t[i], t[i - 1]
        row = [int(a[0] - p[0]), int(a[1] - a[0]), int(a[2] - a[1]), int(a[3] - a[2]), int(a[4] - a[3])]
        e2 = [int(a[5] - a[4]), int(a[6] - a[5]), int(a[7] - a[6]), int(a[8] - a[5])] if a[5] > 0 else []
        e6 = [int(a[10] - a[4]), int(a[11] - a[10]), int(a[12] - a[11]), int(a[13] - a[10])] if a[10] > 0 else []
        print(f"{i:4d} {row[0]:7d} | {row[1]:8d} {row[2]:9d} {row[3]:10d} {row[4]:10d} | {e2} {e6}")
