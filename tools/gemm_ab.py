"""A/B of GEMM scheduling modes inside one process (alternating, so clocks / temperature affect both arms alike).
    python tools/gemm_ab.py [--modes 0,8]      # debug bit 3 (8) = static round-robin persistent grid, 0 = CLC dynamic"""
import argparse
import os
import statistics
import sys

import torch

os.environ.setdefault("LLMFP8_ALLOW_TUNING", "1")      # this tool drives the GEMM tuning knobs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llm_fp8_b200 import _lib, ops  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--modes", default="0,8")
ap.add_argument("--rounds", type=int, default=5)
a = ap.parse_args()
modes = [int(m) for m in a.modes.split(",")]
lib = _lib.load()
dev = "cuda"
M, H, I, QKV = 6144, 4096, 14336, 6144
one = torch.ones(1, device=dev)


def t(fn, iters=6):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


for lname, N, K in [("qkv", QKV, H), ("proj", H, H), ("fc1", 2 * I, H), ("fc2", H, I)]:
    xq = torch.randint(0, 120, (M, K), device=dev, dtype=torch.uint8)
    wq = torch.randint(0, 120, (N, K), device=dev, dtype=torch.uint8)
    dyq = torch.randint(0, 120, (M, N), device=dev, dtype=torch.uint8)
    X, W, DY = ops.QTensor(xq, 0, one), ops.QTensor(wq, 0, one), ops.QTensor(dyq, 1, one)
    sfx = torch.full((ops.mx_scale_bytes(M, K),), 127, device=dev, dtype=torch.uint8)
    sfw = torch.full((ops.mx_scale_bytes(N, K),), 127, device=dev, dtype=torch.uint8)
    sfdyc = torch.full((ops.mx_scale_bytes(N, M),), 127, device=dev, dtype=torch.uint8)
    sfxc = torch.full((ops.mx_scale_bytes(K, M),), 127, device=dev, dtype=torch.uint8)
    cases = [("G1 fprop", lambda: ops.fp8_gemm(X, False, W, False, M, N, K)),
             ("G2 dgrad", lambda: ops.fp8_gemm(DY, False, W, True, M, K, N)),
             ("G3 wgrad", lambda: ops.fp8_gemm(DY, True, X, True, N, K, M)),
             ("G1m fprop", lambda: ops.mxfp8_gemm(xq, sfx, False, wq, sfw, False, M, N, K)),
             ("G3m wgrad", lambda: ops.mxfp8_gemm(dyq, sfdyc, True, xq, sfxc, True, N, K, M))]
    flops = 2.0 * M * N * K
    for cname, fn in cases:
        res = {m: [] for m in modes}
        for m in modes:
            lib.llmfp8_gemm_debug(m)
            t(fn, 3)
        for _ in range(a.rounds):
            for m in modes:
                lib.llmfp8_gemm_debug(m)
                res[m].append(t(fn))
        print(f"{cname:10s} {lname:5s}", "  ".join(f"mode{m}: {flops / statistics.median(v) / 1e9:6.0f} TF" for m, v in res.items()), flush=True)
lib.llmfp8_gemm_debug(0)
