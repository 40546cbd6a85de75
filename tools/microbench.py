"""Per-kernel microbenchmarks on one B200: quantise kernels in GB/s (algorithmic bytes), GEMMs in TFLOP/s,
next to on-GPU comparison points (torch bf16 matmul, torch._scaled_mm = cuBLASLt FP8) — comparison points only, never a fallback.

    python tools/microbench.py [--model llama8b] [--out gpurun_out/micro.jsonl]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

os.environ.setdefault("LLMFP8_ALLOW_TUNING", "1")      # this tool drives the GEMM tuning knobs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import llm_fp8_b200 as te  # noqa: E402
from llm_fp8_b200 import _lib, ops  # noqa: E402

SHAPES = {  # H, I, qkv_out, M
    "llama1b": (2048, 8192, 3072, 8192),
    "llama3b": (3072, 8192, 5120, 8192),
    "llama8b": (4096, 14336, 6144, 6144),
    "qwen14b": (5120, 13824, 7168, 8192),
}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


class NvmlClocks:
    """SM clock / power / throttle reasons sampled every 5 ms by a thread through NVML while one kernel is being timed, so that every
    line this tool emits carries its own clock record (bench.ClockSampler's nvidia-smi loop is too coarse for a 10-launch timing)."""

    def __init__(self, index=0):
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = str(e)[:100]

    def __enter__(self):
        import threading
        self.sm, self.pw, self.reasons, self._stop = [], [], 0, False
        if self.ok:
            def loop():
                import time
                while not self._stop:
                    try:
                        self.sm.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                        self.pw.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
                        self.reasons |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                    except Exception:  # noqa: BLE001
                        pass
                    time.sleep(0.005)
            self._t = threading.Thread(target=loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop = True
        if self.ok:
            self._t.join()

    def record(self):
        if not self.ok or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable"]}
        nv = self.nv
        names = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown), ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown), ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap)]
        s = sorted(self.sm)
        return {"sm_mhz": s[len(s) // 2], "sm_mhz_min": s[0], "sm_max_mhz": self.max_mhz, "power_w_max": max(self.pw) if self.pw else None,
                "samples": len(s), "reasons": [n for n, bit in names if self.reasons & bit]}


_clocks = None
_last_clocks = None
_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    _flush.zero_()


def timeit(fn, iters=10, warm=3, flush=True):
    global _clocks, _last_clocks
    if _clocks is None:
        _clocks = NvmlClocks(torch.cuda.current_device())
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    with _clocks:
        for _ in range(iters):
            if flush:
                flush_l2()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
    _last_clocks = _clocks.record()
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama8b")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "micro.jsonl"))
    ap.add_argument("--only", default="")
    ap.add_argument("--bn", type=int, default=0, help="force the GEMM tile width (0 = heuristic)")
    ap.add_argument("--cg", type=int, default=0, help="force the CTA group (1 | 2; 0 = heuristic)")
    ap.add_argument("--debug", type=int, default=0)
    ap.add_argument("--tag", default="")
    ap.add_argument("--kfilter", default="", help="only run memory-bound kernels whose name contains this substring")
    ap.add_argument("--experimental", action="store_true", help="also time the code paths that have not been validated on hardware yet")
    ap.add_argument("--ms", default="", help="comma-separated token counts to sweep instead of the model's default M (BASELINE.json config 5: 512..16384)")
    ap.add_argument("--no-bwd", action="store_true", help="GEMM sweep: fprop forms only")
    args = ap.parse_args()
    H, I, QKV, M0 = SHAPES[args.model]
    hbm, bf16_tf, src = peaks()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    out = open(args.out, "a")

    from llm_fp8_b200 import _lib
    _lib.load().llmfp8_gemm_force_bn(args.bn)
    _lib.load().llmfp8_gemm_force_cg(args.cg)
    _lib.load().llmfp8_gemm_debug(args.debug)

    def emit(rec):
        rec["model"] = args.model
        rec["bn"] = args.bn
        rec["cg"] = args.cg
        rec["debug"] = args.debug
        rec["clocks"] = _last_clocks          # sampled while THIS kernel was being timed
        if args.tag:
            rec["tag"] = args.tag
        print(json.dumps(rec), flush=True)
        out.write(json.dumps(rec) + "\n")
        out.flush()

    for M in ([int(v) for v in args.ms.split(",")] if args.ms else [M0]):
        run_one(args, emit, M, H, I, QKV, hbm, bf16_tf, src)
    out.close()


def run_one(args, emit, M, H, I, QKV, hbm, bf16_tf, src):
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(1234)
    scale = torch.tensor([32.0], device=dev)
    amax = torch.zeros(1, device=dev)

    def mem(name, fn, nbytes):
        if args.kfilter and args.kfilter not in name:
            return
        med, best = timeit(fn)
        emit({"kernel": name, "ms": med, "ms_best": best, "GBps": nbytes / med / 1e6, "frac_hbm": nbytes / med / 1e6 / hbm, "peak": src, "bytes": nbytes, "M": M})

    if not args.only or "quant" in args.only:
        x = torch.randn(M, H, device=dev, generator=g).bfloat16()
        gamma = torch.ones(H, device=dev, dtype=torch.bfloat16)
        w1 = torch.randn(2 * I, H, device=dev, generator=g).bfloat16() * 0.02
        fc1 = torch.randn(M, 2 * I, device=dev, generator=g).bfloat16()
        da = torch.randn(M, I, device=dev, generator=g).bfloat16()
        mem("K2 cast_amax x[M,H]", lambda: ops.quantize_raw(x, 0, scale, amax), 3 * M * H)
        mem("K2 cast_amax W1[2I,H]", lambda: ops.quantize_raw(w1, 0, scale, amax), 3 * 2 * I * H)
        mem("K2b cast_amax_bgrad dY[M,H]", lambda: ops.quantize_bgrad_raw(x, 1, scale, amax), 3 * M * H)
        mem("K1 rmsnorm_quant [M,H]", lambda: ops.rmsnorm_quant_raw(x, gamma, 1e-5, 0, scale, amax), 3 * M * H + 4 * M)
        mem("K3 swiglu_quant [M,2I]", lambda: ops.swiglu_quant_raw(fc1, 0, scale, amax), 5 * M * I)
        mem("K3b dswiglu_quant [M,2I]", lambda: ops.dswiglu_quant_raw(da, fc1, 1, scale, amax), 8 * M * I)
        rs = ops.rmsnorm_rsigma(x, 1e-5)
        mem("K5 rmsnorm_bwd [M,H]", lambda: ops.rmsnorm_bwd(x, x, gamma, rs), 6 * M * H)
        mem("K5 rmsnorm_bwd + dres [M,H]", lambda: ops.rmsnorm_bwd(x, x, gamma, rs, add=x), 8 * M * H)
        wide = torch.randn(M, QKV, device=dev, generator=g).bfloat16()
        qv = wide[:, :H]
        mem("K2 cast_amax query slice of [M,QKV] (pitch)", lambda: ops.quantize_raw(qv, 0, scale, amax), 3 * M * H)
        mem("K2m mx_quantize row+col query slice of [M,QKV] (pitch)", lambda: ops.mx_quantize(qv, True, True), int(4.0625 * M * H))
        del wide, qv
        mem("K2m mx_quantize row+col x[M,H]", lambda: ops.mx_quantize(x, True, True), int(4.0625 * M * H))
        mem("K2m mx_quantize row x[M,H]", lambda: ops.mx_quantize(x, True, False), int(3.03125 * M * H))
        mem("K2m mx_quantize row+col W1[2I,H]", lambda: ops.mx_quantize(w1, True, True), int(4.0625 * 2 * I * H))
        mem("K1m rsigma + mx_quantize row+col [M,H]", lambda: ops.mx_quantize(x, True, True, ops.rmsnorm_rsigma(x, 1e-5), gamma), int(4.0625 * M * H))
        # MXFP8 activation path: the two-kernel sequences and (--experimental) their one-pass replacements (csrc/mx_act_quantize.cu)
        mem("K3m swiglu_bf16 + mx_quantize row+col [M,I] (two kernels)", lambda: ops.mx_quantize(ops.swiglu_bf16(fc1), True, True),
            int(10.0625 * M * I))
        mem("K3mb dswiglu_bf16 + mx_quantize row+col [M,2I] (two kernels)", lambda: ops.mx_quantize(ops.dswiglu_bf16(da, fc1)[0], True, True),
            int(18.125 * M * I))
        mem("K3m mx_swiglu_quantize row+col [M,I] (one pass)", lambda: ops.mx_swiglu_quantize(fc1, True, True), int(6.0625 * M * I))
        if I % 128 == 0:
            mem("K3mb mx_dswiglu_quantize row+col [M,2I] (one pass)", lambda: ops.mx_dswiglu_quantize(da, fc1, True, True), int(10.125 * M * I))
        mem("copy bf16 [M,2I] (torch, reference point)", lambda: fc1.clone(), 4 * M * 2 * I)
        mem("copy bf16 [M,H] (torch, reference point for the short kernels)", lambda: x.clone(), 4 * M * H)
        del x, w1, fc1, da

    if "sustained" in args.only:
        # the twelve GEMMs of one decoder layer (fprop, dgrad, wgrad of qkv / proj / fc1 / fc2) back to back for ~3 s: what the GEMMs alone
        # sustain under the board's power cap — the ceiling for the in-step GEMM rate bench.py reports (roofline.achieved)
        import time
        one = torch.ones(1, device=dev)
        ops_list, fl = [], 0.0
        for lname, N, K in [("qkv", QKV, H), ("proj", H, H), ("fc1", 2 * I, H), ("fc2", H, I)]:
            xb = torch.randn(M, K, device=dev, generator=g).bfloat16()
            wb = (torch.randn(N, K, device=dev, generator=g) * 0.02).bfloat16()
            dyb = (torch.randn(M, N, device=dev, generator=g) * 1e-3).bfloat16()
            xm, wm, dym = ops.mx_quantize(xb, True, True), ops.mx_quantize(wb, True, True), ops.mx_quantize(dyb, True, True)
            X, W, DY = ops.QTensor(xm.row, 0, one), ops.QTensor(wm.row, 0, one), ops.QTensor(dym.row, 1, one)
            for mx in (False, True):
                if mx:
                    fns = [lambda xm=xm, wm=wm, N=N, K=K: ops.mxfp8_gemm(xm.row, xm.sf_row, False, wm.row, wm.sf_row, False, M, N, K),
                           lambda dym=dym, wm=wm, N=N, K=K: ops.mxfp8_gemm(dym.row, dym.sf_row, False, wm.col, wm.sf_col, True, M, K, N),
                           lambda dym=dym, xm=xm, N=N, K=K: ops.mxfp8_gemm(dym.col, dym.sf_col, True, xm.col, xm.sf_col, True, N, K, M)]
                else:
                    fns = [lambda X=X, W=W, N=N, K=K: ops.fp8_gemm(X, False, W, False, M, N, K),
                           lambda DY=DY, W=W, N=N, K=K: ops.fp8_gemm(DY, False, W, True, M, K, N),
                           lambda DY=DY, X=X, N=N, K=K: ops.fp8_gemm(DY, True, X, True, N, K, M)]
                ops_list.append((mx, fns))
            fl += 3 * 2.0 * M * N * K
        for want_mx in (True, False):
            fns = [f for mx, fs in ops_list if mx == want_mx for f in fs]
            for f in fns:
                f()
            torch.cuda.synchronize()
            global _clocks, _last_clocks
            if _clocks is None:
                _clocks = NvmlClocks(torch.cuda.current_device())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with _clocks:
                t0, n = time.time(), 0
                e0.record()
                while time.time() - t0 < 3.0:
                    for f in fns:
                        f()
                    n += 1
                    if n % 8 == 0:
                        torch.cuda.synchronize()
                e1.record()
                torch.cuda.synchronize()
            _last_clocks = _clocks.record()
            ms = e0.elapsed_time(e1) / n
            emit({"kernel": f"sustained: the 12 GEMMs of a layer back to back, {'mxfp8' if want_mx else 'per-tensor'}, {n} rounds", "M": M, "ms": ms,
                  "TFLOPs": fl / ms / 1e9, "frac_2x_bf16_sustained_proxy": None, "frac_fp8_nominal": fl / ms / 1e9 / 4500.0})

    if "optim" in args.only:
        # fused AdamW + MXFP8 requantise of the fc1 weight [2I, H] against what it replaces: torch.optim.AdamW's step on that weight + the block
        # quantiser's pass over the updated weight in the next forward
        w = (torch.randn(2 * I, H, device=dev, generator=g) * 0.02).bfloat16().requires_grad_(True)
        w.grad = (torch.randn(2 * I, H, device=dev, generator=g) * 1e-4).bfloat16()
        mm, vv = torch.zeros_like(w), torch.zeros_like(w)
        out = ops._mx_out(2 * I, H, True, True, w.device)
        lib = _lib.load()
        st = torch.cuda.current_stream().cuda_stream
        n = 2 * I * H

        def fused():
            lib.llmfp8_adamw_mx_step(w.data_ptr(), w.grad.data_ptr(), mm.data_ptr(), vv.data_ptr(), 2 * I, H, 1e-4, 0.9, 0.95, 1e-8, 0.1, 0.1, 0.05,
                                     out.row.data_ptr(), out.sf_row.data_ptr(), out.col.data_ptr(), out.sf_col.data_ptr(), st)
        topt = torch.optim.AdamW([w], lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1)
        mem("O1 fused AdamW + MX requantise W1[2I,H] (one pass)", fused, int(16.0625 * n))
        mem("O1 torch.optim.AdamW step W1[2I,H] (bf16 state; comparison point)", lambda: topt.step(), int(14 * n))
        mem("O1 torch.optim.AdamW step + mx_quantize row+col (what the fused pass replaces)", lambda: (topt.step(), ops.mx_quantize(w.detach(), True, True)), int(18.0625 * n))

    if "lmhead" in args.only:
        # lm_head + cross-entropy at the model's head shape, next to the reference's sequence in torch
        # (bf16 matmul -> logits [M, V] -> fp32 cross-entropy), forward and forward + backward; bf16 tcgen05 GEMM vs cuBLAS at the head's shapes
        V = {"llama1b": 128256, "llama3b": 128256, "llama8b": 128256, "qwen14b": 152064}[args.model]
        Mh = min(M, 4096)
        xh = (torch.randn(Mh, H, device=dev, generator=g)).bfloat16().requires_grad_(True)
        wh = (torch.randn(V, H, device=dev, generator=g) * 0.02).bfloat16().requires_grad_(True)
        lab = torch.randint(0, V, (Mh,), device=dev, generator=g)
        fl = 2.0 * Mh * V * H

        def rec_tf(name, fn, flops):
            med, best = timeit(fn, iters=5, warm=2, flush=False)
            emit({"kernel": f"{name} M={Mh} V={V} H={H}", "M": Mh, "ms": med, "ms_best": best, "TFLOPs": flops / med / 1e9,
                  "frac_bf16_measured": flops / med / 1e9 / bf16_tf})

        def ours_fb():
            xh.grad = wh.grad = None
            te.lm_head_cross_entropy(xh, wh, lab).backward()

        def torch_fb():
            xh.grad = wh.grad = None
            torch.nn.functional.cross_entropy((xh @ wh.t()).float(), lab).backward()
        with torch.no_grad():
            rec_tf("lm_head loss fwd (ours: GEMM + CE statistics in the epilogue, no logits)", lambda: te.lm_head_cross_entropy(xh.detach(), wh.detach(), lab), fl)
            rec_tf("lm_head loss fwd (torch: bf16 matmul + fp32 cross_entropy)", lambda: torch.nn.functional.cross_entropy((xh @ wh.t()).float(), lab), fl)
            rec_tf("bf16 GEMM logits (ours, tcgen05 kind::f16)", lambda: te.bf16_gemm(xh.detach(), False, wh.detach(), False, Mh, V, H), fl)
            rec_tf("bf16 GEMM logits (torch.matmul, cuBLAS)", lambda: xh @ wh.t(), fl)
        rec_tf("lm_head loss fwd+bwd (ours: 4 GEMM passes)", ours_fb, 4 * fl)
        rec_tf("lm_head loss fwd+bwd (torch: 3 GEMM passes + logits in fp32)", torch_fb, 3 * fl)
        del xh, wh

    if not args.only or "gemm" in args.only:
        one = torch.ones(1, device=dev)
        layers = [("qkv", QKV, H), ("proj", H, H), ("fc1", 2 * I, H), ("fc2", H, I)]
        for lname, N, K in layers:
            xq = torch.randint(0, 120, (M, K), device=dev, dtype=torch.uint8)
            wq = torch.randint(0, 120, (N, K), device=dev, dtype=torch.uint8)
            dyq = torch.randint(0, 120, (M, N), device=dev, dtype=torch.uint8)
            X, W, DY = ops.QTensor(xq, 0, one), ops.QTensor(wq, 0, one), ops.QTensor(dyq, 1, one)
            flops = 2.0 * M * N * K

            def rec(name, fn, extra=None):
                if args.no_bwd and ("dgrad" in name or "wgrad" in name):
                    return
                med, best = timeit(fn, iters=8, warm=3, flush=False)
                r = {"kernel": f"{name} {lname} M={M} N={N} K={K}", "M": M, "ms": med, "ms_best": best, "TFLOPs": flops / med / 1e9,
                     "frac_2x_bf16_measured": flops / med / 1e9 / (2 * bf16_tf), "frac_fp8_nominal": flops / med / 1e9 / 4500.0}
                if extra:
                    r.update(extra)
                emit(r)

            rec("G1 fprop fp8", lambda: ops.fp8_gemm(X, False, W, False, M, N, K))
            if lname in ("proj", "fc2"):
                resid = torch.randn(M, N, device=dev, dtype=torch.bfloat16)
                rec("G1 fprop fp8 + residual", lambda: ops.fp8_gemm(X, False, W, False, M, N, K, residual=resid))
            rec("G2 dgrad fp8", lambda: ops.fp8_gemm(DY, False, W, True, M, K, N))
            rec("G3 wgrad fp8", lambda: ops.fp8_gemm(DY, True, X, True, N, K, M))
            sfx = torch.full((ops.mx_scale_bytes(M, K),), 127, device=dev, dtype=torch.uint8)
            sfw = torch.full((ops.mx_scale_bytes(N, K),), 127, device=dev, dtype=torch.uint8)
            rec("G1m fprop mxfp8", lambda: ops.mxfp8_gemm(xq, sfx, False, wq, sfw, False, M, N, K))
            if lname in ("proj", "fc2"):
                rec("G1m fprop mxfp8 + residual", lambda: ops.mxfp8_gemm(xq, sfx, False, wq, sfw, False, M, N, K, residual=resid))
            sfdy = torch.full((ops.mx_scale_bytes(M, N),), 127, device=dev, dtype=torch.uint8)
            sfwc = torch.full((ops.mx_scale_bytes(K, N),), 127, device=dev, dtype=torch.uint8)
            rec("G2m dgrad mxfp8", lambda: ops.mxfp8_gemm(dyq, sfdy, False, wq, sfwc, True, M, K, N))
            sfdyc = torch.full((ops.mx_scale_bytes(N, M),), 127, device=dev, dtype=torch.uint8)
            sfxc = torch.full((ops.mx_scale_bytes(K, M),), 127, device=dev, dtype=torch.uint8)
            rec("G3m wgrad mxfp8", lambda: ops.mxfp8_gemm(dyq, sfdyc, True, xq, sfxc, True, N, K, M))
            # comparison points (library): bf16 matmul and cuBLASLt FP8
            xb = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
            wb = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
            rec("ref torch.matmul bf16 (cuBLAS)", lambda: xb @ wb.t())
            try:
                x8, w8 = xq.view(torch.float8_e4m3fn), wq.view(torch.float8_e4m3fn)
                rec("ref torch._scaled_mm e4m3 (cuBLASLt)", lambda: torch._scaled_mm(x8, w8.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16))
            except Exception as e:  # noqa: BLE001
                emit({"kernel": f"ref torch._scaled_mm {lname}", "error": str(e)[:200]})
            try:   # cuBLASLt block-scaled MXFP8 (comparison point)
                e8 = torch.float8_e8m0fnu
                rec("ref torch._scaled_mm mxfp8 (cuBLASLt)", lambda: torch._scaled_mm(x8, w8.t(), scale_a=sfx.view(e8), scale_b=sfw.view(e8), out_dtype=torch.bfloat16))
            except Exception as e:  # noqa: BLE001
                emit({"kernel": f"ref torch._scaled_mm mxfp8 {lname}", "error": str(e)[:300]})
            del xq, wq, dyq, xb, wb


if __name__ == "__main__":
    main()
