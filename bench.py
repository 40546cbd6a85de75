#!/usr/bin/env python
"""bench.py — headline benchmark of the FP8 / MXFP8 linear-layer hot path (BASELINE.json metric).

A *step* = one forward + backward pass of the linear-layer path of a whole decoder stack (qkv, proj, fc1, fc2 GEMMs in
fprop/dgrad/wgrad form, the fused RMSNorm / SwiGLU / quantise kernels around them) over one batch of synthetic tokens,
driven through the plugin surface the reference's TELlamaDecoderLayer uses (LayerNormLinear / Linear / LayerNormMLP under
fp8_autocast).  Default workload = BASELINE.json configs[3]: Llama-3.1-8B shapes, MXFP8BlockScaling, batch 12 x seq 512 per
GPU, 32 layers, random-init weights.  Metric: linear TFLOP/s = 6 * tokens * P_linear / step time.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--scenario mxfp8|default|hybrid] [--model llama-3.1-8b] [--impl reference]

N > 1: launched under torchrun, one rank per GPU; each rank owns `batch` sequences (weak scaling), the delayed-scaling
amax pools are MAX-all-reduced and the weight gradients averaged over NCCL, overlapped with backward.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "fp8_linear_path_tflops"
UNIT = "TFLOP/s"
FP8_NOMINAL_TFLOPS = 4500.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama-3.1-8b")
    ap.add_argument("--scenario", default="mxfp8", choices=["mxfp8", "default", "hybrid"],
                    help="reference decoder swap: mxfp8 = te_llama_mxfp8.py, default = te_llama.py, hybrid = te_llama_hybrid.py")
    ap.add_argument("--batch", type=int, default=12, help="sequences per GPU (reference semantics: per-device batch)")
    ap.add_argument("--global-batch", type=int, default=0,
                    help="STRONG scaling: total sequences, split over the ranks (batch = global-batch / world); 0 = weak scaling with --batch per GPU")
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--layers", type=int, default=0, help="0 = the model's own depth")
    ap.add_argument("--graph", type=int, default=int(os.environ.get("LLMFP8_BENCH_GRAPH", "0")),
                    help="1: capture the whole forward+backward step in one CUDA graph and replay it (N=1 only; same kernels, no host launch cost)")
    ap.add_argument("--no-fuse-residual", action="store_true", help="run the plain h = h + f(h) layer form (separate elementwise passes)")
    ap.add_argument("--optimizer", default="none", choices=["none", "fused", "torch"],
                    help="NOT the headline (SURVEY.md section 8 leaves the optimizer outside the path): also step AdamW inside the timed region — "
                         "'fused' = llm_fp8_b200.FusedAdamW (its pass over the weights emits their MXFP8 copies, the forward skips its casts), "
                         "'torch' = torch.optim.AdamW")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--cpu-sample-tokens", type=int, default=1536)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "bf16": d["bf16_tflops"], "bf16_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "src": "measured"}
    return {"hbm": 6650.0, "bf16": 1590.0, "bf16_sustained": 1400.0, "src": "fallback"}


def workload_name(a, layers):
    opt = "" if getattr(a, "optimizer", "none") == "none" else f" + AdamW step ({a.optimizer})"
    return f"{a.model} linear path x{layers} layers, scenario={a.scenario}, batch {a.batch} x seq {a.seq} per GPU, fwd+bwd{opt}"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index: int):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(gpu_index)],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, pw) if p > 0.5 * max(pw)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_port_baseline(a, tokens: int):
    """Oracle ('port') fake-quant linear path of ONE decoder layer on the host cores, bounded sample."""
    import torch
    from llm_fp8_b200.layer import MODEL_SHAPES, linear_params_per_layer
    from oracle import fp8_oracle as fo
    H, I, heads, kv, hd, _ = MODEL_SHAPES[a.model]
    P = linear_params_per_layer(H, I, heads, kv, hd)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(7)
    qkv = heads * hd + 2 * kv * hd
    p = {"ln1": torch.ones(H), "ln2": torch.ones(H), "wqkv": torch.randn(qkv, H, generator=g) * 0.02, "wproj": torch.randn(H, heads * hd, generator=g) * 0.02,
         "w1": torch.randn(2 * I, H, generator=g) * 0.02, "b1": torch.zeros(2 * I), "w2": torch.randn(H, I, generator=g) * 0.02, "b2": torch.zeros(H)}
    for v in p.values():
        v.requires_grad_(True)
    x = torch.randn(tokens, H, generator=g).requires_grad_(True)
    mx = a.scenario == "mxfp8"
    bwd = fo.E4M3 if mx else fo.E5M2
    best = None
    for it in range(2):
        t0 = time.perf_counter()
        y = fo.fq_layer_linear_path(x, p, 1e-5, 1e-5, fo.E4M3, bwd, mx)
        y.backward(torch.ones_like(y))
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
        x.grad = None
        for v in p.values():
            v.grad = None
    return {"value": 6.0 * tokens * P / best / 1e12, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle fake-quant ({'MXFP8' if mx else 'per-tensor FP8'}) linear path, 1 layer of {a.model}, {tokens} tokens, fwd+bwd, fp32 matmul, best of 2"}


def reference_arm(a):
    """`--impl reference`: the reference's own CPU-capable code path for these layers — the plain HF bf16 modules it builds
    without --use_te (train_fp8.py:118-124) — on the host cores, one decoder layer's linear path per step on a bounded
    token sample.  (Its TE path has no CPU implementation and TE is not installable here: DESIGN.md.)"""
    import torch
    import torch.nn as nn
    from llm_fp8_b200.layer import MODEL_SHAPES, linear_params_per_layer
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    H, I, heads, kv, hd, L = MODEL_SHAPES[a.model]
    P = linear_params_per_layer(H, I, heads, kv, hd)
    tokens = a.cpu_sample_tokens
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    kind = "reference"
    try:
        from transformers import LlamaConfig
        from transformers.models.llama.modeling_llama import LlamaMLP, LlamaRMSNorm
        cfg = LlamaConfig(hidden_size=H, intermediate_size=I, num_attention_heads=heads, num_key_value_heads=kv, head_dim=hd, rms_norm_eps=1e-5,
                          num_hidden_layers=1, vocab_size=32)
        ln1, ln2, mlp = LlamaRMSNorm(H, eps=1e-5), LlamaRMSNorm(H, eps=1e-5), LlamaMLP(cfg)
        how = "HF transformers LlamaRMSNorm / nn.Linear q,k,v,o / LlamaMLP, bf16 (the reference's non-TE path, train_fp8.py:118-124)"
    except Exception as e:  # noqa: BLE001
        kind = "port"

        class _MLP(nn.Module):
            def __init__(self):
                super().__init__()
                self.gate_proj, self.up_proj, self.down_proj = nn.Linear(H, I, bias=False), nn.Linear(H, I, bias=False), nn.Linear(I, H, bias=False)

            def forward(self, x):
                return self.down_proj(nn.functional.silu(self.gate_proj(x)) * self.up_proj(x))
        ln1, ln2, mlp = nn.RMSNorm(H, eps=1e-5), nn.RMSNorm(H, eps=1e-5), _MLP()
        how = f"torch bf16 modules (transformers unavailable: {str(e)[:60]})"
    q, k, v, o = (nn.Linear(H, heads * hd, bias=False), nn.Linear(H, kv * hd, bias=False), nn.Linear(H, kv * hd, bias=False),
                  nn.Linear(heads * hd, H, bias=False))
    mods = nn.ModuleList([ln1, ln2, mlp, q, k, v, o]).to(torch.bfloat16)
    x = torch.randn(1, tokens, H).bfloat16().requires_grad_(True)

    def step():
        h = x
        y = ln1(h)
        qq, kk, vv = q(y), k(y), v(y)
        h = h + o(qq) + 0.0 * (kk.sum() + vv.sum())      # attention core = identity on the query slice (as in our arm); k, v still get gradients
        h = h + mlp(ln2(h))
        h.float().square().mean().backward()
        x.grad = None
        mods.zero_grad(set_to_none=True)

    for _ in range(max(1, min(a.warmup, 2))):
        step()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step()
    dt = (time.perf_counter() - t0) / a.steps
    val = 6.0 * tokens * P / dt / 1e12
    sample = f"{how}; 1 decoder layer's linear path of {a.model}, {tokens} tokens per step, fwd+bwd"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(a, a.layers or L), "cpu_sample": sample},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ our arm
def _traffic(table, name):
    """roofline.traffic = dram__bytes_read.sum + dram__bytes_write.sum per launch of that kernel, from the committed ncu --set full
    capture (profiles/ncu_traffic.json), or null; which capture it came from goes beside it."""
    t = (table or {}).get(name)
    if not t:
        return {"traffic": None}
    return {"traffic": t.get("dram_bytes_per_launch_avg"), "traffic_unit": "B per launch", "traffic_source": {k: v for k, v in t.items() if k != "dram_bytes_per_launch_avg"}}


def main():
    a = parse()
    if a.impl == "reference":
        reference_arm(a)
        return
    import torch
    import torch.distributed as dist

    import llm_fp8_b200 as te
    from llm_fp8_b200 import _lib, dp
    from llm_fp8_b200.layer import MODEL_SHAPES, LinearPathStack

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: llm_fp8_b200 has no CPU path (use --impl reference for the CPU arm)")
    if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
        os.environ["NCCL_DEBUG"] = "WARN"          # NCCL prints its version banner on stdout; stdout carries the JSON line only
    rank, world, local = dp.init_from_env()
    dev = torch.device("cuda", local)
    lib = _lib.load()
    pk = peaks()
    scaling = "weak"
    if a.global_batch > 0:                 # "also report strong scaling (global M fixed, M_local = M / n)"
        if a.global_batch % world:
            raise SystemExit(f"--global-batch {a.global_batch} is not divisible by {world} ranks")
        a.batch, scaling = a.global_batch // world, "strong"

    model = LinearPathStack(a.model, a.scenario, a.layers or None, device=dev, fuse_residual=False if a.no_fuse_residual else None)
    L = model.n_layers
    H = model.H
    M = a.batch * a.seq
    flops_step = model.linear_flops(M)                 # per rank
    # LLMFP8_NVLS_REDUCE=1 (opt-in, unmeasured): weight gradients written by the wgrad GEMMs into a symmetric arena and summed in the
    # NVSwitch fabric by our own kernel (csrc/nvls_reduce.cu) instead of NCCL
    arena = dp.NvlsGradArena(dp.gemm_weight_buffers(model)) if world > 1 and os.environ.get("LLMFP8_NVLS_REDUCE", "0") == "1" else None
    reducer = dp.GradAllReducer(model.parameters(), nvls=arena) if world > 1 else None

    g = torch.Generator(device=dev).manual_seed(1234 + rank)           # 42 + rank seeding in the reference: train_multi_gpu.py:1003
    x_dev = torch.randn(a.batch, a.seq, H, device=dev, generator=g).bfloat16()
    x_host = x_dev.cpu().pin_memory()
    # e2e: the loss of every step is read back into pinned memory.  The read is pipelined one step deep, the way a training loop logs
    # its loss without stalling the launch queue: step i's copy is asynchronous, and the host waits for it (and looks at the value)
    # only after it has issued step i+1.  The timed region ends with a full synchronize, so every read is inside it.
    loss_host = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_evt = [None, None]
    e2e_state = {"i": 0, "last": float("nan")}

    def read_loss(loss_dev):
        i = e2e_state["i"]
        loss_host[i & 1].copy_(loss_dev.detach().reshape(1), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        loss_evt[i & 1] = ev
        prev = loss_evt[(i + 1) & 1]
        if prev is not None:
            prev.synchronize()
            e2e_state["last"] = float(loss_host[(i + 1) & 1][0])
        e2e_state["i"] = i + 1

    def step(x_in):
        x = x_in.detach().requires_grad_(True)
        y = model(x)
        loss = y.float().square().mean()
        (loss * reducer.loss_scale if reducer is not None and reducer.prescaled else loss).backward()
        if reducer is not None:
            reducer.finish()
        return loss

    optimizer = None
    if a.optimizer == "fused":
        optimizer = te.FusedAdamW(model, lr=1e-5, betas=(0.9, 0.95), weight_decay=0.01)
    elif a.optimizer == "torch":
        optimizer = torch.optim.AdamW(model.parameters(), lr=1e-5, betas=(0.9, 0.95), weight_decay=0.01)

    def zero():
        if optimizer is not None:
            optimizer.step()                 # inside the timed region, after the gradient reduction
        for p in model.parameters():
            p.grad = None

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = lib.llmfp8_launch_count()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        sync_all()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms / steps, (lib.llmfp8_launch_count() - n0) // steps

    def resident_step():
        step(x_dev)
        zero()

    # e2e input staging: every step's tokens come from pinned host memory; the copy runs on a side stream into one of two device
    # buffers so that it overlaps the previous step's compute (the host runs about one step ahead of the GPU)
    x_stage = [torch.empty_like(x_dev), torch.empty_like(x_dev)]
    x_free = [None, None]                      # event: the step that last read this buffer has finished
    copy_stream = torch.cuda.Stream(device=dev)
    stage_state = {"j": 0}

    def stage_input():
        j = stage_state["j"]
        stage_state["j"] = j + 1
        buf = x_stage[j & 1]
        with torch.cuda.stream(copy_stream):
            if x_free[j & 1] is not None:
                copy_stream.wait_event(x_free[j & 1])
            buf.copy_(x_host, non_blocking=True)                       # H2D from pinned memory, inside the timed region
            ready = torch.cuda.Event()
            ready.record(copy_stream)
        torch.cuda.current_stream().wait_event(ready)
        return buf, j & 1

    def e2e_step():
        buf, slot = stage_input()
        loss = step(buf)
        done = torch.cuda.Event()
        done.record()
        x_free[slot] = done
        read_loss(loss)                                                # D2H read of the step's result (pinned, pipelined one step)
        zero()

    # warm-up (also fills the delayed-scaling history enough for real scales)
    for i in range(max(a.warmup, 3)):
        resident_step()
        if arena is not None and i == 0:
            arena.check()                  # fail after one step, not after hundreds of timed-out rank meetings

    graphed = False
    eager_resident_step, eager_e2e_step = resident_step, e2e_step
    # world > 1 only as a probe (LLMFP8_GRAPH_DP=1, tools/gpu_round2_scale.sh): NCCL's / our gradient reduces are captured with the step, so
    # every rank replays the same launch sequence with no host jitter between ranks — unmeasured
    if a.graph and (world == 1 or os.environ.get("LLMFP8_GRAPH_DP", "0") == "1"):
        # the same step, captured once into a CUDA graph (every kernel of forward + backward, the weight-gradient writes included)
        # and replayed: the host issues one launch per step instead of ~1500
        try:
            n0 = lib.llmfp8_launch_count()
            x_static = x_dev.clone()
            zero()
            cg = torch.cuda.CUDAGraph()
            with torch.cuda.graph(cg):
                loss_static = step(x_static)
            captured_launches = int(lib.llmfp8_launch_count() - n0)
            torch.cuda.synchronize()

            def resident_step():            # noqa: F811
                cg.replay()

            def e2e_step():                 # noqa: F811
                x_static.copy_(x_host, non_blocking=True)
                cg.replay()
                read_loss(loss_static)
            graphed = True
            for _ in range(2):
                resident_step()
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            print(f"[bench] CUDA-graph capture failed, staying eager: {type(e).__name__}: {str(e)[:300]}", file=sys.stderr, flush=True)
            torch.cuda.synchronize()
            graphed = False
            resident_step, e2e_step = eager_resident_step, eager_e2e_step
    sampler = ClockSampler(local) if rank == 0 else None
    ms_step, launches = timed(resident_step, a.steps)
    if graphed:
        launches = captured_launches
    e2e_step()
    ms_e2e, _ = timed(e2e_step, a.steps)
    clocks = sampler.stop() if sampler else None

    value = world * flops_step / (ms_step * 1e-3) / 1e12
    e2e_value = world * flops_step / (ms_e2e * 1e-3) / 1e12

    # ---- where a step's time goes on the main stream: forward | backward (its GEMMs compete with the gradient all-reduces for SMs,
    # HBM and power at N > 1) | the wait for the reductions still outstanding when backward has returned (the exposed tail).  Three
    # untimed steps, events on the launching stream, max over ranks — the split the scaling loss is attributed with.
    def timeline_step():
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        x = x_dev.detach().requires_grad_(True)
        ev[0].record()
        y = model(x)
        loss = y.float().square().mean()
        ev[1].record()
        (loss * reducer.loss_scale if reducer is not None and reducer.prescaled else loss).backward()
        ev[2].record()
        if reducer is not None:
            reducer.finish()
        ev[3].record()
        zero()
        return ev
    sync_all()
    tl = [timeline_step() for _ in range(3)]
    sync_all()
    parts = [sum(e[i].elapsed_time(e[i + 1]) for e in tl) / len(tl) for i in range(3)]
    if world > 1:
        t = torch.tensor(parts, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        parts = t.tolist()
    timeline = {"forward_ms": parts[0], "backward_ms": parts[1], "reduce_tail_ms": parts[2], "how": "CUDA events on the main stream, 3 steps, max over ranks"}

    # ---- roofline leg: every C-ABI launch of one more step bracketed by CUDA events on the launching stream
    roof, roof_q, shares = None, None, None
    if not a.no_roofline and rank != 0:
        resident_step()                      # every rank runs the step (its gradient / amax all-reduces must match rank 0's)
    if not a.no_roofline and rank == 0:
        # per-launch timing needs the launches one after the other: the wgrad side stream (functional.WGRAD_SIDE_STREAM) is switched off
        # for this ONE untimed step, otherwise a kernel's event pair also spans the time it waits for SMs held by the overlapping GEMM
        from llm_fp8_b200 import functional as _fn
        _side, _fn.WGRAD_SIDE_STREAM = _fn.WGRAD_SIDE_STREAM, False
        try:
            _lib.profile_start()
            eager_resident_step()
            recs = _lib.profile_stop()
        finally:
            _fn.WGRAD_SIDE_STREAM = _side
        tot = sum(r[3] for r in recs) or 1e-9
        by = {}
        canon = {"llmfp8_fp8_gemm_res": "llmfp8_fp8_gemm", "llmfp8_mxfp8_gemm_res": "llmfp8_mxfp8_gemm", "llmfp8_mx_quantize_ld": "llmfp8_mx_quantize",
                 "llmfp8_quant_cast_amax_2d": "llmfp8_quant_cast_amax", "llmfp8_rmsnorm_bwd_add": "llmfp8_rmsnorm_bwd"}   # same kernels, extended entry points
        for n, u, w, ms in recs:
            n = canon.get(n, n)
            d = by.setdefault(n, {"unit": u, "work": 0.0, "ms": 0.0, "launches": 0})
            d["work"] += w; d["ms"] += ms; d["launches"] += 1
        shares = {n.replace("llmfp8_", ""): round(d["ms"] / tot, 4) for n, d in sorted(by.items(), key=lambda kv: -kv[1]["ms"])}
        # what the step spends OUTSIDE this library's kernels (torch ops of the stand-in loss and attention slice, launch gaps): the timed step
        # minus the sum of the per-launch times of the serialised profile step
        shares["_library_kernels_ms"] = round(tot, 3)
        shares["_step_ms_minus_library_kernels_ms"] = round(ms_step - tot, 3)
        gname = "llmfp8_mxfp8_gemm" if a.scenario == "mxfp8" else "llmfp8_fp8_gemm"
        traffic = None
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp))
        if gname in by:
            d = by[gname]
            ach = d["work"] / (d["ms"] * 1e-3) / 1e12
            peak = 2.0 * pk["bf16_sustained"]
            roof = {"bound": "tensor", "kernel": gname.replace("llmfp8_", "") + " (fp8_gemm_kernel: tcgen05.mma, all fprop/dgrad/wgrad launches of the step)",
                    "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "peak_note": f"2 x {pk['src']} sustained bf16 cuBLAS TFLOP/s (MEASURED_PEAKS.json has no FP8 figure; FP8 dense = 2 x bf16 on this part)",
                    "frac_of_nominal_fp8_4500": ach / FP8_NOMINAL_TFLOPS, "frac_of_measured_bf16": ach / pk["bf16_sustained"],
                    "power_cap_note": "the board is at its 1000 W cap: the twelve GEMMs of a layer back to back, alone, sustain 2288 TF (mxfp8) / 2554 TF "
                                      "(per-tensor) at Llama-3.1-8B shapes (profiles/r02/k_micro_sustained_gemms.jsonl) - the in-step rate is at that ceiling",
                    "launches": d["launches"], "avg_launch_ms": d["ms"] / d["launches"], "share_of_step": d["ms"] / tot,
                    **_traffic(traffic, gname)}
        qn = max((n for n in by if by[n]["unit"] == "byte"), key=lambda n: by[n]["ms"], default=None)
        if qn:
            d = by[qn]
            ach = d["work"] / (d["ms"] * 1e-3) / 1e9
            roof_q = {"bound": "hbm", "kernel": qn.replace("llmfp8_", ""), "achieved": ach, "peak": pk["hbm"], "unit": "GB/s", "frac": ach / pk["hbm"],
                      "peak_note": f"{pk['src']} copy bandwidth", "launches": d["launches"], "avg_launch_ms": d["ms"] / d["launches"],
                      "share_of_step": d["ms"] / tot, **_traffic(traffic, qn),
                      # every memory-bound kernel of the step (algorithmic bytes / measured launch time): the largest one by time alone hides the mix —
                      # since round 2 the big activation tensors go through the fused SwiGLU / dSwiGLU + quantiser kernels, and what is left under
                      # `mx_quantize` is the weights (0.8 of HBM) and the [M, H]-sized tensors (0.6)
                      "by_kernel": {n.replace("llmfp8_", ""): {"GBps": round(v["work"] / (v["ms"] * 1e-3) / 1e9, 1), "frac": round(v["work"] / (v["ms"] * 1e-3) / 1e9 / pk["hbm"], 3),
                                                               "launches": v["launches"], "share_of_step": round(v["ms"] / tot, 4)}
                                    for n, v in sorted(by.items(), key=lambda kv: -kv[1]["ms"]) if v["unit"] == "byte" and v["ms"] > 0}}

    if arena is not None:
        arena.check()                      # a rank meeting that timed out invalidates the run: raise instead of printing a number
        if reducer.nvls_launches == 0:
            raise RuntimeError("LLMFP8_NVLS_REDUCE=1 but no gradient went through the in-switch reduce")
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_port_baseline(a, a.cpu_sample_tokens)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "fp8_e4m3" + ("+e8m0" if a.scenario == "mxfp8" else "/e5m2") + " (fp32 accumulate, bf16 out)",
            "data": "synthetic",
            "config": {"workload": workload_name(a, L), "tokens_per_gpu": M, "global_tokens": M * world, "layers": L,
                       "linear_params_per_layer": model.params_per_layer, "flops_per_step_per_gpu": flops_step,
                       "parallelism": f"dp{world}", "grad_reduce": (("in-switch sum (multimem.ld_reduce / st, own kernel) of 1/N-prescaled gradients" if arena is not None else "sum of 1/N-prescaled gradients") if reducer is not None and reducer.prescaled else "avg") if world > 1 else None,
                       "launch": "cuda-graph replay" if graphed else "eager stream launches",
                       "residual_adds": (("fused: forward add after the GEMM, gradient in the RMSNorm backward" if a.scenario == "mxfp8"
                                         else "fused: GEMM epilogue / RMSNorm backward") if model.layers[0].fuse_residual else "separate elementwise passes"), "l2": "inputs larger than L2 (per-step working set: weights + activations of the stack >> 126 MB)",
                       "frac_of_nominal_fp8_peak": value / world / FP8_NOMINAL_TFLOPS,
                       "experiments": {k: os.environ[k] for k in ("LLMFP8_GEMM_DEBUG", "LLMFP8_MX_EPILOGUE_RESIDUAL", "LLMFP8_MX_FUSED_ACT",
                                                                  "LLMFP8_WGRAD_STREAM", "LLMFP8_FUSE_RESIDUAL", "LLMFP8_JOIN_SPLIT_WEIGHTS", "LLMFP8_PDL", "LLMFP8_GEMM_FORCE_CG", "LLMFP8_GEMM_FORCE_BN", "NCCL_CGA_CLUSTER_SIZE", "NCCL_MAX_CTAS", "LLMFP8_GRAD_BUCKET_MB", "LLMFP8_NVLS_REDUCE", "LLMFP8_NVLS_CTAS", "LLMFP8_NVLS_CHUNK_MB", "LLMFP8_GRAPH_DP", "LLMFP8_NVLS_AMAX") if k in os.environ}},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": x_host.numel() * 2, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches), "clocks": clocks, "timeline": timeline, "roofline": roof, "roofline_quantize": roof_q, "kernel_time_shares": shares,
            "cpu_baseline": cpu,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
