#!/bin/bash
# K1 / K1r / K2 / K2b through shared-memory rings (cp.async.bulk / TMA) vs the register-staged kernels (LLMFP8_QUANT_V1=1), K5 with the deeper ring
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02q_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r02q_pytest_gpu.log
timeout 300 python tools/microbench.py --only quant --out gpurun_out/r02q_micro_ring.jsonl > gpurun_out/r02q_micro_ring.log 2>&1; echo "micro ring exit $?"
LLMFP8_QUANT_V1=1 LLMFP8_K5_V1=1 timeout 300 python tools/microbench.py --only quant --tag v1 --out gpurun_out/r02q_micro_v1.jsonl > gpurun_out/r02q_micro_v1.log 2>&1; echo "micro v1 exit $?"
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B --scenario default > gpurun_out/r02q_bench_default_ring.json 2> gpurun_out/r02q_bench_default_ring.err; echo "bench default ring exit $?"
LLMFP8_QUANT_V1=1 LLMFP8_K5_V1=1 $B --scenario default > gpurun_out/r02q_bench_default_v1.json 2> gpurun_out/r02q_bench_default_v1.err; echo "bench default v1 exit $?"
$B > gpurun_out/r02q_bench_mxfp8_ring.json 2> gpurun_out/r02q_bench_mxfp8_ring.err; echo "bench mxfp8 ring exit $?"
LLMFP8_QUANT_V1=1 LLMFP8_K5_V1=1 $B > gpurun_out/r02q_bench_mxfp8_v1.json 2> gpurun_out/r02q_bench_mxfp8_v1.err; echo "bench mxfp8 v1 exit $?"
python - <<'PY'
import json
rows = {}
for tag, f in (("ring", "gpurun_out/r02q_micro_ring.jsonl"), ("v1", "gpurun_out/r02q_micro_v1.jsonl")):
    try:
        for l in open(f):
            d = json.loads(l)
            rows.setdefault(d["kernel"], {})[tag] = d
    except Exception as e:
        print(f, "unreadable", e)
for k, v in rows.items():
    a, b = v.get("v1"), v.get("ring")
    if a and b:
        print(f"{k[:72]:72s} v1 {a['ms']*1000:7.1f} us {a.get('frac_hbm', 0):.3f}   ring {b['ms']*1000:7.1f} us {b.get('frac_hbm', 0):.3f}")
for t in ("default_ring", "default_v1", "mxfp8_ring", "mxfp8_v1"):
    try:
        d = json.loads(open(f"gpurun_out/r02q_bench_{t}.json").read().strip().splitlines()[-1])
        print(f"{t:13s} {d['ms_per_step']:7.2f} ms/step  value {d['value']:7.1f}  e2e {d['e2e']['value']:7.1f}  shares {d['kernel_time_shares']}")
    except Exception as e:
        print(t, "unreadable", e)
PY
