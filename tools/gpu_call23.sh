#!/bin/bash
# K5 (rmsnorm_bwd) streamed through a cp.async.bulk ring vs the register-prefetch kernel (LLMFP8_K5_V1=1); the whole GPU suite on the tree with the
# two-CTAs-per-SM activation kernels as the only form; bench; ncu of K5
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02p_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r02p_pytest_gpu.log
timeout 200 python tools/microbench.py --only quant --kfilter "K5" --out gpurun_out/r02p_micro_k5_ring.jsonl > gpurun_out/r02p_micro_k5_ring.log 2>&1; echo "micro ring exit $?"
LLMFP8_K5_V1=1 timeout 200 python tools/microbench.py --only quant --kfilter "K5" --tag v1 --out gpurun_out/r02p_micro_k5_v1.jsonl > gpurun_out/r02p_micro_k5_v1.log 2>&1; echo "micro v1 exit $?"
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02p_bench_ring.json 2> gpurun_out/r02p_bench_ring.err; echo "bench ring exit $?"
LLMFP8_K5_V1=1 $B > gpurun_out/r02p_bench_k5v1.json 2> gpurun_out/r02p_bench_k5v1.err; echo "bench v1 exit $?"
$B --scenario default > gpurun_out/r02p_bench_default_ring.json 2> gpurun_out/r02p_bench_default_ring.err; echo "bench default ring exit $?"
LLMFP8_K5_V1=1 $B --scenario default > gpurun_out/r02p_bench_default_k5v1.json 2> gpurun_out/r02p_bench_default_k5v1.err; echo "bench default v1 exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r02p_micro_k5_ring.jsonl", "gpurun_out/r02p_micro_k5_v1.jsonl"):
    try:
        for l in open(f):
            d = json.loads(l)
            print(f"{f[-12:-6]} {d['kernel'][:60]:60s} {d['ms']*1000:7.1f} us  {d.get('frac_hbm', 0):.3f} of HBM  sm {d.get('clocks', {}).get('sm_mhz')}")
    except Exception as e:
        print(f, "unreadable", e)
for t in ("ring", "k5v1", "default_ring", "default_k5v1"):
    try:
        d = json.loads(open(f"gpurun_out/r02p_bench_{t}.json").read().strip().splitlines()[-1])
        print(f"{t:13s} {d['ms_per_step']:7.2f} ms/step  value {d['value']:7.1f}  e2e {d['e2e']['value']:7.1f}  shares {d['kernel_time_shares']}")
    except Exception as e:
        print(t, "unreadable", e)
PY
timeout 300 ncu --set full --clock-control none -k regex:"rmsnorm_bwd" -s 4 -c 2 -f -o gpurun_out/prof_k5 python tools/microbench.py --only quant --kfilter "K5" --out gpurun_out/r02p_scratch.jsonl > gpurun_out/r02p_ncu_k5.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/prof_k5.ncu-rep --page raw --csv > gpurun_out/prof_k5_raw.csv 2>/dev/null; rm -f gpurun_out/prof_k5.ncu-rep gpurun_out/r02p_scratch.jsonl
