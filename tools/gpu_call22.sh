#!/bin/bash
# two-CTAs-per-SM activation + MX quantiser kernels (mx_act_quantize2.cu): parity, microbench and in-step A/B against the one-CTA form (LLMFP8_ACT_V1=1)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_quantize.py tests/test_gpu_mx_modules.py -x -q -m gpu -k "swiglu or act or mlp" > gpurun_out/r02o_pytest_act2.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r02o_pytest_act2.log
timeout 200 python tools/microbench.py --only quant --kfilter "K3m" --out gpurun_out/r02o_micro_act2.jsonl > gpurun_out/r02o_micro_act2.log 2>&1; echo "micro v2 exit $?"
LLMFP8_ACT_V1=1 timeout 200 python tools/microbench.py --only quant --kfilter "K3m" --tag v1 --out gpurun_out/r02o_micro_act1.jsonl > gpurun_out/r02o_micro_act1.log 2>&1; echo "micro v1 exit $?"
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02o_bench_act2.json 2> gpurun_out/r02o_bench_act2.err; echo "bench v2 exit $?"
LLMFP8_ACT_V1=1 $B > gpurun_out/r02o_bench_act1.json 2> gpurun_out/r02o_bench_act1.err; echo "bench v1 exit $?"
$B > gpurun_out/r02o_bench_act2_b.json 2> gpurun_out/r02o_bench_act2_b.err; echo "bench v2 (again) exit $?"
python - <<'PY'
import json
for f in ("gpurun_out/r02o_micro_act2.jsonl", "gpurun_out/r02o_micro_act1.jsonl"):
    try:
        for l in open(f):
            d = json.loads(l)
            print(f"{f[-11:-6]} {d['kernel'][:70]:70s} {d['ms']*1000:7.1f} us  {d.get('frac_hbm', 0):.3f} of HBM  sm {d.get('clocks', {}).get('sm_mhz')}")
    except Exception as e:
        print(f, "unreadable", e)
for t in ("act2", "act1", "act2_b"):
    try:
        d = json.loads(open(f"gpurun_out/r02o_bench_{t}.json").read().strip().splitlines()[-1])
        print(f"{t:7s} {d['ms_per_step']:7.2f} ms/step  value {d['value']:7.1f}  e2e {d['e2e']['value']:7.1f}  shares {d['kernel_time_shares']}")
    except Exception as e:
        print(t, "unreadable", e)
PY
# ncu of the two new kernels at the 8B shape (one launch each, after the microbench's warm-up launches)
timeout 300 ncu --set full --clock-control none -k regex:"quantize2_kernel" -s 6 -c 2 -f -o gpurun_out/prof_act2 python tools/microbench.py --only quant --kfilter "(one pass)" --out gpurun_out/r02o_micro_ncu_scratch.jsonl > gpurun_out/r02o_ncu_act2.log 2>&1; echo "ncu exit $?"
ncu -i gpurun_out/prof_act2.ncu-rep --page raw --csv > gpurun_out/prof_act2_raw.csv 2>/dev/null; rm -f gpurun_out/prof_act2.ncu-rep gpurun_out/r02o_micro_ncu_scratch.jsonl
