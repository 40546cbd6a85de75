#!/bin/bash
# round 2, fourth call (1 GPU): new quantiser test, quantiser microbench (with clock records), GEMM A/B (5-stage per-tensor), ncu probe of the fc1 GEMM
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_quantize.py tests/test_gpu_mx_modules.py -q -m gpu -p no:cacheprovider > gpurun_out/r02d_pytest.log 2>&1
echo "pytest exit $?  $(tail -1 gpurun_out/r02d_pytest.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02d_pytest.log | head
timeout 300 python tools/microbench.py --model llama8b --only quant --out gpurun_out/r02d_micro_quant.jsonl > gpurun_out/r02d_micro_quant.log 2>&1; echo "microbench quant exit $?"
timeout 200 python tools/gemm_ab.py --modes 0,64,128 --rounds 3 > gpurun_out/r02d_gemm_ab_stages.txt 2>&1; echo "gemm A/B exit $?"
timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_mxfp8.json 2> gpurun_out/r02d_bench_mxfp8.err; echo "bench exit $?"
LLMFP8_GEMM_DEBUG=128 timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02d_bench_mxfp8_backoff.json 2> gpurun_out/r02d_bench_mxfp8_backoff.err; echo "bench backoff exit $?"
python tools/gemm_probe.py > gpurun_out/r02d_probe_plain.log 2>&1; echo "probe plain exit $?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_kernel -s 1 -c 7 -o gpurun_out/r02d_probe_gemm python tools/gemm_probe.py > gpurun_out/r02d_ncu_probe.log 2>&1
echo "ncu probe exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r02d_micro_quant.jsonl"):
    d = json.loads(l)
    print(f"{d['kernel'][:62]:62s} {d['ms']*1e3:8.1f} us  {d['GBps']:7.0f} GB/s  {d['frac_hbm']:.2f}  sm {d['clocks'].get('sm_mhz')}")
d = json.loads(open("gpurun_out/r02d_bench_mxfp8.json").read().strip().splitlines()[-1])
print("bench mxfp8", d["value"], d["ms_per_step"], d["clocks"])
PY
cat gpurun_out/r02d_gemm_ab_stages.txt | tail -22
