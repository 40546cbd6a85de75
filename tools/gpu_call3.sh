#!/bin/bash
# round 2, third call (1 GPU): suite on the new defaults, bench lines, MX scale-handling upper bounds, ncu of the fused activation kernels
mkdir -p gpurun_out
timeout 500 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/r02c_pytest_gpu.log 2>&1
echo "pytest gpu exit $?  $(tail -1 gpurun_out/r02c_pytest_gpu.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02c_pytest_gpu.log | head -20
B="timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02c_bench_mxfp8.json 2> gpurun_out/r02c_bench_mxfp8.err; echo "bench mxfp8 exit $?"
LLMFP8_PDL=1 $B > gpurun_out/r02c_bench_mxfp8_pdl.json 2> gpurun_out/r02c_bench_mxfp8_pdl.err; echo "bench mxfp8 pdl exit $?"
LLMFP8_MX_FUSED_ACT=0 LLMFP8_WGRAD_STREAM=0 LLMFP8_MX_EPILOGUE_RESIDUAL=0 $B > gpurun_out/r02c_bench_mxfp8_olddefaults.json 2> gpurun_out/r02c_bench_mxfp8_olddefaults.err; echo "bench mxfp8 old defaults exit $?"
$B --scenario default > gpurun_out/r02c_bench_default.json 2> gpurun_out/r02c_bench_default.err; echo "bench default exit $?"
timeout 200 python tools/gemm_ab.py --modes 0,1,2,3 --rounds 3 > gpurun_out/r02c_gemm_ab_sf.txt 2>&1; echo "gemm A/B exit $?"
CMD="python tools/microbench.py --model llama8b --only quant --kfilter K3m --experimental --out gpurun_out/r02c_micro_tmp.jsonl"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"mx_swiglu_quantize_kernel|mx_dswiglu_quantize_kernel" -c 4 -o gpurun_out/r02c_prof_act $CMD > gpurun_out/r02c_ncu_act.log 2>&1
echo "ncu act exit $?"
for f in gpurun_out/r02c_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1].split('r02c_bench_')[1]:40s} value {d['value']:8.1f} {d['unit']}  {d['ms_per_step']:7.2f} ms  e2e {d['e2e']['value']:8.1f}  sm {d['clocks']['sm_mhz']}")
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
cat gpurun_out/r02c_gemm_ab_sf.txt | tail -25
