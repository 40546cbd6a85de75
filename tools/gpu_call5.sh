#!/bin/bash
# round 2, fifth call (1 GPU): ordered first TMEM loads in the OVL epilogue (debug 256 = previous behaviour)
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_mx_modules.py -q -m gpu -p no:cacheprovider > gpurun_out/r02e_pytest.log 2>&1
echo "pytest exit $?  $(tail -1 gpurun_out/r02e_pytest.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02e_pytest.log | head
timeout 200 python tools/gemm_ab.py --modes 0,256 --rounds 3 > gpurun_out/r02e_gemm_ab.txt 2>&1; echo "gemm A/B exit $?"
B="timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02e_bench_mxfp8.json 2> gpurun_out/r02e_bench_mxfp8.err; echo "bench exit $?"
LLMFP8_GEMM_DEBUG=256 $B > gpurun_out/r02e_bench_mxfp8_unordered.json 2> gpurun_out/r02e_bench_mxfp8_unordered.err; echo "bench unordered exit $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_kernel -s 5 -c 2 -o gpurun_out/r02e_probe_gemm python tools/gemm_probe.py > gpurun_out/r02e_ncu_probe.log 2>&1
echo "ncu probe exit $?"
python - <<'PY'
import json
for t in ("mxfp8", "mxfp8_unordered"):
    d = json.loads(open(f"gpurun_out/r02e_bench_{t}.json").read().strip().splitlines()[-1])
    print("bench", t, round(d["value"], 1), round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"], d.get("timeline"))
PY
grep "G.m" gpurun_out/r02e_gemm_ab.txt
# first hardware run of the bf16 GEMM + lm_head epilogues (last: a hang here must not cost the measurements above)
timeout 300 python -m pytest tests/test_gpu_lm_head.py -q -m gpu -p no:cacheprovider -x > gpurun_out/r02e_pytest_lm_head.log 2>&1
echo "pytest lm_head exit $?  $(tail -1 gpurun_out/r02e_pytest_lm_head.log)"; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02e_pytest_lm_head.log | head -20
