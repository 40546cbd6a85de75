#!/bin/bash
# round 2 (1 GPU): 448-column MX super-tiles
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu -p no:cacheprovider -k "super_tiles" > gpurun_out/r02i_pytest_st.log 2>&1
echo "pytest super-tiles exit $?  $(tail -1 gpurun_out/r02i_pytest_st.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02i_pytest_st.log | head -5
timeout 500 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_mx_modules.py tests/test_gpu_modules.py -q -m gpu -p no:cacheprovider > gpurun_out/r02i_pytest.log 2>&1
echo "pytest exit $?  $(tail -1 gpurun_out/r02i_pytest.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02i_pytest.log | head
# 256 = no super-tiles (overlapped accumulators), 0 = super-tiles where B is K-major (fprop), 512 = also dgrad / wgrad under CTA pairs
timeout 300 python tools/gemm_ab.py --modes 256,0,512 --rounds 3 > gpurun_out/r02i_gemm_ab.txt 2>&1; echo "gemm A/B exit $?"; grep "G.m" gpurun_out/r02i_gemm_ab.txt
B="timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
LLMFP8_GEMM_DEBUG=256 $B > gpurun_out/r02i_bench_mxfp8_nost.json 2> gpurun_out/r02i_bench_mxfp8_nost.err; echo "bench no-st exit $?"
$B > gpurun_out/r02i_bench_mxfp8.json 2> gpurun_out/r02i_bench_mxfp8.err; echo "bench exit $?"
LLMFP8_GEMM_DEBUG=512 $B > gpurun_out/r02i_bench_mxfp8_stall.json 2> gpurun_out/r02i_bench_mxfp8_stall.err; echo "bench st-all exit $?"
python - <<'PY'
import json
for t in ("mxfp8_nost", "mxfp8", "mxfp8_stall"):
    try:
        d = json.loads(open(f"gpurun_out/r02i_bench_{t}.json").read().strip().splitlines()[-1])
        print("bench", t, round(d["value"], 1), round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"], round(d["roofline"]["achieved"]) if d.get("roofline") else None)
    except Exception as e:
        print(t, "unreadable", e)
PY
python tools/gemm_trace.py mx > gpurun_out/r02i_gemm_trace.txt 2>&1; head -10 gpurun_out/r02i_gemm_trace.txt
