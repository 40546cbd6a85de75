#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_mx_modules.py -x -q -m gpu -p no:cacheprovider > gpurun_out/r02h_pytest.log 2>&1
echo "pytest exit $?  $(tail -1 gpurun_out/r02h_pytest.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02h_pytest.log | head
timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02h_bench_mxfp8.json 2> gpurun_out/r02h_bench_mxfp8.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02h_bench_mxfp8.json").read().strip().splitlines()[-1])
print("bench", round(d["value"], 1), round(d["ms_per_step"], 2), d["clocks"]["sm_mhz"], d["timeline"], d["kernel_time_shares"], d["roofline"]["achieved"], d["roofline_quantize"])
PY
