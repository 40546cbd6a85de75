#!/bin/bash
# HEAD as the driver will run it: smoke, bench (defaults), reference arm
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02y_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r02y_smoke.log
timeout 400 python bench.py > gpurun_out/r02y_bench_mxfp8.json 2> gpurun_out/r02y_bench_mxfp8.err; echo "bench exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02y_bench_mxfp8.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"], d["gpu_launches"], d["roofline"]["frac"], d["roofline_quantize"], d["cpu_baseline"], d["clocks"], d["kernel_time_shares"])
PY
