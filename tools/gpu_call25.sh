#!/bin/bash
# the cleaned tree (ring kernels only): whole GPU suite, smoke, quantiser microbench, headline bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02r_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02r_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02r_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r02r_smoke.log
timeout 300 python tools/microbench.py --only quant --out gpurun_out/r02r_micro_quant.jsonl > gpurun_out/r02r_micro_quant.log 2>&1; echo "micro exit $?"
timeout 400 python bench.py > gpurun_out/r02r_bench_mxfp8.json 2> gpurun_out/r02r_bench_mxfp8.err; echo "bench exit $?"
timeout 300 python bench.py --scenario default --no-cpu-baseline > gpurun_out/r02r_bench_default.json 2> gpurun_out/r02r_bench_default.err; echo "bench default exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r02r_micro_quant.jsonl"):
    d = json.loads(l); print(f"{d['kernel'][:72]:72s} {d['ms']*1000:7.1f} us {d.get('frac_hbm', 0):.3f}")
for t in ("mxfp8", "default"):
    d = json.loads(open(f"gpurun_out/r02r_bench_{t}.json").read().strip().splitlines()[-1])
    print(t, d["value"], d["ms_per_step"], d["e2e"], d.get("roofline"), d.get("roofline_quantize", {}).get("frac"), d["clocks"])
PY
