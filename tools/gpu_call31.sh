#!/bin/bash
# dgamma column reduce on 32 partial-row lanes (was 8): K5 tests, K5 microbench, whole suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02x_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02x_pytest_gpu.log
timeout 200 python tools/microbench.py --only quant --kfilter "K5" --out gpurun_out/r02x_micro_k5.jsonl > gpurun_out/r02x_micro_k5.log 2>&1; echo "micro exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r02x_micro_k5.jsonl"):
    d = json.loads(l); print(f"{d['kernel'][:60]:60s} {d['ms']*1000:7.1f} us {d.get('frac_hbm', 0):.3f}")
PY
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"colreduce|rmsnorm_bwd" -c 12 python tools/microbench.py --only quant --kfilter "K5" --out gpurun_out/r02x_scratch.jsonl 2>&1 | grep -E "colreduce|rmsnorm_bwd|gpu__time" | head -24; rm -f gpurun_out/r02x_scratch.jsonl
