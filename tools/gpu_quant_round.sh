#!/bin/bash
# quantise-kernel iteration: bit-exact tests, then the quantise microbench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_quantize.py tests/test_gpu_modules.py tests/test_golden.py -q -m gpu -p no:cacheprovider --timeout 300 > gpurun_out/test_gpu_quantize.log 2>&1
echo "quant tests exit $?"; tail -8 gpurun_out/test_gpu_quantize.log
timeout 300 python tools/loss_band.py 2>&1 | grep -v Warning | tail -8
timeout 600 python tools/microbench.py --model llama8b --only quant --out gpurun_out/micro_quant.jsonl > gpurun_out/micro_quant.log 2>&1
echo "micro exit $?"
for v in 1 2 3; do
  LLMFP8_TUNE_DSWIGLU=$v timeout 300 python tools/microbench.py --model llama8b --only quant --kfilter dswiglu --tag dsw$v --out gpurun_out/micro_quant.jsonl > gpurun_out/micro_dsw$v.log 2>&1
done
python - <<'PY'
import json
for l in open("gpurun_out/micro_quant.jsonl"):
    d=json.loads(l); print(f'{d.get("tag",""):5s}{d["kernel"]:60s} {d["ms"]*1e3:8.1f} us {d["GBps"]:7.0f} GB/s {d["frac_hbm"]:.3f}')
PY
