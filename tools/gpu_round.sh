#!/bin/bash
# One GPU-box pass: per-file GPU tests (separate processes so one CUDA fault cannot poison the rest), then microbench.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in test_gpu_quantize test_gpu_gemm test_gpu_modules; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu -p no:cacheprovider --timeout 120 -x > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
  tail -5 gpurun_out/$f.log
done
timeout 600 python tools/microbench.py --model llama8b --out gpurun_out/micro.jsonl > gpurun_out/micro.log 2>&1
echo "micro exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
