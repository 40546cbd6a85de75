#!/bin/bash
# bench pass on one GPU: all three scenarios, then the module tests
mkdir -p gpurun_out
for sc in mxfp8 default hybrid; do
  timeout 900 python bench.py --steps 3 --warmup 3 --scenario $sc > gpurun_out/bench_$sc.json 2> gpurun_out/bench_$sc.err
  echo "bench $sc exit $?" >> gpurun_out/summary2.txt
  tail -c 600 gpurun_out/bench_$sc.err
done
timeout 600 python -m pytest tests/test_gpu_modules.py -q -m gpu -p no:cacheprovider --timeout 120 > gpurun_out/test_gpu_modules.log 2>&1
echo "modules exit $?" >> gpurun_out/summary2.txt
cat gpurun_out/summary2.txt
