#!/bin/bash
# One GPU-box pass after a kernel change: GPU tests (one process per file), bench lines (fused / graph / plain variants, three scenarios),
# ncu launch list + full captures, microbench.  Everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in test_gpu_quantize test_gpu_gemm test_gpu_modules; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu -p no:cacheprovider --timeout 180 > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$f.log
done
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
run_bench() {  # name, args...
  local name=$1; shift
  timeout 600 python bench.py "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  echo "bench $name exit $?" >> gpurun_out/summary.txt
  tail -c 400 gpurun_out/bench_$name.err
}
run_bench mxfp8 --steps 5 --warmup 3
run_bench mxfp8_graph --steps 5 --warmup 3 --graph 1 --no-cpu-baseline
run_bench mxfp8_plain --steps 5 --warmup 3 --no-fuse-residual --no-cpu-baseline
run_bench default --steps 5 --warmup 3 --scenario default
run_bench hybrid --steps 5 --warmup 3 --scenario hybrid --no-cpu-baseline
run_bench reference --impl reference --steps 2 --warmup 1
bash tools/gpu_ncu.sh mxfp8 > gpurun_out/ncu_driver.log 2>&1
timeout 600 python tools/microbench.py --model llama8b --out gpurun_out/micro.jsonl > gpurun_out/micro.log 2>&1
echo "micro exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
