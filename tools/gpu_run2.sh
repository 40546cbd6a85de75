#!/bin/bash
# Second pass: the driver's own test command, the residual-epilogue microbench, bench lines at every BASELINE.json config, Qwen2.5-14B M sweep.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider --timeout 240 > gpurun_out/pytest_gpu.log 2>&1
echo "pytest -m gpu exit $?" >> gpurun_out/summary.txt; tail -3 gpurun_out/pytest_gpu.log
run_bench() {  # name, args...
  local name=$1; shift
  timeout 600 python bench.py "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  echo "bench $name exit $?" >> gpurun_out/summary.txt
  tail -c 300 gpurun_out/bench_$name.err
}
run_bench mxfp8 --steps 5 --warmup 3
run_bench cfg2_llama1b_default --model llama-3.2-1b --scenario default --batch 16 --steps 5 --warmup 3 --no-cpu-baseline
run_bench cfg3_llama3b_hybrid --model llama-3.2-3b --scenario hybrid --batch 16 --steps 5 --warmup 3 --no-cpu-baseline
run_bench qwen14b_mxfp8 --model qwen2.5-14b --scenario mxfp8 --batch 16 --layers 24 --steps 3 --warmup 3 --no-cpu-baseline
timeout 300 python tools/microbench.py --model llama8b --only gemm --out gpurun_out/micro_gemm.jsonl > gpurun_out/micro_gemm.log 2>&1
echo "micro gemm exit $?" >> gpurun_out/summary.txt
timeout 600 python tools/microbench.py --model qwen14b --ms 512,1024,2048,4096,8192,16384 --no-bwd --out gpurun_out/micro_qwen14b_sweep.jsonl > gpurun_out/micro_qwen.log 2>&1
echo "micro qwen sweep exit $?" >> gpurun_out/summary.txt
timeout 300 python tools/loss_band.py > gpurun_out/loss_band.txt 2>&1
echo "loss band exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
