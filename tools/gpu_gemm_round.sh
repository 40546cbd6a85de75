#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py -q -m gpu -p no:cacheprovider --timeout 120 > gpurun_out/test_gpu_gemm.log 2>&1
echo "gemm tests exit $?"; tail -15 gpurun_out/test_gpu_gemm.log
for cfg in "2 256 0" "2 192 0" "1 256 0"; do
  set -- $cfg
  timeout 600 python tools/microbench.py --model llama8b --only gemm --cg $1 --bn $2 --debug $3 --out gpurun_out/micro_gemm.jsonl > gpurun_out/micro_gemm_$1_$2_$3.log 2>&1
  echo "micro cg=$1 bn=$2 debug=$3 exit $?"
done
