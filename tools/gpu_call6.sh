#!/bin/bash
mkdir -p gpurun_out
python tools/gemm_trace.py > gpurun_out/r02f_gemm_trace.txt 2>&1; echo "trace exit $?"
timeout 300 python -m pytest tests/test_gpu_gemm.py -q -m gpu -p no:cacheprovider > gpurun_out/r02f_pytest.log 2>&1; echo "pytest exit $? $(tail -1 gpurun_out/r02f_pytest.log)"
timeout 200 python tools/microbench.py --model llama8b --only lmhead --out gpurun_out/r02f_micro_lmhead.jsonl > gpurun_out/r02f_micro_lmhead.log 2>&1; echo "micro lmhead exit $?"; tail -8 gpurun_out/r02f_micro_lmhead.log | cut -c1-300
head -60 gpurun_out/r02f_gemm_trace.txt
