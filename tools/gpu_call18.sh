#!/bin/bash
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_optim.py -q -m gpu -p no:cacheprovider > gpurun_out/r02l_pytest_optim.log 2>&1
echo "pytest optim exit $?  $(tail -1 gpurun_out/r02l_pytest_optim.log)"; grep -E "^FAILED|^ERROR|^E  " gpurun_out/r02l_pytest_optim.log | head -12 | cut -c1-250
timeout 200 python tools/microbench.py --model llama8b --only optim --out gpurun_out/r02l_micro_optim.jsonl > gpurun_out/r02l_micro_optim.log 2>&1; echo "micro optim exit $?"; tail -5 gpurun_out/r02l_micro_optim.log | cut -c1-330
