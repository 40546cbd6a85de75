#!/bin/bash
mkdir -p gpurun_out
timeout 80 python -m pytest tests/test_gpu_modules.py -x -q -m gpu -p no:cacheprovider > gpurun_out/test_gpu_modules.log 2>&1
echo "test_gpu_modules exit $?"; tail -12 gpurun_out/test_gpu_modules.log
