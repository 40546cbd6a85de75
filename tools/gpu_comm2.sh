#!/bin/bash
# the C ABI's NCCL collectives on two GPUs (tests/test_gpu_comm.py)
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_comm.py -x -q -m gpu -p no:cacheprovider > gpurun_out/test_gpu_comm.log 2>&1
echo "test_gpu_comm exit $?"; tail -15 gpurun_out/test_gpu_comm.log
