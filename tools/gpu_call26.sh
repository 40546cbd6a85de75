#!/bin/bash
# llmfp8_rmsnorm_bwd_noalloc: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r02s_pytest_noalloc_alone.log 2>&1; echo "alone exit $?"; tail -5 gpurun_out/r02s_pytest_noalloc_alone.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02s_pytest_gpu.log 2>&1; echo "suite exit $?"; tail -5 gpurun_out/r02s_pytest_gpu.log
