#!/bin/bash
# round 2, second call (2 GPUs): the whole GPU suite (incl. the new MXFP8 module-level oracle tests), then tools/gpu_nvls2.sh
mkdir -p gpurun_out
timeout 500 python -m pytest tests/ -q -m gpu -p no:cacheprovider --deselect tests/test_gpu_dp2.py > gpurun_out/r02b_pytest_gpu.log 2>&1
echo "pytest gpu exit $?  $(tail -1 gpurun_out/r02b_pytest_gpu.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r02b_pytest_gpu.log | head -20
bash tools/gpu_nvls2.sh
