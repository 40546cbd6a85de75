#!/bin/bash
# First hardware run of the in-switch gradient reduce (csrc/nvls_reduce.cu, dp.NvlsGradArena) on TWO GPUs: the gated test, then an
# A/B of the data-parallel bench line (NCCL vs own kernel) on the same box.  Every step under its own timeout (the kernel gives up
# after 20 s / 10 s by itself and the process group after 60 / 90 s), so a protocol bug cannot hang the box:
#   /usr/local/graft/bin/gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_nvls2.sh'
mkdir -p gpurun_out
# data-parallel correctness of the FP8 path (tests/test_gpu_dp2.py): scales identical across ranks, gradients == single GPU on the concatenated batch
timeout 500 python -m pytest tests/test_gpu_dp2.py tests/test_gpu_comm.py -q -s -m gpu -p no:cacheprovider > gpurun_out/r02_test_dp2.log 2>&1
echo "test_dp2 exit $?"; grep -E "^\[dp2\]|passed|failed" gpurun_out/r02_test_dp2.log
timeout 280 python -m pytest tests/test_gpu_comm.py -x -q -m gpu -p no:cacheprovider -k nvls > gpurun_out/r02_test_nvls2.log 2>&1
rc=$?; echo "test_nvls exit $rc"; tail -25 gpurun_out/r02_test_nvls2.log
[ $rc -ne 0 ] && exit $rc
LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 \
    tools/nvls_bench.py --out gpurun_out/r02_nvls_bench_n2.jsonl > gpurun_out/r02_nvls_bench_n2.log 2>&1
echo "nvls_bench exit $?"; tail -20 gpurun_out/r02_nvls_bench_n2.log
# per-tensor scenario with the fused in-switch amax + scale update (and the in-switch gradient reduce)
LLMFP8_NVLS_AMAX=1 LLMFP8_NVLS_REDUCE=1 LLMFP8_DIST_TIMEOUT_S=90 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
    --master-port 29522 bench.py --gpus 2 --steps 4 --warmup 3 --scenario default --no-roofline > gpurun_out/r02_nvls_amax_default_n2.json 2> gpurun_out/r02_nvls_amax_default_n2.err
echo "bench default nvls amax+reduce exit $?"; tail -c 400 gpurun_out/r02_nvls_amax_default_n2.json; tail -3 gpurun_out/r02_nvls_amax_default_n2.err
for mode in 0 1; do
  LLMFP8_NVLS_REDUCE=$mode LLMFP8_DIST_TIMEOUT_S=90 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29521 bench.py --gpus 2 --steps 4 --warmup 3 --scenario mxfp8 --no-roofline > gpurun_out/r02_nvls${mode}_n2.json 2> gpurun_out/r02_nvls${mode}_n2.err
  echo "bench nvls=$mode exit $?"; tail -c 600 gpurun_out/r02_nvls${mode}_n2.json; tail -5 gpurun_out/r02_nvls${mode}_n2.err
done
