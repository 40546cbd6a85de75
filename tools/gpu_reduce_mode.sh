#!/bin/bash
# gradient all-reduce as SUM of prescaled gradients (lets NCCL pick its NVLS algorithms) vs AVG, N GPUs, mxfp8; NCCL's algorithm choice logged
N=${1:-8}
mkdir -p gpurun_out
run() {  # tag, env...
  tag=$1; shift
  env "$@" LLMFP8_DIST_TIMEOUT_S=90 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 3 --warmup 3 --scenario mxfp8 --no-roofline > gpurun_out/reduce_${tag}_n$N.json 2> gpurun_out/reduce_${tag}_n$N.err
  echo "$tag n=$N exit $?"; tail -c 400 gpurun_out/reduce_${tag}_n$N.err | grep -v "OMP\|\*\*\*"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/reduce_${tag}_n$N.json").read().strip().splitlines()[-1])
    print("   ${tag}", round(d["value"]), "TF total", round(d["ms_per_step"],1), "ms/step  e2e", round(d["e2e"]["value"]))
except Exception as e: print("   parse failed", e)
PY
}
run sum LLMFP8_GRAD_REDUCE=sum NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,TUNING NCCL_DEBUG_FILE=gpurun_out/nccl_sum_%p.log
grep -h "AllReduce:" gpurun_out/nccl_sum_*.log | sed 's/.*AllReduce:/AllReduce:/' | sort | uniq -c | sort -rn | head -12 > gpurun_out/nccl_sum_summary.txt
grep -h -A3 "Algorithm   |" $(ls gpurun_out/nccl_sum_*.log | head -1) | grep "AllReduce\|Algorithm" | head -8 >> gpurun_out/nccl_sum_summary.txt
rm -f gpurun_out/nccl_sum_*.log
cat gpurun_out/nccl_sum_summary.txt
run avg LLMFP8_GRAD_REDUCE=avg
