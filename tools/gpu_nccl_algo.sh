#!/bin/bash
# which all-reduce algorithm carries the weight gradients: default (with NCCL's own INFO log), NVLS forced, Ring forced — N GPUs, mxfp8
N=${1:-4}
mkdir -p gpurun_out
run() {  # tag, env...
  tag=$1; shift
  env "$@" LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 3 --warmup 3 --scenario mxfp8 --no-roofline > gpurun_out/algo_${tag}_n$N.json 2> gpurun_out/algo_${tag}_n$N.err
  echo "$tag n=$N exit $?"; tail -c 400 gpurun_out/algo_${tag}_n$N.err | grep -v "OMP\|\*\*\*"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/algo_${tag}_n$N.json").read().strip().splitlines()[-1])
    print("   ${tag}", round(d["value"]), "TF total", round(d["ms_per_step"],1), "ms/step  e2e", round(d["e2e"]["value"]))
except Exception as e: print("   parse failed", e)
PY
}
run default NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,ENV,TUNING NCCL_DEBUG_FILE=gpurun_out/nccl_info_%p.log
grep -h -i "nvls\|algo\|multicast" gpurun_out/nccl_info_*.log | sort | uniq -c | sort -rn | head -30 > gpurun_out/nccl_info_summary.txt
ls gpurun_out/nccl_info_*.log | tail -n +2 | xargs rm -f
run nvls NCCL_ALGO=NVLS
run ring NCCL_ALGO=Ring
cat gpurun_out/nccl_info_summary.txt | head -20
