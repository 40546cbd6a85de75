#!/bin/bash
# data-parallel bench at N GPUs under different NCCL CTA budgets (SM contention with the persistent GEMMs)
N=${1:-8}
mkdir -p gpurun_out
run() {  # tag, env...
  tag=$1; shift
  env "$@" LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 3 --warmup 3 --scenario mxfp8 --no-roofline > gpurun_out/sweep_${tag}_n$N.json 2> gpurun_out/sweep_${tag}_n$N.err
  echo "$tag n=$N exit $?"; tail -c 300 gpurun_out/sweep_${tag}_n$N.err | grep -v OMP
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/sweep_${tag}_n$N.json").read().strip().splitlines()[-1])
    print("   ${tag}", round(d["value"]), "TF total", round(d["ms_per_step"],1), "ms/step  e2e", round(d["e2e"]["value"]), d["clocks"])
except Exception as e: print("   parse failed", e)
PY
}
for cfg in ${2:-16 8 4}; do
  run maxctas$cfg NCCL_MAX_CTAS=$cfg
done
