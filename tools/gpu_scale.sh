#!/bin/bash
# data-parallel bench at N GPUs (weak scaling), all three scenarios for N given as $1 (default 2)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
for sc in ${2:-mxfp8 default}; do
  LLMFP8_DIST_TIMEOUT_S=90 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --scenario $sc > gpurun_out/scale_${sc}_n$N.json 2> gpurun_out/scale_${sc}_n$N.err
  echo "scale $sc n=$N exit $?"
  tail -c 400 gpurun_out/scale_${sc}_n$N.err
  python tools/show_bench.py gpurun_out/scale_${sc}_n$N.json 2>&1 | head -3
done
