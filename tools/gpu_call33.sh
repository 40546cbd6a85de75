#!/bin/bash
# two GPUs: does a smaller NCCL CTA budget give the backward GEMMs their SMs back?  (at N = 2 an all-reduce moves 1 x the buffer per rank, against 1.75 x at N = 8)
mkdir -p gpurun_out
run() {  # tag, env...
  tag=$1; shift
  env "$@" LLMFP8_DIST_TIMEOUT_S=90 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
      bench.py --gpus 2 --steps 5 --warmup 3 --no-roofline --no-cpu-baseline > gpurun_out/r02z_n2_$tag.json 2> gpurun_out/r02z_n2_$tag.err; echo "$tag exit $?"
}
run default LLMFP8_NOP=1
run maxctas8 NCCL_MAX_CTAS=8
run maxctas4 NCCL_MAX_CTAS=4
run default_b LLMFP8_NOP=1
python - <<'PY'
import json
for t in ("default", "maxctas8", "maxctas4", "default_b"):
    try:
        d = json.loads(open(f"gpurun_out/r02z_n2_{t}.json").read().strip().splitlines()[-1])
        tl = d.get("timeline") or {}
        print(f"{t:10s} {d['value']:8.1f} TF/s total  {d['ms_per_step']:7.2f} ms/step  fwd/bwd/tail {[round(tl.get(k, 0), 1) for k in ('forward_ms', 'backward_ms', 'reduce_tail_ms')]}  sm {d['clocks']['sm_mhz']}")
    except Exception as e:
        print(t, "unreadable", e)
PY
