#!/bin/bash
# BASELINE.json configs[2] (Llama-3.2-3B, te_llama_hybrid recipe: delayed scaling, the amax MAX exchange on the path; 16 x 512 tokens per GPU) at 1 and 8 GPUs
# on one box with the defaults (coalesced gradient groups) — the earlier line (6.51x) predates the groups.
mkdir -p gpurun_out
A="--model llama-3.2-3b --scenario hybrid --batch 16 --steps 5 --warmup 3 --no-roofline --no-cpu-baseline"
timeout 150 python bench.py --gpus 1 $A > gpurun_out/r02w_cfg3b_n1.json 2> gpurun_out/r02w_cfg3b_n1.err; echo "n=1 exit $?"
LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 \
    bench.py --gpus 8 $A > gpurun_out/r02w_cfg3b_n8.json 2> gpurun_out/r02w_cfg3b_n8.err; echo "n=8 exit $?"
python - <<'PY'
import json
base = None
for n in (1, 8):
    try:
        d = json.loads(open(f"gpurun_out/r02w_cfg3b_n{n}.json").read().strip().splitlines()[-1])
        base = base or d["value"]
        tl = d.get("timeline") or {}
        print(f"n={n}: {d['value']:8.1f} TF/s total  {d['ms_per_step']:7.2f} ms/step  x{d['value'] / base:5.2f}  "
              f"fwd/bwd/tail {[round(tl.get(k, 0), 1) for k in ('forward_ms', 'backward_ms', 'reduce_tail_ms')]}  sm {d['clocks']['sm_mhz']}")
    except Exception as e:
        print(f"n={n}: unreadable", e)
PY
