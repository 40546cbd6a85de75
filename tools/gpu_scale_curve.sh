#!/bin/bash
# the shipped tree at N = 1, 2, 4, 8 on ONE 8-GPU box, launched the way the driver launches it (defaults only)
mkdir -p gpurun_out
S=gpurun_out/r02n_scale_curve.txt; : > $S
timeout 200 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/r02n_scale_n1.json 2> gpurun_out/r02n_scale_n1.err; echo "n=1 exit $?" >> $S
for n in 2 4 8; do
  LLMFP8_DIST_TIMEOUT_S=90 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n \
      bench.py --gpus $n --steps 5 --warmup 3 --no-roofline > gpurun_out/r02n_scale_n$n.json 2> gpurun_out/r02n_scale_n$n.err; echo "n=$n exit $?" >> $S
done
python - >> $S <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/r02n_scale_n{n}.json").read().strip().splitlines()[-1])
        base = base or d["value"]
        tl = d.get("timeline") or {}
        print(f"n={n}: {d['value']:8.1f} TF/s total  {d['ms_per_step']:7.2f} ms/step  x{d['value'] / base:5.2f}  efficiency {d['value'] / base / n:5.3f}  "
              f"fwd/bwd/tail {[round(tl.get(k, 0), 1) for k in ('forward_ms', 'backward_ms', 'reduce_tail_ms')]}  sm {d['clocks']['sm_mhz']}")
    except Exception as e:
        print(f"n={n}: unreadable", e)
PY
cat $S
