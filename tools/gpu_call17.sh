#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/microbench.py --model llama8b --only sustained --out gpurun_out/r02k_micro_sustained.jsonl > gpurun_out/r02k_micro_sustained.log 2>&1; echo "exit $?"
python - <<'PY'
import json
for l in open("gpurun_out/r02k_micro_sustained.jsonl"):
    d = json.loads(l); print(d["kernel"][:80], round(d["ms"], 3), "ms", round(d["TFLOPs"]), "TF", d["clocks"])
PY
