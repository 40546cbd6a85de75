#!/bin/bash
# GEMM iteration: parity tests (short timeouts: a scheduler bug hangs), then dynamic (CLC) vs static tile scheduling
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -q -m gpu -p no:cacheprovider --timeout 60 -x > gpurun_out/test_gpu_gemm.log 2>&1
rc=$?; echo "gemm tests exit $rc"; tail -12 gpurun_out/test_gpu_gemm.log
[ $rc -ne 0 ] && exit 1
for dbg in 0 8; do
  timeout 300 python tools/microbench.py --model llama8b --only gemm --debug $dbg --tag dbg$dbg --out gpurun_out/micro_gemm.jsonl > gpurun_out/micro_gemm_$dbg.log 2>&1
  echo "micro debug=$dbg exit $?"
done
python - <<'PY'
import json, collections
rows = collections.OrderedDict()
for l in open("gpurun_out/micro_gemm.jsonl"):
    d = json.loads(l)
    if "TFLOPs" in d: rows.setdefault(d["kernel"], {})[d.get("tag", "")] = d["TFLOPs"]
for k, v in rows.items():
    print(f"{k:64s}", "  ".join(f"{t}={x:7.0f}" for t, x in v.items()))
PY
