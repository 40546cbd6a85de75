#!/bin/bash
# block-scaled GEMM: overlapped accumulators 224 apart (32 shared TMEM columns, debug bit 8) against 192 apart (64 shared, the default)
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu -k "overlapped" > gpurun_out/r02u_pytest_ovl.log 2>&1; rc=$?; echo "pytest ovl exit $rc"; tail -5 gpurun_out/r02u_pytest_ovl.log
[ $rc -ne 0 ] && exit 0
export LLMFP8_ALLOW_TUNING=1
for dbg in 0 256 0 256; do
  timeout 300 python tools/microbench.py --only gemm --debug $dbg --tag dbg$dbg --no-bwd --out gpurun_out/r02u_micro_gemm_dbg${dbg}.jsonl >> gpurun_out/r02u_micro_gemm.log 2>&1; echo "micro debug $dbg exit $?"
done
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02u_bench_192.json 2> gpurun_out/r02u_bench_192.err; echo "bench 192 exit $?"
LLMFP8_GEMM_DEBUG=256 $B > gpurun_out/r02u_bench_224.json 2> gpurun_out/r02u_bench_224.err; echo "bench 224 exit $?"
$B > gpurun_out/r02u_bench_192_b.json 2> gpurun_out/r02u_bench_192_b.err; echo "bench 192 (again) exit $?"
LLMFP8_GEMM_DEBUG=256 $B > gpurun_out/r02u_bench_224_b.json 2> gpurun_out/r02u_bench_224_b.err; echo "bench 224 (again) exit $?"
python - <<'PY'
import json
for dbg in (0, 256):
    try:
        for l in open(f"gpurun_out/r02u_micro_gemm_dbg{dbg}.jsonl"):
            d = json.loads(l)
            if "mxfp8" in d["kernel"] and d["kernel"].startswith("G"):
                print(f"debug {dbg:3d} {d['kernel'][:60]:60s} {d.get('TFLOPs', 0):8.1f} TF  sm {d.get('clocks', {}).get('sm_mhz')}")
    except Exception as e:
        print(dbg, "unreadable", e)
for t in ("192", "224", "192_b", "224_b"):
    try:
        d = json.loads(open(f"gpurun_out/r02u_bench_{t}.json").read().strip().splitlines()[-1])
        print(f"{t:6s} {d['ms_per_step']:7.2f} ms/step  value {d['value']:7.1f}  gemm avg {d['roofline']['avg_launch_ms']*1000:6.1f} us  {d['roofline']['achieved']:6.0f} TF")
    except Exception as e:
        print(t, "unreadable", e)
PY
