#!/bin/bash
mkdir -p gpurun_out
B="timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02m_bench_noopt.json 2> gpurun_out/r02m_bench_noopt.err; echo "no optimizer exit $?"
$B --optimizer fused > gpurun_out/r02m_bench_fused.json 2> gpurun_out/r02m_bench_fused.err; echo "fused exit $?"; tail -3 gpurun_out/r02m_bench_fused.err
$B --optimizer torch > gpurun_out/r02m_bench_torch.json 2> gpurun_out/r02m_bench_torch.err; echo "torch exit $?"; tail -3 gpurun_out/r02m_bench_torch.err
python - <<'PY'
import json
for t in ("noopt", "fused", "torch"):
    try:
        d = json.loads(open(f"gpurun_out/r02m_bench_{t}.json").read().strip().splitlines()[-1])
        print(f"{t:6s} {d['ms_per_step']:7.2f} ms/step  value {d['value']:7.1f}  launches {d['gpu_launches']}  shares {d['kernel_time_shares']}")
    except Exception as e:
        print(t, "unreadable", e)
PY
