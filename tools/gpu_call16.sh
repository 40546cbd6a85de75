#!/bin/bash
# same box, same GEMM code: is the in-step GEMM rate a property of the kernel or of what runs between the GEMMs (power budget)?
mkdir -p gpurun_out
B="timeout 170 python bench.py --steps 10 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02j_bench_new.json 2> gpurun_out/r02j_bench_new.err; echo "new defaults exit $?"
LLMFP8_MX_FUSED_ACT=0 LLMFP8_WGRAD_STREAM=0 LLMFP8_MX_EPILOGUE_RESIDUAL=0 $B > gpurun_out/r02j_bench_old.json 2> gpurun_out/r02j_bench_old.err; echo "round-1 defaults exit $?"
LLMFP8_WGRAD_STREAM=0 $B > gpurun_out/r02j_bench_nowgradstream.json 2> gpurun_out/r02j_bench_nowgradstream.err; echo "no wgrad stream exit $?"
$B > gpurun_out/r02j_bench_new2.json 2> gpurun_out/r02j_bench_new2.err; echo "new defaults again exit $?"
python - <<'PY'
import json
for t in ("new", "old", "nowgradstream", "new2"):
    d = json.loads(open(f"gpurun_out/r02j_bench_{t}.json").read().strip().splitlines()[-1])
    r = d["roofline"]; sh = d["kernel_time_shares"]
    print(f"{t:14s} {d['value']:7.1f} TF/s {d['ms_per_step']:7.2f} ms | GEMM launches avg {r['avg_launch_ms']*1e3:6.1f} us = {r['achieved']:5.0f} TF, sum {r['avg_launch_ms']*r['launches']:6.1f} ms | library kernels {sh['_library_kernels_ms']:6.1f} ms | sm {d['clocks']['sm_mhz']} MHz {d['clocks']['power_w_max']} W")
PY
