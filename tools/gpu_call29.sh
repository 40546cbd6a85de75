#!/bin/bash
# final tree (32-column accumulator overlap as the default, no-alloc K5 entry): whole GPU suite, smoke, default bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r02v_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/r02v_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02v_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r02v_smoke.log
timeout 400 python bench.py > gpurun_out/r02v_bench_mxfp8.json 2> gpurun_out/r02v_bench_mxfp8.err; echo "bench exit $?"; cat gpurun_out/r02v_bench_mxfp8.json | cut -c1-400
timeout 300 python bench.py --impl reference > gpurun_out/r02v_bench_reference.json 2> gpurun_out/r02v_bench_reference.err; echo "reference exit $?"; cut -c1-300 gpurun_out/r02v_bench_reference.json
