#!/bin/bash
# Round-end rehearsal on one GPU: exactly what the driver runs (pytest -m gpu, smoke, bench default flags, bench --impl reference)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest -m gpu exit $?" > gpurun_out/summary.txt; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "bench reference exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py > gpurun_out/bench_mxfp8.json 2> gpurun_out/bench_mxfp8.err; echo "bench exit $?" >> gpurun_out/summary.txt
timeout 600 python bench.py --scenario default --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench default exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
