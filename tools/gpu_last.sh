#!/bin/bash
mkdir -p gpurun_out
(timeout 55 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err; echo "bench exit $?" > gpurun_out/last_bench.txt) &
timeout 55 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/pytest_last.log 2>&1
echo "pytest exit $?"; tail -3 gpurun_out/pytest_last.log
wait
cat gpurun_out/last_bench.txt; tail -c 300 gpurun_out/bench_last.err; head -c 700 gpurun_out/bench_last.json
