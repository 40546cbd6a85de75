#!/bin/bash
# Final pass of round 2 on one B200 (the tree as it ships), part A: GPU suite (default and PDL build), smoke, bench at every BASELINE.json
# config (with the CPU baseline in the headline line), reference arm, microbench.  Part B (ncu evidence): tools/gpu_final_r02_ncu.sh.
mkdir -p gpurun_out
timeout 500 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/final_pytest_gpu.log 2>&1
echo "pytest gpu exit $?  $(tail -1 gpurun_out/final_pytest_gpu.log)"; grep -E "^FAILED|^ERROR" gpurun_out/final_pytest_gpu.log | head
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?"
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/final_bench_mxfp8.json 2> gpurun_out/final_bench_mxfp8.err; echo "bench mxfp8 exit $?"
timeout 300 python bench.py --steps 20 --warmup 3 --scenario default --no-cpu-baseline > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err; echo "bench default exit $?"
timeout 300 python bench.py --steps 20 --warmup 3 --scenario hybrid --no-cpu-baseline > gpurun_out/final_bench_hybrid.json 2> gpurun_out/final_bench_hybrid.err; echo "bench hybrid exit $?"
LLMFP8_PDL=1 timeout 500 python -m pytest tests/ -q -m gpu -p no:cacheprovider > gpurun_out/final_pytest_gpu_pdl.log 2>&1
echo "pytest gpu (PDL build) exit $?  $(tail -1 gpurun_out/final_pytest_gpu_pdl.log)"
LLMFP8_PDL=1 timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_mxfp8_pdl.json 2> gpurun_out/final_bench_mxfp8_pdl.err; echo "bench mxfp8 PDL exit $?"
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_mxfp8_b.json 2> gpurun_out/final_bench_mxfp8_b.err; echo "bench mxfp8 (again) exit $?"
LLMFP8_PDL=1 timeout 400 python bench.py --steps 20 --warmup 3 --scenario default --no-cpu-baseline > gpurun_out/final_bench_default_pdl.json 2> gpurun_out/final_bench_default_pdl.err; echo "bench default PDL exit $?"
timeout 300 python bench.py --model llama-3.2-1b --scenario default --batch 16 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_cfg1b_default.json 2> gpurun_out/final_bench_cfg1b.err; echo "bench cfg 1B exit $?"
timeout 300 python bench.py --model llama-3.2-3b --scenario hybrid --batch 16 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_cfg3b_hybrid.json 2> gpurun_out/final_bench_cfg3b.err; echo "bench cfg 3B exit $?"
timeout 300 python bench.py --model qwen2.5-14b --batch 16 --layers 24 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_qwen14b_mxfp8.json 2> gpurun_out/final_bench_qwen14b.err; echo "bench qwen14b exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "bench reference exit $?"
timeout 400 python tools/microbench.py --model llama8b --out gpurun_out/final_micro_llama8b.jsonl > gpurun_out/final_micro.log 2>&1; echo "microbench exit $?"
timeout 200 python tools/microbench.py --model llama8b --only lmhead --out gpurun_out/final_micro_llama8b.jsonl >> gpurun_out/final_micro.log 2>&1; echo "microbench lm_head exit $?"
timeout 400 python tools/microbench.py --model qwen14b --only gemm --no-bwd --ms 512,1024,2048,4096,8192,16384 --out gpurun_out/final_micro_qwen14b_msweep.jsonl > gpurun_out/final_micro_qwen.log 2>&1; echo "microbench qwen sweep exit $?"
for f in gpurun_out/final_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1].split('final_bench_')[1]:28s} value {d['value']:8.1f} {d['unit']}  {d.get('ms_per_step', 0):7.2f} ms  e2e {d['e2e']['value']:8.1f}  sm {(d.get('clocks') or {}).get('sm_mhz')}")
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
done
du -sh gpurun_out
