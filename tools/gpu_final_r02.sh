#!/bin/bash
# Final pass of round 2 on one B200 (the tree as it ships): GPU suite, headline bench lines (with the CPU baselines), reference arm, then the
# ncu evidence for profiles/: launch list of a 2-layer step and --set full captures of EVERY GEMM launch of one step (fprop, dgrad, wgrad of
# qkv / proj / fc1 / fc2) and of every quantiser, for the MXFP8 and the per-tensor scenario.  Summarise here with: python tools/ncu_summary.py r02
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
timeout 300 python bench.py --model llama-3.2-1b --scenario default --batch 16 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_cfg1b_default.json 2> gpurun_out/final_bench_cfg1b.err; echo "bench cfg 1B exit $?"
timeout 300 python bench.py --model llama-3.2-3b --scenario hybrid --batch 16 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_cfg3b_hybrid.json 2> gpurun_out/final_bench_cfg3b.err; echo "bench cfg 3B exit $?"
timeout 300 python bench.py --model qwen2.5-14b --batch 16 --layers 24 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/final_bench_qwen14b_mxfp8.json 2> gpurun_out/final_bench_qwen14b.err; echo "bench qwen14b exit $?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "bench reference exit $?"
timeout 400 python tools/microbench.py --model llama8b --out gpurun_out/final_micro_llama8b.jsonl > gpurun_out/final_micro.log 2>&1; echo "microbench exit $?"
timeout 200 python tools/microbench.py --model llama8b --only lmhead --out gpurun_out/final_micro_llama8b.jsonl >> gpurun_out/final_micro.log 2>&1; echo "microbench lm_head exit $?"
timeout 400 python tools/microbench.py --model qwen14b --only gemm --no-bwd --ms 512,1024,2048,4096,8192,16384 --out gpurun_out/final_micro_qwen14b_msweep.jsonl > gpurun_out/final_micro_qwen.log 2>&1; echo "microbench qwen sweep exit $?"
for sc in mxfp8 default; do
  sfx=""; [ $sc = default ] && sfx="_pt"
  CMD="python bench.py --layers 2 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --scenario $sc"
  if [ $sc = mxfp8 ]; then
    timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/final_ncu_list.log 2>&1; echo "ncu list exit $?"
  fi
  # 3 warm-up steps x 24 GEMM launches are skipped; the next 24 are one whole timed-region step of the 2-layer stack
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_kernel -s 72 -c 24 -f -o gpurun_out/prof_gemm$sfx $CMD > gpurun_out/final_ncu_gemm$sfx.log 2>&1; echo "ncu gemm $sc exit $?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"quantize_kernel|swiglu|rmsnorm|cast_amax" -s 90 -c 40 -f -o gpurun_out/prof_quant$sfx $CMD > gpurun_out/final_ncu_quant$sfx.log 2>&1; echo "ncu quant $sc exit $?"
done
ls -la gpurun_out | head -40
