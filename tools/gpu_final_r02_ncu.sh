#!/bin/bash
# Final pass, part B: ncu evidence for profiles/.  Launch list of a 2-layer step; --set full captures of EVERY GEMM launch (fprop, dgrad, wgrad of
# qkv / proj / fc1 / fc2) and every quantiser launch of one step of a 1-layer stack, MXFP8 and per-tensor.  The reports are exported to CSV
# on the box and deleted (they exceed gpurun's 64 MiB return limit).  Summarise here with: python tools/ncu_summary.py r02
mkdir -p gpurun_out
CMD2="python bench.py --layers 2 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD2 > gpurun_out/final_ncu_list.log 2>&1; echo "ncu list exit $?"
for sc in mxfp8 default; do
  sfx=""; [ $sc = default ] && sfx="_pt"
  CMD="python bench.py --layers 1 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --scenario $sc"
  # 3 warm-up steps x 12 GEMM launches are skipped; the next 12 are the timed step of the 1-layer stack
  timeout 500 ncu --set full --clock-control none -k regex:fp8_gemm_kernel -s 36 -c 12 -f -o gpurun_out/prof_gemm$sfx $CMD > gpurun_out/final_ncu_gemm$sfx.log 2>&1; echo "ncu gemm $sc exit $?"
  ncu -i gpurun_out/prof_gemm$sfx.ncu-rep --page raw --csv > gpurun_out/prof_gemm${sfx}_raw.csv 2>/dev/null; rm -f gpurun_out/prof_gemm$sfx.ncu-rep
  timeout 500 ncu --set full --clock-control none -k regex:"quantize_kernel|swiglu|rmsnorm|cast_amax" -s 66 -c 26 -f -o gpurun_out/prof_quant$sfx $CMD > gpurun_out/final_ncu_quant$sfx.log 2>&1; echo "ncu quant $sc exit $?"
  ncu -i gpurun_out/prof_quant$sfx.ncu-rep --page raw --csv > gpurun_out/prof_quant${sfx}_raw.csv 2>/dev/null; rm -f gpurun_out/prof_quant$sfx.ncu-rep
done
du -sh gpurun_out; ls -la gpurun_out | grep -E "prof_|launches"
