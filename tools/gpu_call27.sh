#!/bin/bash
# ncu evidence of the shipped tree after the ring / two-CTA quantiser kernels: launch list of a 2-layer step, --set full of every quantiser launch
# of one step (MXFP8 and per-tensor).  GEMM captures of call 15 stay valid (gemm.cu unchanged).  Summarise: python tools/ncu_summary.py r02b
mkdir -p gpurun_out
CMD2="python bench.py --layers 2 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 $CMD2 > gpurun_out/r02t_plain.json 2> gpurun_out/r02t_plain.err; echo "plain run exit $?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD2 > gpurun_out/r02t_ncu_list.log 2>&1; echo "ncu list exit $?"
for sc in mxfp8 default; do
  sfx=""; [ $sc = default ] && sfx="_pt"
  CMD="python bench.py --layers 1 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --scenario $sc"
  timeout 500 ncu --set full --clock-control none -k regex:"quantize_kernel|swiglu|rmsnorm|cast_amax" -s 66 -c 26 -f -o gpurun_out/prof_quant$sfx $CMD > gpurun_out/r02t_ncu_quant$sfx.log 2>&1; echo "ncu quant $sc exit $?"
  ncu -i gpurun_out/prof_quant$sfx.ncu-rep --page raw --csv > gpurun_out/prof_quant${sfx}_raw.csv 2>/dev/null; rm -f gpurun_out/prof_quant$sfx.ncu-rep
done
du -sh gpurun_out
