#!/bin/bash
# ncu pass: launch list of a short bench run, then --set full captures of the GEMM and the MX quantiser
mkdir -p gpurun_out
CMD="python bench.py --layers 2 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline --scenario ${1:-mxfp8}"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "list exit $?"
$CMD > gpurun_out/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_kernel -s 48 -c 6 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm exit $?"
$CMD > gpurun_out/ncu_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"mx_quantize_kernel|swiglu|rmsnorm|cast_amax" -s 60 -c 12 -o gpurun_out/prof_quant $CMD > gpurun_out/ncu_quant.log 2>&1
echo "quant exit $?"
ls -la gpurun_out
