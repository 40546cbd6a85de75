#!/bin/bash
# ncu --set full of the quantise kernels inside the microbench (one launch of each, after warm-up), stall reasons included
mkdir -p gpurun_out
CMD="python tools/microbench.py --model llama8b --only quant --out gpurun_out/micro_ncu_scratch.jsonl"
$CMD > gpurun_out/ncu_q_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"dswiglu_quant_kernel|swiglu_quant_kernel|mx_quantize_kernel|rmsnorm_bwd_kernel|rmsnorm_fwd_kernel|cast_amax" \
    --launch-skip 30 --launch-count 60 -o gpurun_out/prof_quant2 -f $CMD > gpurun_out/ncu_q.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_quant2.ncu-rep --page raw --csv > gpurun_out/prof_quant2_raw.csv 2> gpurun_out/ncu_q_export.log
rm -f gpurun_out/prof_quant2.ncu-rep
ls -la gpurun_out
