#!/bin/bash
# SASS evidence for profiles/: per object file, how many tcgen05 / TMA / multimem instructions the sm_100a code carries
# (mnemonics as listed in /opt/skills/guides/B200_PROFILING.md).  Runs without a GPU:  bash tools/sass_summary.sh > profiles/r02/sass_counts.txt
cd "$(dirname "$0")/../llm_fp8_b200/csrc" || exit 1
echo "# cuobjdump -sass <object> | grep -c <mnemonic>   ($(/usr/local/cuda/bin/nvcc --version | tail -1), -gencode arch=compute_100a,code=sm_100a)"
printf "%-22s %8s %8s %8s %8s %8s %8s %8s %8s %8s %8s\n" object UTCQMMA UTCHMMA UTCCP LDTM UTMALDG UTMASTG UBLKCP UCGABAR LDGMC SYNCS
for o in gemm.o mx_quantize.o mx_act_quantize.o adamw_mx.o quantize.o rmsnorm.o swiglu.o nvls_reduce.o lm_head.o; do
  s=$(/usr/local/cuda/bin/cuobjdump -sass $o 2>/dev/null)
  c() { echo "$s" | grep -c "$1"; }
  printf "%-22s %8s %8s %8s %8s %8s %8s %8s %8s %8s %8s\n" $o $(c UTCQMMA) $(c UTCHMMA) $(c UTCCP) $(c LDTM) $(c UTMALDG) $(c UTMASTG) $(c UBLKCP) $(c UCGABAR) $(c LDGMC) $(c "SYNCS")
done
echo
echo "# gemm.o: distinct tensor-core / TMA instruction forms"
/usr/local/cuda/bin/cuobjdump -sass gemm.o | grep -oE "UTC[A-Z]*MMA[.A-Z0-9_]*|UTCCP[.A-Z0-9_]*|LDTM[.A-Z0-9_]*|UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|UTCBAR[.A-Z0-9_]*" | sort | uniq -c | sort -rn
