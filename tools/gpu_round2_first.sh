#!/bin/bash
# First GPU call of the next round (one B200, 25-30 min of box time: 16 bench lines of ~1 min, two GPU suites, ncu): everything that was
# written after the round-1 GPU budget ran out.  Every step has its own timeout; results are written as they come, so a cut-off call
# still leaves gpurun_out/r02_* behind.
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash tools/gpu_round2_first.sh'
# 1. full GPU suite on the default tree (overlapped-accumulator MX GEMM is the default since d2c7555);
# 2. the gated experiments' own tests (LLMFP8_EXPERIMENTAL=1): RES + OVL epilogue (debug bit 5), SwiGLU/dSwiGLU fused with the MX
#    quantiser (LLMFP8_MX_FUSED_ACT), wgrad side stream (LLMFP8_WGRAD_STREAM);
# 3. bench lines: default tree, then each experiment switched on alone, then together (A/B on the same box, same clocks);
# 4. GEMM A/B at the 8B shapes (debug 16 = single accumulator vs 0 = overlapped) and the ncu launch list + one full capture.
mkdir -p gpurun_out
S=gpurun_out/r02_summary.txt
: > $S
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,power.limit --format=csv > gpurun_out/r02_smi.txt 2>&1

timeout 400 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/r02_pytest_gpu.log 2>&1
echo "pytest gpu exit $?  $(tail -1 gpurun_out/r02_pytest_gpu.log)" >> $S

LLMFP8_EXPERIMENTAL=1 timeout 200 python -m pytest tests/test_gpu_gemm.py -q -m gpu -p no:cacheprovider -k "residual_overlapped" > gpurun_out/r02_exp_res_ovl.log 2>&1
echo "exp RES+OVL exit $?  $(tail -1 gpurun_out/r02_exp_res_ovl.log)" >> $S
LLMFP8_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_quantize.py tests/test_golden.py -q -m gpu -p no:cacheprovider -k "mx_swiglu or mx_dswiglu or mx_act or denormal_quotients" > gpurun_out/r02_exp_fused_act.log 2>&1
echo "exp fused act exit $?  $(tail -1 gpurun_out/r02_exp_fused_act.log)" >> $S
LLMFP8_EXPERIMENTAL=1 timeout 200 python -m pytest tests/test_gpu_modules.py -q -m gpu -p no:cacheprovider -k "wgrad_side_stream" > gpurun_out/r02_exp_wgrad_stream.log 2>&1
echo "exp wgrad stream exit $?  $(tail -1 gpurun_out/r02_exp_wgrad_stream.log)" >> $S
LLMFP8_EXPERIMENTAL=1 timeout 200 python -m pytest tests/test_gpu_modules.py -q -m gpu -p no:cacheprovider -k "weight_prefetch" > gpurun_out/r02_exp_wprefetch.log 2>&1
echo "exp weight prefetch exit $?  $(tail -1 gpurun_out/r02_exp_wprefetch.log)" >> $S

# the programmatic-dependent-launch build: the whole GPU suite on it, then its bench line
LLMFP8_PDL=1 timeout 400 python -m pytest tests/ -x -q -m gpu -p no:cacheprovider > gpurun_out/r02_pytest_gpu_pdl.log 2>&1
echo "pytest gpu (PDL build) exit $?  $(tail -1 gpurun_out/r02_pytest_gpu_pdl.log)" >> $S

B="timeout 170 python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r02_bench_mxfp8.json 2> gpurun_out/r02_bench_mxfp8.err; echo "bench mxfp8 exit $?" >> $S
LLMFP8_WGRAD_STREAM=1 $B > gpurun_out/r02_bench_mxfp8_wgradstream.json 2> gpurun_out/r02_bench_mxfp8_wgradstream.err; echo "bench mxfp8 +wgrad stream exit $?" >> $S
LLMFP8_MX_EPILOGUE_RESIDUAL=1 LLMFP8_GEMM_DEBUG=32 $B > gpurun_out/r02_bench_mxfp8_resovl.json 2> gpurun_out/r02_bench_mxfp8_resovl.err; echo "bench mxfp8 +RES/OVL exit $?" >> $S
LLMFP8_MX_FUSED_ACT=1 $B > gpurun_out/r02_bench_mxfp8_fusedact.json 2> gpurun_out/r02_bench_mxfp8_fusedact.err; echo "bench mxfp8 +fused act exit $?" >> $S
LLMFP8_MX_FUSED_ACT=1 LLMFP8_WGRAD_STREAM=1 LLMFP8_MX_EPILOGUE_RESIDUAL=1 LLMFP8_GEMM_DEBUG=32 $B > gpurun_out/r02_bench_mxfp8_all.json 2> gpurun_out/r02_bench_mxfp8_all.err; echo "bench mxfp8 all experiments exit $?" >> $S
LLMFP8_WEIGHT_PREFETCH=1 $B > gpurun_out/r02_bench_mxfp8_wprefetch.json 2> gpurun_out/r02_bench_mxfp8_wprefetch.err; echo "bench mxfp8 +weight prefetch exit $?" >> $S
LLMFP8_WEIGHT_PREFETCH=1 $B --scenario default > gpurun_out/r02_bench_default_wprefetch.json 2> gpurun_out/r02_bench_default_wprefetch.err; echo "bench default +weight prefetch exit $?" >> $S
LLMFP8_PDL=1 $B > gpurun_out/r02_bench_mxfp8_pdl.json 2> gpurun_out/r02_bench_mxfp8_pdl.err; echo "bench mxfp8 PDL build exit $?" >> $S
LLMFP8_PDL=1 $B --scenario default > gpurun_out/r02_bench_default_pdl.json 2> gpurun_out/r02_bench_default_pdl.err; echo "bench default PDL build exit $?" >> $S
LLMFP8_JOIN_SPLIT_WEIGHTS=0 $B > gpurun_out/r02_bench_mxfp8_catweights.json 2> gpurun_out/r02_bench_mxfp8_catweights.err; echo "bench mxfp8 torch.cat qkv weights exit $?" >> $S
LLMFP8_GEMM_DEBUG=16 $B > gpurun_out/r02_bench_mxfp8_singleacc.json 2> gpurun_out/r02_bench_mxfp8_singleacc.err; echo "bench mxfp8 single accumulator exit $?" >> $S
$B --scenario default > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err; echo "bench default exit $?" >> $S
LLMFP8_WGRAD_STREAM=1 $B --scenario default > gpurun_out/r02_bench_default_wgradstream.json 2> gpurun_out/r02_bench_default_wgradstream.err; echo "bench default +wgrad stream exit $?" >> $S

timeout 200 python tools/microbench.py --model llama8b --only quant --kfilter K3m --experimental --out gpurun_out/r02_micro_fused_act.jsonl > gpurun_out/r02_micro_fused_act.log 2>&1; echo "microbench fused act exit $?" >> $S
timeout 120 python tools/gemm_ab.py --modes 16,0 --rounds 4 > gpurun_out/r02_gemm_ab_ovl.txt 2>&1; echo "gemm A/B exit $?" >> $S

CMD="python bench.py --layers 2 --steps 1 --warmup 3 --no-cpu-baseline --no-roofline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
echo "ncu list exit $?" >> $S
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fp8_gemm_kernel -s 48 -c 6 -o gpurun_out/r02_prof_gemm $CMD > gpurun_out/r02_ncu_gemm.log 2>&1
echo "ncu gemm exit $?" >> $S

for f in gpurun_out/r02_bench_*.json; do python - "$f" <<'EOF' >> $S
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1].split('r02_bench_')[1]:40s} value {d['value']:8.1f} {d['unit']}  {d['ms_per_step']:7.2f} ms  e2e {d['e2e']['value']:8.1f}  sm {d['clocks']['sm_mhz']}")
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
EOF
done
cat $S
