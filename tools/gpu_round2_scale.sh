#!/bin/bash
# 8-GPU probes for the >= 7x weak-scaling target (round 1: 6.87x).  ~2 min of box time per line, charged 8x:
#   /usr/local/graft/bin/gpurun --gpus 8 --timeout 1500 -- 'bash tools/gpu_round2_scale.sh'
# Run AFTER tools/gpu_round2_first.sh has validated the switches on one GPU; drop the lines whose switch failed there.
N=${1:-8}
mkdir -p gpurun_out
S=gpurun_out/r02_scale_summary.txt
: > $S
# Every line costs ~2 min x 8 GPUs = 16 GPU-minutes of the round's 180: run a SHORT list.  LINES="tag tag ..." selects (default below);
# LINES=all runs everything (only with budget to spare).
LINES=${LINES:-"base nvls32 nvls96 bucket256 cfg3b strong"}
run() {  # tag, env assignments...
  tag=$1; shift
  case " $LINES " in *" all "*|*" $tag "*) ;; *) return 0;; esac
  env "$@" LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 4 --warmup 3 --scenario mxfp8 --no-roofline > gpurun_out/r02_scale_${tag}_n$N.json 2> gpurun_out/r02_scale_${tag}_n$N.err
  echo "$tag n=$N exit $?" >> $S
  python - >> $S <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_scale_${tag}_n$N.json").read().strip().splitlines()[-1])
    tl = d.get("timeline") or {}
    print("   ${tag}", round(d["value"]), "TF total", round(d["ms_per_step"], 1), "ms/step  e2e", round(d["e2e"]["value"]),
          "fwd/bwd/tail ms", [round(tl.get(k, 0), 1) for k in ("forward_ms", "backward_ms", "reduce_tail_ms")], d["clocks"])
except Exception as e:
    print("   ${tag}: parse failed", e)
PY
}
# single-GPU reference on the same box (ratio denominators must come from the same box)
timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-roofline > gpurun_out/r02_scale_base_n1.json 2> gpurun_out/r02_scale_base_n1.err; echo "base n=1 exit $?" >> $S
python - >> $S <<PY
import json
d = json.loads(open("gpurun_out/r02_scale_base_n1.json").read().strip().splitlines()[-1])
tl = d.get("timeline") or {}
print("   n=1", round(d["value"]), "TF", round(d["ms_per_step"], 1), "ms/step  fwd/bwd/tail ms", [round(tl.get(k, 0), 1) for k in ("forward_ms", "backward_ms", "reduce_tail_ms")], d["clocks"])
PY
run base LLMFP8_NOOP=1
run pdl LLMFP8_PDL=1                                   # launch gaps matter more when NCCL kernels interleave
run nowgradstream LLMFP8_WGRAD_STREAM=0
run cga2 NCCL_CGA_CLUSTER_SIZE=2                       # NCCL CTAs in pairs: fewer TPCs half-occupied (our GEMM needs both SMs of a TPC)
run cga1 NCCL_CGA_CLUSTER_SIZE=1
run cg1 LLMFP8_GEMM_FORCE_CG=1                         # single-CTA GEMM tiles: do they coexist better with the all-reduce kernels?
run bucket256 LLMFP8_GRAD_BUCKET_MB=256                # coalesced gradient groups: ~2 NCCL launches per decoder layer instead of 6
run bucket512 LLMFP8_GRAD_BUCKET_MB=512                # ~1 per layer
run bucket128 LLMFP8_GRAD_BUCKET_MB=128
run bucket1024 LLMFP8_GRAD_BUCKET_MB=1024
run bucket0 LLMFP8_GRAD_BUCKET_MB=0                    # one collective per gradient (the round-1 behaviour)
# in-switch gradient reduce (own multimem kernel, csrc/nvls_reduce.cu) — only after tools/gpu_nvls2.sh passed on two GPUs
run nvls32 LLMFP8_NVLS_REDUCE=1
run nvls8 LLMFP8_NVLS_REDUCE=1 LLMFP8_NVLS_CTAS=8
run nvls96 LLMFP8_NVLS_REDUCE=1 LLMFP8_NVLS_CTAS=96
# BASELINE.json configs[2]: Llama-3.2-3B, te_llama_hybrid recipe (delayed scaling: the amax MAX exchange is on the path), batch 16 x 512 per GPU
run2() {  # tag, n, extra bench args...
  tag=$1; n=$2; shift 2
  case " $LINES " in *" all "*|*" $tag "*) ;; *) return 0;; esac
  if [ "$n" = 1 ]; then
    timeout 200 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-roofline "$@" > gpurun_out/r02_scale_${tag}_n$n.json 2> gpurun_out/r02_scale_${tag}_n$n.err
  else
    LLMFP8_DIST_TIMEOUT_S=90 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29513 \
        bench.py --gpus $n --steps 4 --warmup 3 --no-roofline "$@" > gpurun_out/r02_scale_${tag}_n$n.json 2> gpurun_out/r02_scale_${tag}_n$n.err
  fi
  echo "$tag n=$n exit $?" >> $S
  python - >> $S <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_scale_${tag}_n$n.json").read().strip().splitlines()[-1])
    tl = d.get("timeline") or {}
    print("   ${tag} n=$n", round(d["value"]), "TF total", round(d["ms_per_step"], 1), "ms/step", d["scaling"], d["config"]["workload"][:60],
          "fwd/bwd/tail ms", [round(tl.get(k, 0), 1) for k in ("forward_ms", "backward_ms", "reduce_tail_ms")])
except Exception as e:
    print("   ${tag}: parse failed", e)
PY
}
run2 cfg3b 1 --model llama-3.2-3b --scenario hybrid --batch 16
run2 cfg3b $N --model llama-3.2-3b --scenario hybrid --batch 16
# strong scaling: 24 sequences in total, 24 on one GPU against 24 / N per GPU
run2 strong 1 --global-batch 24
run2 strong $N --global-batch 24
# whole data-parallel step (collectives included) replayed from one CUDA graph: no host launch jitter between ranks
case " $LINES " in *" all "*|*" graphdp "*) ;; *) cat $S; exit 0;; esac
tag=graphdp; env LLMFP8_GRAPH_DP=1 LLMFP8_DIST_TIMEOUT_S=90 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 4 --warmup 3 --scenario mxfp8 --no-roofline --graph 1 > gpurun_out/r02_scale_${tag}_n$N.json 2> gpurun_out/r02_scale_${tag}_n$N.err
echo "$tag n=$N exit $?  $(tail -c 300 gpurun_out/r02_scale_${tag}_n$N.json | tr -d '\n' | grep -o '"ms_per_step": [0-9.]*')" >> $S
cat $S
