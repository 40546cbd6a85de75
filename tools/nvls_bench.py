"""The in-switch gradient reduce (csrc/nvls_reduce.cu) against NCCL's all-reduce(SUM) on the same box, alone (no GEMMs running):
time per call and bus bandwidth (2(n-1)/n * bytes / t) over gradient sizes from the Llama-3.1-8B layer, for a
few CTA counts.  One rank per GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 tools/nvls_bench.py [--out f.jsonl]

Timed on the device with CUDA events on the launching stream, 3 warm-up + 10 timed back-to-back calls, max over ranks.  Not a bench.py
number: the step-level effect (SMs, HBM traffic and power taken from the GEMMs) is what tools/gpu_round2_scale.sh measures."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fp8_b200 import dp  # noqa: E402

SHAPES = {"k_or_v [1024, 4096]": (1024, 4096), "proj [4096, 4096]": (4096, 4096), "qkv [6144, 4096]": (6144, 4096),
          "fc2 [4096, 14336]": (4096, 14336), "fc1 [28672, 4096]": (28672, 4096)}


def timed(fn, stream, iters=10, warm=3):
    for _ in range(warm):
        fn()
    stream.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(iters):
        fn()
    e1.record(stream)
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3      # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--ctas", default="8,32,96")
    a = ap.parse_args()
    rank, world, local = dp.init_from_env(backend="nccl")
    dev = torch.device("cuda", local)
    ws = [torch.nn.Parameter(torch.zeros(s, dtype=torch.bfloat16, device=dev)) for s in SHAPES.values()]
    lines = []
    stream = torch.cuda.current_stream()
    for ctas in [int(c) for c in a.ctas.split(",")]:
        arena = dp.NvlsGradArena([(w, [w]) for w in ws], max_ctas=ctas, timeout_ms=20000)
        for (name, shape), home in zip(SHAPES.items(), arena.homes):
            nbytes = home.numel() * 2
            where = arena.locate(home)
            us_nvls = timed(lambda: arena.reduce(where, home, stream), stream)
            rec = {"gradient": name, "bytes": nbytes, "n_gpus": world, "nvls_ctas": ctas, "us_nvls": round(us_nvls, 1),
                   "busbw_nvls_GBps": round(2 * (world - 1) / world * nbytes / us_nvls / 1e3, 1)}
            if ctas == int(a.ctas.split(",")[0]):
                plain = torch.zeros(shape, dtype=torch.bfloat16, device=dev)
                us_nccl = timed(lambda: dist.all_reduce(plain), stream)
                rec.update(us_nccl=round(us_nccl, 1), busbw_nccl_GBps=round(2 * (world - 1) / world * nbytes / us_nccl / 1e3, 1))
            lines.append(rec)
        arena.check()
        arena.release()
        del arena
    if rank == 0:
        text = "\n".join(json.dumps(r) for r in lines)
        print(text)
        if a.out:
            with open(a.out, "w") as f:
                f.write(text + "\n")
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
