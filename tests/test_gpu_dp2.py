"""Data-parallel correctness of the FP8 path on TWO GPUs (the reference's contract: DDP-averaged gradients over
token-sharded ranks, train_multi_gpu.py:355-378, :1003).  Two NCCL ranks step a LinearPathStack under delayed scaling on their halves
of a global batch; asserted:
  (i)  after every step the pools' `scale` and `amax_history` are BIT-IDENTICAL on both ranks (the amax MAX exchange);
  (ii) the reduced weight gradients equal the single-GPU gradients on the concatenated batch — the scales are the same by (i), so the
       fp8 bytes are the same and only the summation order of dW differs (bf16 rounding of two partial sums): band 2^-6 of max|ref|;
  (iii) the same under MXFP8 (no amax exchange; block scales are local) and with gradient accumulation under no_sync().
Skipped on a one-GPU box; tools/gpu_dp2.sh runs it under `gpurun --gpus 2` and keeps the log under profiles/."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu
STEPS = 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batch(step, B, S, H):
    g = torch.Generator().manual_seed(500 + step)
    return (torch.randn(B, S, H, generator=g) * (1.0 + 0.5 * step)).bfloat16()


def _loss(y):
    return y.float().pow(2).mean()


def _run_single(model_name, scenario, layers, B, S, accum):
    """one process, whole global batch (in `accum` micro-batches): the gradients the data-parallel run must reproduce"""
    import llm_fp8_b200 as te
    from llm_fp8_b200 import fp8
    from llm_fp8_b200.layer import LinearPathStack
    te.reset_fp8_state()
    model = LinearPathStack(model_name, scenario, layers, device="cuda", seed=77)
    out = []
    for step in range(STEPS):
        model.zero_grad(set_to_none=True)
        x = _batch(step, B, S, model.H).cuda()
        for mb in x.chunk(accum):
            (_loss(model(mb)) / accum).backward()
        out.append([p.grad.detach().float().cpu() for p in model.parameters()])
    te.flush_all()
    state = [(p.scale[:p.used].cpu(), p.amax_history[:, :p.used].cpu()) for p in fp8.all_pools()]
    return out, state


def _worker(rank, world, port, out, model_name, scenario, layers, B, S, accum):
    import torch.distributed as dist

    import llm_fp8_b200 as te
    from llm_fp8_b200 import dp, fp8
    from llm_fp8_b200.layer import LinearPathStack
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dp.init_from_env("nccl")
    dev = torch.device("cuda", rank)
    te.reset_fp8_state()
    model = LinearPathStack(model_name, scenario, layers, device=dev, seed=77)          # same seed: replicated weights
    reducer = dp.GradAllReducer(model.parameters())
    res = {"grads": [], "state_equal": [], "names": [n for n, _ in model.named_parameters()]}
    for step in range(STEPS):
        model.zero_grad(set_to_none=True)
        x = _batch(step, B, S, model.H)
        mbs = x.chunk(accum)
        for i, mb in enumerate(mbs):
            mine = dp.shard_tokens(mb.shape[0], rank, world)
            loss = _loss(model(mb[mine.start:mine.stop].to(dev))) / accum * reducer.loss_scale
            if i < len(mbs) - 1:
                with reducer.no_sync():
                    loss.backward()
            else:
                loss.backward()
        reducer.finish()
        res["grads"].append([p.grad.detach().float().cpu() for p in model.parameters()])
        # (i) every rank holds the same delayed-scaling state once the step's exchange has been applied
        te.flush_all()
        same = True
        for p in fp8.all_pools():
            mine_state = torch.cat([p.scale[:p.used].reshape(-1), p.amax_history[:, :p.used].reshape(-1)]).contiguous()
            gathered = [torch.empty_like(mine_state) for _ in range(world)]
            dist.all_gather(gathered, mine_state)
            same = same and all(torch.equal(gathered[0].view(torch.int32), g.view(torch.int32)) for g in gathered)
        res["state_equal"].append(bool(same))
    res["state"] = [(p.scale[:p.used].cpu(), p.amax_history[:, :p.used].cpu()) for p in fp8.all_pools()]
    res["pools"] = len(fp8.all_pools())
    res["collectives"] = reducer.collectives
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("scenario,accum", [("hybrid", 1), ("default", 1), ("mxfp8", 1), ("hybrid", 2)])
def test_two_gpu_data_parallel_matches_single_gpu(tmp_path, scenario, accum):
    import torch.multiprocessing as mp
    model_name, layers, B, S = "llama-3.2-1b", 2, 4 * accum, 128
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, model_name, scenario, layers, B, S, accum), nprocs=2, join=True)
    got = torch.load(out)
    want, want_state = _run_single(model_name, scenario, layers, B, S, accum)
    assert all(got["state_equal"]), f"delayed-scaling state differs between ranks: {got['state_equal']}"
    if scenario != "mxfp8":
        assert got["pools"] >= 2
        # NOTE: with per-step flush_all() on the ranks but one flush at the end on the single GPU the update SCHEDULE is the same
        # (an update is applied before the next use either way), so the final states must agree bit for bit as well
        for (s0, h0), (s1, h1) in zip(want_state, got["state"]):
            assert torch.equal(s0, s1) and torch.equal(h0, h1), "scales after the MAX exchange != single-GPU scales on the concatenated batch"
    worst = 0.0
    for step in range(STEPS):
        for name, g, w in zip(got["names"], got["grads"][step], want[step]):
            err = (g - w).abs().max().item() / (w.abs().max().item() + 1e-30)
            worst = max(worst, err)
            assert err <= 2 ** -6, f"step {step} {name}: data-parallel gradient differs from the single-GPU one by {err:.3e} of max|ref|"
    print(f"[dp2] {scenario} accum={accum}: ranks' scaling state identical at every step; worst gradient deviation {worst:.3e} of max|ref| "
          f"({got['collectives']} collectives)")
