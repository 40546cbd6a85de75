"""The C ABI's own NCCL collectives (llmfp8_nccl_*) on two GPUs: amax MAX all-reduce and bf16 gradient SUM all-reduce, in place,
against values computed by hand.  Skipped on a one-GPU box (the kernels-only CI box); tools/gpu_comm2.sh runs it under gpurun --gpus 2."""
import os
import socket

import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    import torch.distributed as dist

    from llm_fp8_b200 import dp
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)          # only to hand the NCCL id round
    comm = dp.CAbiComm(rank, world, torch.device("cuda", rank))
    amax = torch.tensor([1.0 + rank, 5.0 - rank, 0.0, 3.5], device="cuda")
    comm.all_reduce_max_f32(amax)
    g = torch.Generator().manual_seed(3)
    base = torch.randn(1 << 20, generator=g)
    grad = (base * (rank + 1)).bfloat16().cuda()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    comm.all_reduce_sum_bf16(grad, stream=side)                             # on a side stream, as the overlapped reducer would
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    want = sum((base * (r + 1)).bfloat16().float() for r in range(world))
    ok_sum = torch.allclose(grad.float().cpu(), want, rtol=2 ** -7, atol=1e-6)
    ok_max = torch.equal(amax.cpu(), torch.tensor([float(world), 5.0, 0.0, 3.5]))
    comm.destroy()
    if rank == 0:
        torch.save({"ok_sum": ok_sum, "ok_max": ok_max}, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_cabi_nccl_collectives_two_gpus(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    assert got["ok_max"] and got["ok_sum"], got


def test_cabi_nccl_single_rank_roundtrip():
    """world = 1: the communicator comes up on one GPU and both collectives are identities."""
    from llm_fp8_b200 import dp
    comm = dp.CAbiComm(0, 1, torch.device("cuda", 0))
    a = torch.tensor([2.0, 0.5], device="cuda")
    g = torch.arange(64, device="cuda").bfloat16()
    comm.all_reduce_max_f32(a)
    comm.all_reduce_sum_bf16(g)
    torch.cuda.synchronize()
    assert a.tolist() == [2.0, 0.5] and torch.equal(g.cpu(), torch.arange(64).bfloat16())
    comm.destroy()


# ------------------------------------------------------------------------------------------------ in-switch gradient reduce (NVLS)
def _nvls_worker(rank, world, port, out):
    """(1) llmfp8_nvls_allreduce_sum_bf16 on arena homes of awkward sizes, launched back to back on a side stream (signal pads reused),
    against the sum computed on the host; (2) a 2-block FP8 stack stepped once with the NCCL reducer and once with the in-switch
    reducer: same weight gradients (bf16 rounding of a different summation order apart), and the gradients live in the arena."""
    import torch.distributed as dist

    from llm_fp8_b200 import dp
    from llm_fp8_b200.layer import LinearPathStack
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                      LLMFP8_DIST_TIMEOUT_S="60")
    dp.init_from_env(backend="nccl")
    dev = torch.device("cuda", rank)
    res = {}
    # -- (1) the kernel alone
    shapes = [(1, 8), (3, 40), (257, 1032), (4096, 4096)]
    ws = [torch.nn.Parameter(torch.empty(s, dtype=torch.bfloat16, device=dev)) for s in shapes]
    arena = dp.NvlsGradArena([(w, [w]) for w in ws], chunk_bytes=1 << 20, timeout_ms=20000)
    side = torch.cuda.Stream()
    ok = True
    for rep in range(3):                                  # pads must be clean again after every launch
        gen = torch.Generator().manual_seed(11 + rep)
        bases = [torch.randn(s, generator=gen) for s in shapes]
        for home, b in zip(arena.homes, bases):
            home.copy_((b * (rank + 1)).bfloat16().reshape(-1))
        side.wait_stream(torch.cuda.current_stream())
        for home in arena.homes:
            where = arena.locate(home)
            assert where is not None
            arena.reduce(where, home, side)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        dist.barrier()
        for home, b in zip(arena.homes, bases):
            want = sum((b * (r + 1)).bfloat16().float() for r in range(world)).reshape(-1)
            ok = ok and torch.allclose(home.float().cpu(), want, rtol=2 ** -7, atol=1e-6)
        dist.barrier()
    arena.check()
    arena.release()
    res["kernel_ok"] = ok
    # -- (2) the reducer on the FP8 path
    grads = {}
    for mode in ("nccl", "nvls"):
        model = LinearPathStack("llama-3.2-1b", "mxfp8", 2, device=dev)
        ar = dp.NvlsGradArena(dp.gemm_weight_buffers(model), timeout_ms=20000) if mode == "nvls" else None
        red = dp.GradAllReducer(model.parameters(), nvls=ar)
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        x = torch.randn(2, 256, model.H, device=dev, generator=g).bfloat16().requires_grad_(True)
        (model(x).float().square().mean() * red.loss_scale).backward()
        red.finish()
        torch.cuda.synchronize()
        grads[mode] = [p.grad.float().cpu() for p in model.parameters() if p.dim() == 2]
        if ar is not None:
            ar.check()
            res["in_arena"] = all(ar.locate(p.grad) is not None for p in model.parameters() if p.dim() == 2)
            res["launches"] = red.nvls_launches
            ar.release()
        red.remove()
        dist.barrier()
    # -- (3) the amax exchange fused with the scale update (LLMFP8_NVLS_AMAX): against the oracle's update of the elementwise maximum
    from llm_fp8_b200 import fp8
    from llm_fp8_b200.recipe import DelayedScaling, Format
    from oracle import fp8_oracle as fo
    fp8.reset_fp8_state()
    fp8.NVLS_AMAX = True
    try:
        r = DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=4)
        pool = fp8.get_pool(r, True, dev)
        assert pool._nvls is not None
        off = pool.claim(5)
        gen = torch.Generator().manual_seed(5)
        want_hist = torch.zeros(4, 5)
        want_scale, want_inv, fmax = torch.ones(5), torch.ones(5), torch.full((5,), 448.0)
        ok3 = True
        for it in range(6):                                # more updates than the history is long: the roll wraps
            per_rank = [torch.rand(5, generator=gen) * (10.0 ** (it - 2)) for _ in range(world)]
            per_rank[it % world][2] = 0.0
            pool.amax_history[0, off:off + 5] = per_rank[rank].to(dev)
            pool.pending = True
            pool.flush()
            torch.cuda.synchronize()
            want_hist[0] = torch.stack(per_rank).max(0).values
            fo.amax_scale_update(want_hist, want_scale, want_inv, fmax, 0)
            ok3 = ok3 and torch.equal(pool.scale[off:off + 5].cpu(), want_scale) and torch.equal(pool.scale_inv[off:off + 5].cpu(), want_inv) \
                and torch.equal(pool.amax_history[:, off:off + 5].cpu(), want_hist)
            dist.barrier()
        res["amax_ok"] = ok3
        n = __import__("ctypes").c_uint(0)
        from llm_fp8_b200 import _lib
        _lib.check(_lib.load().llmfp8_nvls_timeouts(__import__("ctypes").byref(n)), "nvls_timeouts")
        res["amax_timeouts"] = n.value
    finally:
        fp8.NVLS_AMAX = False
        fp8.reset_fp8_state()
    res["grad_err"] = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-30)) for a, b in zip(grads["nvls"], grads["nccl"]))
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_nvls_gradient_reduce_two_gpus(tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r0.pt")
    mp.spawn(_nvls_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    assert got["kernel_ok"], got
    assert got["in_arena"] and got["launches"] == 2 * 6, got          # per block: q, k, v (row slices of one home), proj, fc1, fc2
    assert got["grad_err"] <= 2 ** -6, got
    assert got["amax_ok"] and got["amax_timeouts"] == 0, got
