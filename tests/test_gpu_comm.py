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
