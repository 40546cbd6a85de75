"""Data-parallel host logic (llm_fp8_b200/dp.py) on CPU: two gloo ranks, tokens sharded over ranks, the gradient all-reducer
must reproduce the single-process gradient of the concatenated batch, and the amax exchange must leave both ranks with the
same maximum (what fp8.ScalePool.flush does with NCCL on the GPU box).  Mirrors the reference's DDP contract
(train_multi_gpu.py:355-378: averaged gradients; :969-1006: env-var rendezvous)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_fp8_b200 import dp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _model(seed=0):
    torch.manual_seed(seed)
    # one weight above the reducer's direct threshold (in-place all-reduce) and small ones (packed bucket)
    return torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.SiLU(), torch.nn.Linear(128, 64, bias=True))


def _batch():
    g = torch.Generator().manual_seed(5)
    return torch.randn(6, 10, 64, generator=g)          # 6 sequences of 10 tokens


def _worker(rank, world, port, out, prescaled=False, bucket_bytes=0):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r, w, _ = dp.init_from_env(backend="gloo")
    assert (r, w) == (rank, world) and dist.get_backend() == "gloo"
    model = _model()
    reducer = dp.GradAllReducer(model.parameters(), direct_numel=64 * 128, prescaled=prescaled, bucket_bytes=bucket_bytes)
    assert reducer.prescaled == prescaled and reducer.loss_scale == (1.0 / world if prescaled else 1.0)
    x = _batch()
    mine = dp.shard_tokens(x.shape[0], rank, world)
    xs = x[mine.start:mine.stop]
    # mean loss over the LOCAL shard; DDP semantics average over ranks, which equals the global mean for equal shards
    (model(xs).square().mean() * reducer.loss_scale).backward()
    reducer.finish()
    grads = [p.grad.clone() for p in model.parameters()]
    # amax exchange: element-wise MAX of a flat fp32 buffer
    amax = torch.tensor([1.0 + rank, 5.0 - rank, 0.0])
    dist.all_reduce(amax, op=dist.ReduceOp.MAX)
    if rank == 0:
        torch.save({"grads": grads, "amax": amax, "bytes": reducer.bytes_reduced, "collectives": reducer.collectives}, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("prescaled", [False, True])
def test_two_rank_grad_allreduce_matches_single_process(tmp_path, prescaled):
    """prescaled: loss * 1/world then SUM (the reduction NVSwitch can run in the fabric) — same mean as AVG."""
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, prescaled), nprocs=2, join=True)
    got = torch.load(out)
    model = _model()
    model(_batch()).square().mean().backward()
    for g, p in zip(got["grads"], model.parameters()):
        assert torch.allclose(g, p.grad, rtol=1e-5, atol=1e-7)
    assert torch.equal(got["amax"], torch.tensor([2.0, 5.0, 0.0]))
    assert got["bytes"] > 0


def test_two_rank_grad_allreduce_coalesced_buckets(tmp_path):
    """bucket_bytes > 0: the large gradients (both Linear weights here) go out as one coalesced group instead of one collective each;
    same gradients as the single process."""
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), out, True, 1 << 30), nprocs=2, join=True)
    got = torch.load(out)
    model = _model()
    model(_batch()).square().mean().backward()
    for g, p in zip(got["grads"], model.parameters()):
        assert torch.allclose(g, p.grad, rtol=1e-5, atol=1e-7)
    assert got["collectives"] == 2          # one coalesced group for the two weights + the packed small gradients


def test_shard_tokens_partitions_exactly():
    for gb in (1, 7, 12, 16):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                seen += list(dp.shard_tokens(gb, r, world))
            assert seen == list(range(gb))
            sizes = [len(dp.shard_tokens(gb, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_single_process_is_a_noop(monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    assert dp.init_from_env()[:2] == (0, 1)
    m = _model()
    red = dp.GradAllReducer(m.parameters())
    m(_batch()).sum().backward()
    red.finish()
    assert red.bytes_reduced == 0


# ------------------------------------------------------------------------------------------------ joined qkv weights under the reducer
class _SplitProbe(torch.nn.Module):
    """The split query / key / value Parameters of LayerNormLinear (row slices of one buffer, modules._joined_rows) driving a plain
    matmul: what the gradient all-reducer sees on the FP8 path is the same — three Parameters whose gradients are views of one
    wgrad output."""

    def __init__(self):
        super().__init__()
        import llm_fp8_b200 as te
        torch.manual_seed(3)
        self.lnl = te.LayerNormLinear(64, 64 + 2 * 32, bias=False, normalization="RMSNorm",
                                      parameters_split={"query": 64, "key": 32, "value": 32}, params_dtype=torch.float32)

    def forward(self, x):
        w, _ = self.lnl._weight_bias()
        return x @ w.t()


def _split_worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dp.init_from_env(backend="gloo")
    model = _SplitProbe()
    params = [model.lnl.query_weight, model.lnl.key_weight, model.lnl.value_weight]
    reducer = dp.GradAllReducer(params, direct_numel=32 * 64, prescaled=True)      # every slice goes the direct (in-place) way
    x = _batch()
    mine = dp.shard_tokens(x.shape[0], rank, world)
    for _ in range(2):                                                             # two optimizer steps' worth: fresh gradient views each time
        for p in params:
            p.grad = None
        (model(x[mine.start:mine.stop]).square().mean() * reducer.loss_scale).backward()
        reducer.finish()
    if rank == 0:
        torch.save([p.grad.clone() for p in params], out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_allreduce_of_joined_split_weights(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_split_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    model = _SplitProbe()
    model(_batch()).square().mean().backward()
    want = [model.lnl.query_weight.grad, model.lnl.key_weight.grad, model.lnl.value_weight.grad]
    for g, w in zip(got, want):
        assert torch.allclose(g, w, rtol=1e-5, atol=1e-7)


# ------------------------------------------------------------------------------------------------ gradient accumulation (DDP.no_sync)
def _accum_worker(rank, world, port, out, misuse):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dp.init_from_env(backend="gloo")
    model = _model()
    reducer = dp.GradAllReducer(model.parameters(), direct_numel=64 * 128, prescaled=True)
    x = _batch()
    micro = [x[:4], x[4:]]                                  # two micro-batches of one optimizer step, each sharded over the ranks
    err = None
    try:
        for i, xb in enumerate(micro):
            mine = dp.shard_tokens(xb.shape[0], rank, world)
            loss = model(xb[mine.start:mine.stop]).square().sum() * reducer.loss_scale
            if misuse or i == len(micro) - 1:
                loss.backward()
                reducer.finish()
            else:
                with reducer.no_sync():
                    loss.backward()
    except RuntimeError as e:
        err = str(e)
    grads = [p.grad.clone() for p in model.parameters()]
    # second optimizer step after dropping the gradients: the guard must not fire
    if not misuse:
        model.zero_grad(set_to_none=True)
        (model(x[dp.shard_tokens(6, rank, world).start:dp.shard_tokens(6, rank, world).stop]).square().sum() * reducer.loss_scale).backward()
        reducer.finish()
        # zeroed in place: same tensors -> new_step() declares the new accumulation window
        model.zero_grad(set_to_none=False)
        reducer.new_step()
        (model(x[dp.shard_tokens(6, rank, world).start:dp.shard_tokens(6, rank, world).stop]).square().sum() * reducer.loss_scale).backward()
        reducer.finish()
    if rank == 0:
        torch.save({"grads": grads, "err": err, "collectives": reducer.collectives}, out)
    if err is None:
        dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gradient_accumulation_with_no_sync(tmp_path):
    """Two micro-batches per optimizer step: the first under no_sync() (hooks idle), the second reduces the accumulated gradient once ->
    world-mean of the summed micro-batch gradients == single-process gradient (prescaled SUM: loss * 1/world on every micro-batch)."""
    out = str(tmp_path / "r0.pt")
    mp.spawn(_accum_worker, args=(2, _free_port(), out, False), nprocs=2, join=True)
    got = torch.load(out)
    assert got["err"] is None
    model = _model()
    x = _batch()
    # every rank's loss is a SUM over its shard, scaled by 1/world: the reduced gradient is the global sum / world
    ((model(x[:4]).square().sum() + model(x[4:]).square().sum()) / 2).backward()
    for g, p in zip(got["grads"], model.parameters()):
        assert torch.allclose(g, p.grad, rtol=1e-5, atol=1e-6)


def test_accumulating_into_a_reduced_gradient_raises(tmp_path):
    """Without no_sync() the second backward would add its local gradient onto the already reduced one and reduce the sum again
    (world * G1 + G2): the reducer refuses."""
    out = str(tmp_path / "r0.pt")
    mp.spawn(_accum_worker, args=(2, _free_port(), out, True), nprocs=2, join=True)
    got = torch.load(out)
    assert got["err"] is not None and "no_sync" in got["err"]
