"""In-switch gradient reduce (csrc/nvls_reduce.cu, dp.NvlsGradArena, functional.wgrad_out) — everything that can be checked without a
GPU: the C ABI's argument contract, the arena plan, the address lookup, the slice walk of the kernel (every 16-byte vector reduced by
exactly one thread of exactly one rank), the cross-rank meeting protocol over consecutive launches (exhaustive interleavings of a
two-rank model), the wgrad-output hand-over to autograd, and the reducer's routing with two gloo ranks.  Stands in for hardware runs
until tests/test_gpu_comm.py::test_nvls_* has run on two B200s.  Reference behaviour replaced: DDP's gradient all-reduce
(train_multi_gpu.py:355-378)."""
import ctypes as C
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_fp8_b200 import _lib, dp
from llm_fp8_b200 import functional as fn

THREADS, UNROLL = 128, 4          # kThreads / kUnroll of csrc/nvls_reduce.cu


# ------------------------------------------------------------------------------------------------ C ABI contract (no launch happens)
def test_cabi_rejects_bad_arguments_before_launching():
    lib = _lib.load()
    pads = C.c_void_p(0x1000)
    call = lib.llmfp8_nvls_allreduce_sum_bf16
    assert call(0x1000, 0, pads, 9216, 0, 2, 0, 0, None) == 0                 # empty gradient: nothing to do
    assert call(0x1000, 64, pads, 9216, 0, 1, 0, 0, None) == -1               # world 1 has no peers
    assert call(0x1000, 64, pads, 9216, 2, 2, 0, 0, None) == -1               # rank out of range
    assert call(0x1000, 12, pads, 9216, 0, 2, 0, 0, None) == -1               # not whole 16-byte vectors
    assert call(None, 64, pads, 9216, 0, 2, 0, 0, None) == -1                 # no multicast address
    assert call(0x1008, 64, pads, 9216, 0, 2, 0, 0, None) == -1               # misaligned
    assert call(0x1000, 64, None, 9216, 0, 2, 0, 0, None) == -1               # no signal pads
    assert call(0x1000, 1 << 30, pads, 32, 0, 8, 0, 0, None) == -1            # pad too small for 32 CTAs x 8 ranks
    assert b"signal pad" in lib.llmfp8_last_error()


def test_cabi_amax_update_rejects_bad_arguments_before_launching():
    lib = _lib.load()
    pads = C.c_void_p(0x1000)
    call = lib.llmfp8_nvls_amax_scale_update
    ok = (0x2000, 0x3000, 0x4000, 0x5000, 0x6000)          # history, multicast row 0, scale, scale_inv, fmax
    assert call(*ok, 16, 0, 0, pads, 9216, 0, 2, 0, None) == 0                 # no slots: nothing to do
    assert call(*ok, 0, 64, 0, pads, 9216, 0, 2, 0, None) == -1                # empty history
    assert call(*ok, 16, 64, 200, pads, 9216, 0, 2, 0, None) == -1             # margin out of range
    assert call(*ok, 16, 64, 0, pads, 9216, 0, 1, 0, None) == -1               # world 1 has no peers
    assert call(0x2000, None, 0x4000, 0x5000, 0x6000, 16, 64, 0, pads, 9216, 0, 2, 0, None) == -1     # no multicast address
    assert call(0x2000, 0x3002, 0x4000, 0x5000, 0x6000, 16, 64, 0, pads, 9216, 0, 2, 0, None) == -1   # misaligned
    assert call(None, 0x3000, 0x4000, 0x5000, 0x6000, 16, 64, 0, pads, 9216, 0, 2, 0, None) == -1     # no state
    assert call(*ok, 16, 64, 0, None, 9216, 0, 2, 0, None) == -1               # no signal pads
    assert call(*ok, 16, 4096, 0, pads, 64, 0, 8, 0, None) == -1               # 32 CTAs x 8 ranks do not fit a 64-byte pad


def test_pools_stay_in_plain_memory_unless_asked(monkeypatch):
    from llm_fp8_b200 import fp8
    from llm_fp8_b200.recipe import DelayedScaling
    assert fp8.NVLS_AMAX is False                          # opt-in only (LLMFP8_NVLS_AMAX=1)
    fp8.reset_fp8_state()
    pool = fp8.get_pool(DelayedScaling(amax_history_len=4), True, torch.device("cpu", 0))
    assert pool._nvls is None and pool.amax_history.shape == (4, fp8._POOL_CAP)
    fp8.reset_fp8_state()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_grid_is_the_same_function_on_every_rank(world):
    lib = _lib.load()
    for numel in (8, 4096, 1 << 20, 28672 * 4096):
        for cap in (0, 1, 4, 16):
            g = lib.llmfp8_nvls_grid(numel, world, cap)
            per = -(-(numel // 8) // world)
            want = max(1, min(cap or 32, -(-per // (THREADS * UNROLL))))
            assert g == want


# ------------------------------------------------------------------------------------------------ the kernel's slice walk
def _walk(nvec, world, ctas):
    """replays nvls_allreduce_sum_bf16_kernel's index arithmetic -> how often each vector is reduced, and by which rank"""
    hits = [0] * nvec
    owner = [-1] * nvec
    per = -(-nvec // world)
    stride = ctas * THREADS
    for rank in range(world):
        lo, hi = per * rank, min(nvec, per * rank + per)
        for b in range(ctas):
            for t in range(THREADS):
                i = lo + b * THREADS + t
                while i < hi:
                    for u in range(UNROLL):
                        j = i + u * stride
                        if j < hi:
                            hits[j] += 1
                            owner[j] = rank
                    i += stride * UNROLL
    return hits, owner


@pytest.mark.parametrize("nvec,world,ctas", [(1, 2, 1), (7, 2, 1), (513, 2, 1), (4099, 4, 2), (20000, 8, 3), (2048 * 3 + 5, 3, 1)])
def test_every_vector_is_reduced_exactly_once(nvec, world, ctas):
    hits, owner = _walk(nvec, world, ctas)
    assert hits == [1] * nvec
    assert owner == sorted(owner)                       # contiguous slices in rank order


# ------------------------------------------------------------------------------------------------ the meeting protocol
def test_meeting_protocol_two_ranks_consecutive_launches_all_interleavings():
    """Each rank runs, per launch: put(entry) wait(entry) [work] put(exit) wait(exit); put = CAS 0 -> 1 on the PEER's slot for me,
    wait = CAS 1 -> 0 on MY slot for the peer; a failed CAS retries.  Explores every interleaving of two ranks over three launches
    (launch n+1 of a fast rank overlapping launch n of a slow one): the work phases of launch n never start before both ranks have
    entered launch n, no rank leaves launch n's exit before both finished their work, and no state deadlocks."""
    launches = 3
    prog_len = 4 * launches            # steps per rank: 0 put_e, 1 wait_e, 2 put_x, 3 wait_x (work happens between 1 and 2)
    start = ((0, 0), (0, 0))           # (pc of rank 0, pc of rank 1), (slot on rank0 written by rank1, slot on rank1 written by rank0)
    seen, todo = {start}, [start]
    while todo:
        (pcs, slots) = todo.pop()
        if pcs == (prog_len, prog_len):
            assert slots == (0, 0)     # pads are clean again: the next launch can reuse them
            continue
        moved = False
        for r in (0, 1):
            pc = pcs[r]
            if pc == prog_len:
                continue
            kind = pc % 4
            s = list(slots)
            if kind in (0, 2):         # put on the peer's slot (index 1 - r holds "slot on peer written by r")
                idx = 1 - r
                if s[idx] != 0:
                    continue           # CAS fails, retry later
                s[idx] = 1
            else:                      # wait on my slot
                idx = r
                if s[idx] != 1:
                    continue
                s[idx] = 0
            npcs = list(pcs)
            npcs[r] = pc + 1
            # invariants at the transitions that matter
            if kind == 1:              # r passes the entry meeting of launch pc // 4: the peer must have entered that launch
                assert pcs[1 - r] >= (pc // 4) * 4 + 1
            if kind == 3:              # r passes the exit meeting: the peer must have finished its work (issued its exit put)
                assert pcs[1 - r] >= (pc // 4) * 4 + 3
            nxt = (tuple(npcs), tuple(s))
            moved = True
            if nxt not in seen:
                seen.add(nxt)
                todo.append(nxt)
        assert moved, f"deadlock at {pcs} {slots}"
    assert len(seen) > 4 * launches          # the walk did branch (37 states for three launches)


# ------------------------------------------------------------------------------------------------ arena plan / lookup
def test_plan_arena_is_aligned_ordered_and_bounded():
    sizes = [6144 * 4096 * 2, 4096 * 4096 * 2, 28672 * 4096 * 2, 4096 * 14336 * 2] * 3 + [24]
    place, chunks = dp.plan_arena(sizes, 1 << 28)
    assert len(place) == len(sizes)
    ends = {}
    for (ci, off), n in zip(place, sizes):
        assert off % 256 == 0 and off >= ends.get(ci, 0)          # no overlap, in order
        ends[ci] = off + n
        assert ends[ci] <= chunks[ci]
    assert all(c <= (1 << 28) or sum(1 for ci, _ in place if ci == i) == 1 for i, c in enumerate(chunks))   # oversize only when alone
    assert dp.plan_arena(sizes, 1 << 28) == (place, chunks)       # deterministic: every rank derives the same layout


def test_locate_in_chunks():
    bases, sizes = [0x10000, 0x80000], [0x1000, 0x2000]
    assert dp.locate_in_chunks(bases, sizes, 0x10000, 0x1000) == (0, 0)
    assert dp.locate_in_chunks(bases, sizes, 0x80100, 64) == (1, 0x100)
    assert dp.locate_in_chunks(bases, sizes, 0x10ff0, 32) is None          # runs over the end
    assert dp.locate_in_chunks(bases, sizes, 0x8000, 16) is None           # before the first chunk
    assert dp.locate_in_chunks(bases, sizes, 0x20000, 16) is None          # between chunks
    assert dp.locate_in_chunks(bases, sizes, 0x10008, 16) is None          # not 16-byte aligned inside the chunk


# ------------------------------------------------------------------------------------------------ wgrad output hand-over
class _WgradProbe(torch.autograd.Function):
    """what the linear Functions do with functional.wgrad_out: the weight gradient is written into the registered home (or a fresh
    tensor) and returned to autograd"""

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(x, w)
        ctx.wkey, ctx.shape = w.data_ptr(), tuple(w.shape)
        return x @ w.t()

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        out = fn.wgrad_out(ctx.wkey, *ctx.shape, torch.float32)
        dw = dy.t() @ x
        if out is not None:
            out.copy_(dw)
            dw = out
        return dy @ w, dw


def test_wgrad_home_becomes_the_gradient_without_a_copy_and_is_protected_while_live():
    w = torch.nn.Parameter(torch.randn(8, 16))
    arena = torch.zeros(8 * 16 + 32)
    home = arena[32:32 + 8 * 16]
    fn._WGRAD_OUT[w.data_ptr()] = (home, [w])
    try:
        x = torch.randn(4, 16)
        _WgradProbe.apply(x, w).sum().backward()
        want = torch.ones(4, 8).t() @ x
        assert w.grad.data_ptr() == home.data_ptr()                 # autograd adopted the arena view
        assert torch.equal(w.grad, want) and torch.equal(home.view(8, 16), want)
        # second micro-batch while .grad is live: the home must not be overwritten; autograd accumulates into it
        assert fn.wgrad_out(w.data_ptr(), 8, 16, torch.float32) is None
        _WgradProbe.apply(x, w).sum().backward()
        assert torch.allclose(w.grad, 2 * want)
        w.grad = None
        assert fn.wgrad_out(w.data_ptr(), 8, 16, torch.float32) is not None
        assert fn.wgrad_out(w.data_ptr(), 8, 16, torch.bfloat16) is None     # dtype mismatch -> plain allocation
        assert fn.wgrad_out(w.data_ptr(), 4, 16, torch.float32) is None      # shape mismatch
        assert fn.wgrad_out(12345, 8, 16, torch.float32) is None
    finally:
        fn._WGRAD_OUT.pop(w.data_ptr(), None)


def test_gemm_weight_buffers_lists_every_wgrad_output():
    from llm_fp8_b200.layer import DecoderLinearPath, recipes_for
    ar, mr = recipes_for("mxfp8")
    blk = DecoderLinearPath(64, 128, 4, 2, 16, ar, mr, device="cpu")
    bufs = dp.gemm_weight_buffers(blk)
    shapes = sorted(tuple(w.shape) for w, _ in bufs)
    assert shapes == sorted([(64 + 2 * 32, 64), (64, 64), (256, 64), (64, 128)])
    qkv = [ps for w, ps in bufs if len(ps) == 3]
    assert len(qkv) == 1 and qkv[0][0].data_ptr() == [w for w, ps in bufs if len(ps) == 3][0].data_ptr()   # joined buffer starts at query_weight
    covered = {id(p) for _, ps in bufs for p in ps}
    big = {id(p) for p in blk.parameters() if p.dim() == 2}
    assert covered == big


# ------------------------------------------------------------------------------------------------ reducer routing (two gloo ranks)
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _NoStream:
    def wait_stream(self, other):
        pass


class _FakeArena:
    """NvlsGradArena's face towards GradAllReducer, with the in-switch kernel replaced by a gloo all-reduce: checks which gradients
    are routed to it, that they are the arena's memory, and that finish() joins."""

    def __init__(self, homes):
        self.homes = homes
        self.reduced = []
        self.polls = 0
        self._s = _NoStream()

    def poll(self):
        self.polls += 1

    def current_stream(self):
        return self._s

    def new_stream(self):
        return _NoStream()

    def locate(self, g):
        for i, h in enumerate(self.homes):
            if h.data_ptr() <= g.data_ptr() and g.data_ptr() + g.numel() * 4 <= h.data_ptr() + h.numel() * 4:
                return i, g.data_ptr() - h.data_ptr()
        return None

    def reduce(self, where, g, stream):
        self.reduced.append(where)
        dist.all_reduce(g)


def _route_worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dp.init_from_env(backend="gloo")
    torch.manual_seed(0)
    w_in = torch.nn.Parameter(torch.randn(32, 16))       # has a home in the arena
    w_out = torch.nn.Parameter(torch.randn(8, 32))       # has none: NCCL / gloo path
    home = torch.zeros(32 * 16)
    fn._WGRAD_OUT[w_in.data_ptr()] = (home, [w_in])
    arena = _FakeArena([home])
    red = dp.GradAllReducer([w_in, w_out], direct_numel=64, prescaled=True, nvls=arena)
    x = torch.randn(6, 16, generator=torch.Generator().manual_seed(7))[rank * 3:rank * 3 + 3]
    h = _WgradProbe.apply(x, w_in)
    (_WgradProbe.apply(h, w_out).square().mean() * red.loss_scale).backward()
    red.finish()
    assert arena.polls == 1           # the timed-out-meeting counter is looked at once per step
    if rank == 0:
        torch.save({"g_in": w_in.grad.clone(), "g_out": w_out.grad.clone(), "in_home": w_in.grad.data_ptr() == home.data_ptr(),
                    "routed": arena.reduced, "launches": red.nvls_launches, "collectives": red.collectives}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_reducer_routes_arena_gradients_to_the_in_switch_path(tmp_path):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_route_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    w_in = torch.nn.Parameter(torch.randn(32, 16))
    w_out = torch.nn.Parameter(torch.randn(8, 32))
    x = torch.randn(6, 16, generator=torch.Generator().manual_seed(7))
    # the two ranks each take the mean over their 3 rows; the average of the two is the mean over all 6
    ((x @ w_in.t()) @ w_out.t()).square().mean().backward()
    assert got["in_home"] and got["routed"] == [(0, 0)] and got["launches"] == 1
    assert got["collectives"] == 2                                   # one in-switch reduce + one NCCL-path all-reduce
    assert torch.allclose(got["g_in"], w_in.grad, rtol=1e-4, atol=1e-6)
    assert torch.allclose(got["g_out"], w_out.grad, rtol=1e-4, atol=1e-6)


def test_reducer_refuses_avg_with_the_in_switch_path(monkeypatch):
    class W:                                    # world > 1 without a process group
        pass
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    with pytest.raises(RuntimeError, match="plain SUM"):
        dp.GradAllReducer([], prescaled=False, nvls=_FakeArena([]))
