"""Stream ordering of the side-stream code paths, checked on CPU with a happens-before model of CUDA streams and events (tests/_streams.py):
the wgrad side stream (functional._SideWgrad, LLMFP8_WGRAD_STREAM), and the in-switch gradient reducer's side stream (dp.GradAllReducer(nvls=...)).  Every kernel launch of the
MXFP8 decoder path (the torch stand-ins of tests/test_autograd_wiring_cpu.py, wrapped to log what they read and write and on which
stream) must be ordered after the launch that produced its operands, and the main stream must have joined everything wherever
untracked work (autograd, the optimizer) carries on.  A negative control removes one join and the checker must object — so a
pass means something.  These paths have not run on hardware yet; this is what stands in for it."""
import pytest
import torch
import torch.distributed as dist

from _streams import fake_cuda, tensors_in
from test_autograd_wiring_cpu import H, _FakeOps, _layer

from llm_fp8_b200 import dp, ops
from llm_fp8_b200 import functional as fn

TRACKED = ("rmsnorm_rsigma", "mx_quantize", "mxfp8_gemm", "rmsnorm_bwd", "swiglu_bf16", "dswiglu_bf16", "mx_quantize_bgrad")


@pytest.fixture
def tracked_ops(monkeypatch):
    tracker = fake_cuda(monkeypatch)
    monkeypatch.setattr(ops, "rows_ok", _FakeOps.rows_ok)
    monkeypatch.setattr(fn, "MX_FUSED_ACT", False)          # the stand-ins cover the two-kernel activation sequence
    monkeypatch.setattr(fn, "MX_EPILOGUE_RESIDUAL", False)
    for name in TRACKED:
        real = getattr(_FakeOps, name)

        def wrapped(*a, _real=real, _name=name, **kw):
            out = _real(*a, **kw)
            reads = list(tensors_in(a)) + [t for k, v in kw.items() if k != "out" for t in tensors_in(v)]
            tracker.op(_name, reads, list(tensors_in(out)))
            return out
        monkeypatch.setattr(ops, name, wrapped)
    return tracker


def _stack(n):
    layers = torch.nn.ModuleList(_layer(True) for _ in range(n))
    return layers


def _step(layers, tracker, x, seed_grad=True):
    h = x.clone().requires_grad_(True)
    y = h
    for layer in layers:
        y = layer(y)
    tracker.assert_joined()                              # forward ends: the loss (untracked torch kernels) runs on main
    y.float().square().mean().backward()
    tracker.assert_joined()                              # backward ends: the optimizer reads the gradients on main
    return y


def test_wgrad_side_stream_orders_every_operand(tracked_ops, monkeypatch):
    monkeypatch.setattr(fn, "WGRAD_SIDE_STREAM", True)
    layers = _stack(2)
    x = torch.randn(2, 4, H, generator=torch.Generator().manual_seed(1)).bfloat16()
    for _ in range(2):
        _step(layers, tracked_ops, x)
    assert tracked_ops.check() > 50
    assert len(tracked_ops.streams_used()) == 2          # the wgrad GEMMs did go to a side stream


def test_checker_objects_when_a_join_is_missing(tracked_ops, monkeypatch):
    """negative control: without _SideWgrad.join the main stream hands gradients on while the wgrad GEMMs may still run"""
    monkeypatch.setattr(fn, "WGRAD_SIDE_STREAM", True)
    monkeypatch.setattr(fn._SideWgrad, "join", lambda self: None)
    layers = _stack(1)
    x = torch.randn(2, 4, H, generator=torch.Generator().manual_seed(1)).bfloat16()
    with pytest.raises(AssertionError, match="continues without having waited"):
        _step(layers, tracked_ops, x)








# ------------------------------------------------------------------------------------------------ in-switch reducer's side stream
class _Arena:
    """dp.NvlsGradArena's face towards GradAllReducer on the stream model: the reduce kernel is a tracked launch that reads and
    rewrites the gradient on the stream it is given"""

    def __init__(self, tracker):
        self.tracker = tracker
        self.launched = 0

    def current_stream(self):
        return torch.cuda.current_stream()

    def new_stream(self):
        return torch.cuda.Stream()

    def poll(self):
        pass

    def locate(self, g):
        return (0, 0) if g.dim() == 2 else None

    def reduce(self, where, g, stream):
        with torch.cuda.stream(stream):
            self.tracker.op("nvls_reduce", [g], [g])
        self.launched += 1


def test_in_switch_reducer_waits_for_the_wgrad_and_is_joined_by_finish(tracked_ops, monkeypatch):
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(dist, "get_backend", lambda group=None: "gloo")
    monkeypatch.setattr(dist, "all_reduce", lambda t, op=None, group=None, async_op=False: None)     # the packed small gradients
    layers = _stack(2)
    arena = _Arena(tracked_ops)
    red = dp.GradAllReducer(layers.parameters(), prescaled=True, nvls=arena)
    x = torch.randn(2, 4, H, generator=torch.Generator().manual_seed(1)).bfloat16()
    for _ in range(2):
        h = x.clone().requires_grad_(True)
        y = h
        for layer in layers:
            y = layer(y)
        (y.float().square().mean() * red.loss_scale).backward()
        with pytest.raises(AssertionError, match="continues without having waited for nvls_reduce"):
            tracked_ops.assert_joined()                  # before finish(): the reduces are still out on their stream
        red.finish()
        tracked_ops.assert_joined()
        for p in layers.parameters():
            p.grad = None
    assert arena.launched == red.nvls_launches == 2 * 2 * 6       # q, k, v, proj, fc1, fc2 per layer per step
    assert tracked_ops.check() > 100


def test_in_switch_reducer_behind_a_side_stream_wgrad(tracked_ops, monkeypatch):
    """both experiments on: the reduce of a gradient must be ordered after the wgrad GEMM that ran on the OTHER side stream"""
    monkeypatch.setattr(fn, "WGRAD_SIDE_STREAM", True)
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(dist, "get_backend", lambda group=None: "gloo")
    monkeypatch.setattr(dist, "all_reduce", lambda t, op=None, group=None, async_op=False: None)
    layers = _stack(2)
    red = dp.GradAllReducer(layers.parameters(), prescaled=True, nvls=_Arena(tracked_ops))
    x = torch.randn(2, 4, H, generator=torch.Generator().manual_seed(1)).bfloat16()
    h = x.clone().requires_grad_(True)
    y = h
    for layer in layers:
        y = layer(y)
    (y.float().square().mean() * red.loss_scale).backward()
    red.finish()
    tracked_ops.assert_joined()
    assert tracked_ops.check() > 50 and len(tracked_ops.streams_used()) == 3
