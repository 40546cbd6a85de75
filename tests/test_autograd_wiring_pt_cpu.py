"""Host-side autograd wiring of the PER-TENSOR (DelayedScaling) layer path on CPU, the counterpart of test_autograd_wiring_cpu.py for
functional.LinearFn / LayerNormMLPFn: the C-ABI wrappers are replaced by exact torch stand-ins (no quantisation: the "fp8" copies
are the bf16 values, every scale is 1), K4 by the oracle's update, CUDA's stream objects by the happens-before model of
tests/_streams.py.  Under test: which tensors the Functions save, gradient positions, the slot bookkeeping they pass on, the
wgrad-output homes (functional.wgrad_out) and the fused residual path against the reference-shaped one.
Reference structure: TELlamaDecoderLayer.forward with the two DelayedScaling recipes of te_llama.py:39-40, 68-82."""
import pytest
import torch

import llm_fp8_b200 as te
from _streams import fake_cuda, tensors_in
from test_autograd_wiring_cpu import H, HD, HEADS, I, KV, _close, _reference

from llm_fp8_b200 import fp8, ops
from llm_fp8_b200 import functional as fn
from llm_fp8_b200.layer import DecoderLinearPath
from oracle import fp8_oracle as fo


def _q(t, fmt, pool, slot):
    assert 0 <= slot < pool.used, "slot outside what the module claimed"
    return ops.QTensor(t.bfloat16().contiguous(), fmt, pool.scale_inv, slot)


class _FakePt:
    calls = []

    @staticmethod
    def quantize(x, fmt, pool, slot):
        _FakePt.calls.append(("quantize", slot - 0))
        assert x.dim() == 2 and x.stride(1) == 1
        return _q(x, fmt, pool, slot)

    @staticmethod
    def quantize_bgrad(x, fmt, pool, slot):
        return _q(x, fmt, pool, slot), x.float().sum(0).bfloat16()

    @staticmethod
    def rmsnorm_quant(x, gamma, eps, fmt, pool, slot):
        rs = torch.rsqrt(x.float().pow(2).mean(-1) + eps)
        return _q((x.float() * rs[:, None]) * gamma.float(), fmt, pool, slot), rs

    @staticmethod
    def swiglu_quant(fc1, fmt, pool, slot):
        n = fc1.shape[1] // 2
        return _q(torch.nn.functional.silu(fc1[:, :n].float()) * fc1[:, n:].float(), fmt, pool, slot)

    @staticmethod
    def dswiglu_quant(da, fc1, fmt, pool, slot, want_dbias=True):
        n = fc1.shape[1] // 2
        g, u, d = fc1[:, :n].float(), fc1[:, n:].float(), da.float()
        sig = torch.sigmoid(g)
        out = torch.cat((d * u * sig * (1 + g * (1 - sig)), d * g * sig), 1)
        return _q(out, fmt, pool, slot), (out.sum(0).bfloat16() if want_dbias else None)

    @staticmethod
    def fp8_gemm(a, a_mn, b, b_mn, M, N, K, bias=None, out=None, accumulate=False, out_dtype=torch.bfloat16, max_ctas=0, residual=None):
        _FakePt.calls.append("gemm" + (":res" if residual is not None else "") +
                             ("" if out is None else (":home" if out.is_contiguous() else ":pitched")))      # pitched: dx into a zero-padded wide buffer
        A = a.data.float().t() if a_mn else a.data.float()
        B = b.data.float().t() if b_mn else b.data.float()
        assert A.shape == (M, K) and B.shape == (N, K), (A.shape, B.shape, M, N, K)
        d = A @ B.t()
        if bias is not None:
            d = d + bias.float()
        if residual is not None:
            assert residual.shape == (M, N)
            d = d + residual.float()
        if out is not None:
            assert out.shape == (M, N) and not accumulate
            out.copy_(d)
            return out
        return d.to(out_dtype)

    @staticmethod
    def rmsnorm_bwd(dy, x, gamma, rsigma, add=None):
        xh = x.float() * rsigma[:, None]
        tg = dy.float() * gamma.float()
        dx = rsigma[:, None] * (tg - xh * (tg * xh).mean(-1, keepdim=True))
        if add is not None:
            assert add.shape == x.shape and add.is_contiguous()
            dx = dx + add.float()
        return dx.bfloat16(), (dy.float() * xh).sum(0).bfloat16()

    @staticmethod
    def rows_ok(t):
        return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]


NAMES = ("quantize", "quantize_bgrad", "rmsnorm_quant", "swiglu_quant", "dswiglu_quant", "fp8_gemm", "rmsnorm_bwd")


@pytest.fixture
def pt_ops(monkeypatch):
    """-> (stream tracker, K4 call list); every stand-in is also a tracked launch on the current (model) stream"""
    tracker = fake_cuda(monkeypatch)
    _FakePt.calls = []
    monkeypatch.setattr(ops, "rows_ok", _FakePt.rows_ok)
    for name in NAMES:
        real = getattr(_FakePt, name)

        def wrapped(*a, _real=real, _name=name, **kw):
            out = _real(*a, **kw)
            # a QTensor carries a reference to its pool's live scale_inv tensor: that is the ("scale", pool, slot) resource below, not a
            # buffer this launch produces or consumes as a whole
            live = {p.scale_inv.untyped_storage().data_ptr() for p in fp8.all_pools()}
            mine = lambda ts: [t for t in ts if t.untyped_storage().data_ptr() not in live]      # noqa: E731
            reads = mine(list(tensors_in(a)) + [t for k, v in kw.items() if k != "out" for t in tensors_in(v)])
            writes = mine(list(tensors_in(out)))
            for i, arg in enumerate(a):                  # a quantiser reads the scale of ITS slot and max-accumulates into that slot's amax
                if isinstance(arg, fp8.ScalePool):
                    reads.append(("scale", id(arg), a[i + 1]))
                    writes.append(("amax", id(arg), a[i + 1]))
            tracker.op(_name, reads, writes)
            return out
        monkeypatch.setattr(ops, name, wrapped)
    k4 = []

    def update(hist, scale, scale_inv, fmax, T, margin):
        k4.append(T)
        fo.amax_scale_update(hist, scale, scale_inv, fmax, margin)
        pool = next(p for p in fp8.all_pools() if p.amax_history is hist)
        slots = range(pool.used)
        tracker.op("amax_scale_update", [("amax", id(pool), i) for i in slots],
                   [("amax", id(pool), i) for i in slots] + [("scale", id(pool), i) for i in slots])

    monkeypatch.setattr(fp8, "amax_scale_update", update)
    fp8.reset_fp8_state()
    yield tracker, k4
    fp8.reset_fp8_state()


def _recipes():
    return (te.DelayedScaling(fp8_format=te.Format.HYBRID, amax_history_len=16, amax_compute_algo="max"),      # te_llama.py:39
            te.DelayedScaling(fp8_format=te.Format.E4M3, amax_history_len=16, amax_compute_algo="max"))        # te_llama.py:40


def _layer(fuse, seed=3):
    torch.manual_seed(seed)
    ar, mr = _recipes()
    layer = DecoderLinearPath(H, I, HEADS, KV, HD, ar, mr, params_dtype=torch.bfloat16, fuse_residual=fuse)
    with torch.no_grad():
        for n, p in layer.named_parameters():
            if n.endswith("bias"):
                p.normal_(0, 0.05)
            if n.endswith("layer_norm_weight"):
                p.add_(0.1 * torch.randn_like(p))
    return layer


@pytest.mark.parametrize("fuse", [False, True])
def test_per_tensor_decoder_path_matches_torch_reference(pt_ops, fuse):
    tracker, k4 = pt_ops
    layer = _layer(fuse)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 24, H, generator=g).bfloat16().requires_grad_(True)
    dout = torch.randn(2, 24, H, generator=g).bfloat16()
    for step in range(2):                                # the second step runs the pools' lazy flushes
        for p in layer.parameters():
            p.grad = None
        x.grad = None
        out = layer(x)
        out.backward(dout)
    ref, hx, P = _reference(layer, x)
    ref.backward(dout.float())
    _close(out, ref, "output")
    _close(x.grad, hx.grad, "dx")
    for n, p in layer.named_parameters():
        assert p.grad is not None, n
        _close(p.grad, P[n].grad, n)
    calls = _FakePt.calls
    assert calls.count("gemm:res") == (4 if fuse else 0)              # proj + fc2 epilogues, two steps
    assert len(k4) == 4                                               # per step boundary: two recipes x two directions
    assert tracker.check() > 40


def test_per_tensor_wgrads_land_in_their_homes(pt_ops):
    """functional.wgrad_out under LinearFn / LayerNormMLPFn: the joined qkv buffer, proj, fc1 and fc2 each get their registered home as
    the wgrad GEMM's output, autograd adopts it (row slices for the split qkv Parameters), values as without homes"""
    from llm_fp8_b200 import dp
    layer = _layer(True)
    x = torch.randn(2, 8, H, generator=torch.Generator().manual_seed(4)).bfloat16()

    def run():
        for p in layer.parameters():
            p.grad = None
        h = x.clone().requires_grad_(True)
        layer(h).float().square().mean().backward()
        return {n: p.grad.clone() for n, p in layer.named_parameters()}

    plain = run()
    bufs = dp.gemm_weight_buffers(layer)
    arena = torch.zeros(sum(w.numel() for w, _ in bufs), dtype=torch.bfloat16)
    off = 0
    try:
        for w, ps in bufs:
            fn._WGRAD_OUT[w.data_ptr()] = (arena[off:off + w.numel()], list(ps))
            off += w.numel()
        _FakePt.calls.clear()
        homed = run()
        assert _FakePt.calls.count("gemm:home") == 4
        lo, hi = arena.data_ptr(), arena.data_ptr() + arena.numel() * 2
        for n, p in layer.named_parameters():
            assert torch.equal(homed[n], plain[n]), n
            assert (lo <= p.grad.data_ptr() < hi) == (p.dim() == 2), n
    finally:
        fn._WGRAD_OUT.clear()

