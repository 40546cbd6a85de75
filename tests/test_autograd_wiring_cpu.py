"""Host-side autograd wiring of the MXFP8 layer path, checked on CPU with the C-ABI wrappers (llm_fp8_b200.ops) replaced by
exact torch stand-ins (no quantisation: fp32 matmuls, fp32 RMSNorm / SwiGLU, bf16 outputs).  What is under test is the part
that is pure Python: which tensors each autograd Function saves, how many gradients it returns and in which positions, the
residual fusion (`residual=` / `return_input=` / `add_residual=`) and the strided query-slice hand-over.  Numerics of the real
kernels are the GPU tests' business (tests/test_gpu_*.py); nothing here ships.

Reference structure: TELlamaDecoderLayer.forward, te_llama.py:68-82 (h += proj(core(ln_qkv(h))); h += layernorm_mlp(h))."""
import pytest
import torch

import llm_fp8_b200 as te
from llm_fp8_b200 import functional as fn
from llm_fp8_b200 import ops
from llm_fp8_b200.layer import DecoderLinearPath


class _FakeOps:
    """torch stand-ins with the signatures of the ops.* wrappers the MXFP8 Functions call."""
    calls = []

    @staticmethod
    def rows_ok(t):
        return t.dim() == 2 and t.stride(1) == 1 and t.stride(0) >= t.shape[1]

    @staticmethod
    def rmsnorm_rsigma(x, eps):
        _FakeOps.calls.append("rsigma")
        assert x.is_contiguous()
        return torch.rsqrt(x.float().pow(2).mean(-1) + eps)

    @staticmethod
    def mx_quantize(x, rowwise=True, colwise=True, rsigma=None, gamma=None):
        _FakeOps.calls.append("mxq" + ("" if x.is_contiguous() else ":strided"))
        assert x.dim() == 2 and x.stride(1) == 1
        y = x.float()
        if rsigma is not None:
            y = (y * rsigma[:, None]) * gamma.float()
        y = y.bfloat16().contiguous()
        return ops.MXTensor(x.shape[0], x.shape[1], y if rowwise else None, None, y if colwise else None, None)

    @staticmethod
    def mxfp8_gemm(a, sfa, a_mn, b, sfb, b_mn, M, N, K, bias=None, out=None, accumulate=False, out_dtype=torch.bfloat16, max_ctas=0, residual=None):
        _FakeOps.calls.append("gemm" + (":res" if residual is not None else ""))
        A = a.float().t() if a_mn else a.float()
        B = b.float().t() if b_mn else b.float()
        assert A.shape == (M, K) and B.shape == (N, K), (A.shape, B.shape, M, N, K)
        d = A @ B.t()
        if bias is not None:
            d = d + bias.float()
        if residual is not None:
            assert residual.shape == (M, N)
            d = d + residual.float()
        if out is not None:
            assert out.shape == (M, N) and out.stride(1) == 1 and not accumulate
            _FakeOps.calls.append("gemm:out" + (":pitched" if not out.is_contiguous() else ""))
            out.copy_(d)
            return out
        return d.to(out_dtype)

    @staticmethod
    def rmsnorm_bwd(dy, x, gamma, rsigma, add=None):
        _FakeOps.calls.append("rmsbwd" + (":add" if add is not None else ""))
        xh = x.float() * rsigma[:, None]
        tg = dy.float() * gamma.float()
        dx = rsigma[:, None] * (tg - xh * (tg * xh).mean(-1, keepdim=True))
        if add is not None:
            assert add.shape == x.shape and add.is_contiguous()
            dx = dx + add.float()
        return dx.bfloat16(), (dy.float() * xh).sum(0).bfloat16()

    @staticmethod
    def swiglu_bf16(fc1):
        I = fc1.shape[1] // 2
        g, u = fc1[:, :I].float(), fc1[:, I:].float()
        return (torch.nn.functional.silu(g) * u).bfloat16()

    @staticmethod
    def dswiglu_bf16(da, fc1, want_dbias=True):
        I = fc1.shape[1] // 2
        g, u, d = fc1[:, :I].float(), fc1[:, I:].float(), da.float()
        sig = torch.sigmoid(g)
        dg = d * u * sig * (1 + g * (1 - sig))
        du = d * g * sig
        out = torch.cat((dg, du), 1)
        return out.bfloat16(), (out.sum(0).bfloat16() if want_dbias else None)

    @staticmethod
    def mx_quantize_bgrad(x, rowwise=True):
        return _FakeOps.mx_quantize(x, rowwise, True), x.float().sum(0).bfloat16()


@pytest.fixture
def fake_ops(monkeypatch):
    _FakeOps.calls = []
    for name in ("rows_ok", "rmsnorm_rsigma", "mx_quantize", "mxfp8_gemm", "rmsnorm_bwd", "swiglu_bf16", "dswiglu_bf16", "mx_quantize_bgrad"):
        monkeypatch.setattr(ops, name, getattr(_FakeOps, name))
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)      # fp8_autocast's guard; nothing below touches a device
    monkeypatch.setattr(fn, "MX_FUSED_ACT", False)          # the two-kernel activation sequence unless a test turns the fused one on
    monkeypatch.setattr(fn, "MX_EPILOGUE_RESIDUAL", False)
    return _FakeOps


H, I, HEADS, KV, HD = 64, 96, 4, 2, 16


def _layer(fuse):
    torch.manual_seed(3)
    r = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)
    layer = DecoderLinearPath(H, I, HEADS, KV, HD, r, r, params_dtype=torch.bfloat16, fuse_residual=fuse)
    with torch.no_grad():
        for n, p in layer.named_parameters():
            if n.endswith("bias"):
                p.normal_(0, 0.05)
            if n.endswith("layer_norm_weight"):
                p.add_(0.1 * torch.randn_like(p))
    return layer


def _reference(layer, x):
    """Plain torch fp32 statement of the same block (attention core = identity on the query slice)."""
    P = {n: p.detach().float().requires_grad_(True) for n, p in layer.named_parameters()}
    h = x.detach().float().requires_grad_(True)

    def rms(v, g, eps):
        return v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + eps) * g
    wqkv = torch.cat([P["layernorm_qkv.query_weight"], P["layernorm_qkv.key_weight"], P["layernorm_qkv.value_weight"]], 0)
    qkv = rms(h, P["layernorm_qkv.layer_norm_weight"], 1e-5) @ wqkv.t()
    h1 = h + qkv[..., :HEADS * HD] @ P["proj.weight"].t()
    f1 = rms(h1, P["layernorm_mlp.layer_norm_weight"], 1e-5) @ P["layernorm_mlp.fc1_weight"].t() + P["layernorm_mlp.fc1_bias"]
    a = torch.nn.functional.silu(f1[..., :I]) * f1[..., I:]
    out = h1 + a @ P["layernorm_mlp.fc2_weight"].t() + P["layernorm_mlp.fc2_bias"]
    return out, h, P


def _close(got, want, what, tol=3e-2):
    err = (got.float() - want.float()).abs().max().item() / max(want.float().abs().max().item(), 1e-6)
    assert err <= tol, f"{what}: rel err {err:.3e}"


@pytest.mark.parametrize("fuse,epilogue", [(False, False), (True, False), (True, True)])
def test_decoder_linear_path_matches_torch_reference(fake_ops, monkeypatch, fuse, epilogue):
    monkeypatch.setattr(fn, "MX_EPILOGUE_RESIDUAL", epilogue)     # residual in the block-scaled GEMM's epilogue vs an add after it
    layer = _layer(fuse)
    g = torch.Generator().manual_seed(11)
    x = torch.randn(2, 24, H, generator=g).bfloat16().requires_grad_(True)
    dout = torch.randn(2, 24, H, generator=g).bfloat16()
    out = layer(x)
    out.backward(dout)
    ref, hx, P = _reference(layer, x)
    ref.backward(dout.float())
    _close(out, ref, "output")
    _close(x.grad, hx.grad, "dx")
    for n, p in layer.named_parameters():
        assert p.grad is not None, n
        _close(p.grad, P[n].grad, n)
    calls = fake_ops.calls
    assert "mxq:strided" in calls, "the query slice must reach the quantiser as a view (no contiguous copy)"
    if fuse:
        assert calls.count("gemm:res") == (2 if epilogue else 0) and calls.count("rmsbwd:add") == 2, calls
    else:
        assert "gemm:res" not in calls and "rmsbwd:add" not in calls


def test_fused_and_plain_paths_agree(fake_ops):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(3, 8, H, generator=g).bfloat16()
    dout = torch.randn(3, 8, H, generator=g).bfloat16()
    res = []
    for fuse in (False, True):
        layer = _layer(fuse)
        xi = x.clone().requires_grad_(True)
        out = layer(xi)
        out.backward(dout)
        res.append((out.detach(), xi.grad, {n: p.grad for n, p in layer.named_parameters()}))
    _close(res[1][0], res[0][0], "fused vs plain output", tol=2e-2)
    _close(res[1][1], res[0][1], "fused vs plain dx", tol=2e-2)
    for n in res[0][2]:
        _close(res[1][2][n], res[0][2][n], n, tol=2e-2)


def test_linear_residual_and_passthrough_gradients(fake_ops):
    """MXLinearFn alone: residual gradient = dy, pass-through gradient folded into dx, unused pass-through tolerated."""
    torch.manual_seed(0)
    w = (torch.randn(32, H) * 0.1).bfloat16().requires_grad_(True)
    gamma = torch.ones(H).bfloat16().requires_grad_(True)
    x = torch.randn(10, H).bfloat16().requires_grad_(True)
    r = torch.randn(10, 32).bfloat16().requires_grad_(True)
    y = fn.MXLinearFn.apply(x, w, None, None, 0.0, True, r, False)
    dy = torch.randn(10, 32).bfloat16()
    y.backward(dy)
    assert torch.equal(r.grad, dy)
    _close(x.grad, dy.float() @ w.float(), "dx")
    # pass-through used: its gradient arrives in the same backward call
    x2 = x.detach().clone().requires_grad_(True)
    y2, xa = fn.MXLinearFn.apply(x2, w, None, gamma, 1e-5, True, None, True)
    assert xa.data_ptr() == x2.data_ptr()
    extra = torch.randn(10, H).bfloat16()
    (y2.float().sum() + (xa.float() * extra.float()).sum()).backward()
    x3 = x.detach().clone().requires_grad_(True)
    y3 = fn.MXLinearFn.apply(x3, w, None, gamma, 1e-5, True, None, False)
    y3.float().sum().backward()
    _close(x2.grad, x3.grad.float() + extra.float(), "dx with pass-through")
    # pass-through requested but only the alias is consumed
    x4 = x.detach().clone().requires_grad_(True)
    _, xa4 = fn.MXLinearFn.apply(x4, w, None, gamma, 1e-5, True, None, True)
    xa4.float().sum().backward()
    assert torch.equal(x4.grad, torch.ones_like(x4))


def test_weight_copies_are_kept_between_microbatches(fake_ops):
    """is_first_microbatch (TE contract): None casts the weights on every call, True casts and keeps,
    False reuses — same outputs and gradients, fewer quantiser launches; .to() / .cuda() drops the kept copies."""
    torch.manual_seed(1)
    r = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)
    mlp = te.LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16)
    lin = te.Linear(H, 48, bias=False, params_dtype=torch.bfloat16)
    x = torch.randn(6, H).bfloat16().requires_grad_(True)

    def run(first):
        fake_ops.calls.clear()
        x.grad = None
        mlp.zero_grad(); lin.zero_grad()
        with te.fp8_autocast(enabled=True, fp8_recipe=r):
            y = lin(mlp(x, is_first_microbatch=first), is_first_microbatch=first)
        y.float().sum().backward()
        n_q = sum(c.startswith("mxq") for c in fake_ops.calls)
        return y.detach().clone(), x.grad.clone(), mlp.fc1_weight.grad.clone(), lin.weight.grad.clone(), n_q

    base = run(None)
    first = run(True)
    again = run(False)
    assert base[4] == first[4] == again[4] + 3          # fc1, fc2 and the Linear weight are not cast again
    for a, b in zip(base[:4], again[:4]):
        assert torch.equal(a, b)
    assert set(mlp._wcache) == {("w1", "mx"), ("w2", "mx")} and set(lin._wcache) == {("w", "mx")}
    mlp.to(torch.bfloat16)
    assert not mlp._wcache
    assert run(False)[4] == base[4] - 1                 # mlp casts again (and keeps), lin still reuses








@pytest.mark.parametrize("ffn", [128, 96])
def test_fused_activation_branch_is_wired_like_the_two_kernel_sequence(fake_ops, monkeypatch, ffn):
    """functional.MX_FUSED_ACT (LLMFP8_MX_FUSED_ACT=1): SwiGLU / dSwiGLU fused with the block quantiser (ops.mx_swiglu_quantize /
    ops.mx_dswiglu_quantize) must see the same operands and hand on the same tensors as swiglu_bf16 -> mx_quantize and
    dswiglu_bf16 -> mx_quantize; an ffn size that is not a multiple of 128 keeps the two-kernel backward (the C ABI's contract)."""
    used = []

    def mx_swiglu_quantize(fc1, rowwise=True, colwise=True):
        used.append("fwd")
        return _FakeOps.mx_quantize(_FakeOps.swiglu_bf16(fc1), rowwise, colwise)

    def mx_dswiglu_quantize(da, fc1, rowwise=True, colwise=True, want_dbias=True):
        used.append("bwd")
        d, db = _FakeOps.dswiglu_bf16(da, fc1, want_dbias)
        return _FakeOps.mx_quantize(d, rowwise, colwise), db

    monkeypatch.setattr(ops, "mx_swiglu_quantize", mx_swiglu_quantize, raising=False)
    monkeypatch.setattr(ops, "mx_dswiglu_quantize", mx_dswiglu_quantize, raising=False)
    torch.manual_seed(2)
    r = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)
    mlp = te.LayerNormMLP(H, ffn, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16)
    with torch.no_grad():
        mlp.fc1_bias.normal_(0, 0.05)
    x = torch.randn(6, H, generator=torch.Generator().manual_seed(3)).bfloat16()

    def run(flag):
        monkeypatch.setattr(fn, "MX_FUSED_ACT", flag)
        mlp.zero_grad()
        h = x.clone().requires_grad_(True)
        with te.fp8_autocast(enabled=True, fp8_recipe=r):
            y = mlp(h)
        y.float().square().mean().backward()
        return [y.detach().clone(), h.grad.clone()] + [p.grad.clone() for p in mlp.parameters()]

    base = run(False)
    assert used == []
    fused = run(True)
    assert used == (["fwd", "bwd"] if ffn % 128 == 0 else ["fwd"])
    for a, b in zip(base, fused):
        assert torch.equal(a, b)


def test_query_slice_gradient_is_handed_over_without_a_copy(fake_ops):
    """layer._QuerySlice (the attention stand-in) + proj Linear: the dgrad GEMM writes dq straight into columns [0, q) of a zero-padded
    [M, q + 2kv] buffer (functional._padded_dx) and the slice's backward returns that buffer — no torch.cat; values equal the reference."""
    layer = _layer(True)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 24, H, generator=g).bfloat16().requires_grad_(True)
    out = layer(x)
    fake_ops.calls.clear()
    seen = {}
    real_cat = torch.cat

    def spy_cat(ts, dim=0, **kw):
        import traceback
        if any(f.name == "backward" and f.filename.endswith("layer.py") for f in traceback.extract_stack(limit=4)):
            seen["cat"] = True          # called from _QuerySlice.backward (the stand-in kernels concatenate too)
        return real_cat(ts, dim=dim, **kw)
    torch.cat, saved = spy_cat, torch.cat
    try:
        out.float().square().mean().backward()
    finally:
        torch.cat = saved
    assert "gemm:out:pitched" in fake_ops.calls          # proj dgrad wrote into the wide buffer
    assert "cat" not in seen
    ref_out, h, P = _reference(layer, x.detach())
    ref_out.float().square().mean().backward()
    _close(x.grad, h.grad, "dx through the slice hand-over")
