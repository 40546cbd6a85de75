"""Fused AdamW + MXFP8 requantise (csrc/adamw_mx.cu, llm_fp8_b200/optim.py; the reference steps torch.optim.AdamW
on bf16 parameters: train_fp8.py:200-204).  The update against an fp32 torch restatement of torch's AdamW formula (one bf16 rounding per tensor),
the emitted MXFP8 copies BIT-IDENTICAL to the block quantiser run on the updated weight, and the optimizer end to end: the forward after a
step finds its weight casts done (fewer launches, same bytes), and forgets them the moment anything else touches the weights."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_adamw(w, g, m, v, step, lr, b1, b2, eps, wd):
    w, g, m, v = (t.float() for t in (w, g, m, v))
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    denom = v.sqrt() / (1 - b2 ** step) ** 0.5 + eps
    w = w * (1 - lr * wd) - (lr / (1 - b1 ** step)) * m / denom
    return w.bfloat16(), m.bfloat16(), v.bfloat16()


def _close_bf16(got, want, what):
    """equal up to one bf16 ulp on isolated elements (fma contraction / op order), never more; where the two terms of an update cancel
    (b1 m = -(1 - b1) g) the fused multiply-add leaves a residue of fp32 rounding size where separate roundings leave 0: absolute floor"""
    d = (got.float() - want.float()).abs()
    ulp = want.float().abs() * 2 ** -7 + want.float().abs().max() * 2 ** -20
    bad = (d > ulp).nonzero()
    assert bad.numel() == 0, f"{what}: {bad.shape[0]} elements off by more than a bf16 ulp, first at {bad[0].tolist()}: got {got[tuple(bad[0])].item()!r} want {want[tuple(bad[0])].item()!r}"
    assert (d > 0).float().mean().item() < 0.02, f"{what}: {100 * (d > 0).float().mean().item():.2f} % of elements differ"


@pytest.mark.parametrize("R,C", [(128, 128), (256, 384), (992, 1056), (4096, 1024), (96, 160)])
def test_adamw_mx_step_matches_fp32_reference_and_block_quantiser(R, C):
    from llm_fp8_b200 import _lib, ops
    lib = _lib.load()
    g0 = torch.Generator(device="cuda").manual_seed(R + C)
    w = (torch.randn(R, C, device="cuda", generator=g0) * 0.02).bfloat16()
    m = torch.zeros_like(w)
    v = torch.zeros_like(w)
    lr, b1, b2, eps, wd = 3e-3, 0.9, 0.95, 1e-8, 0.1
    st = torch.cuda.current_stream().cuda_stream
    for step in range(1, 4):
        g = (torch.randn(R, C, device="cuda", generator=g0) * 1e-3 * torch.exp(torch.randn(R, 1, device="cuda", generator=g0))).bfloat16()
        want_w, want_m, want_v = _ref_adamw(w, g, m, v, step, lr, b1, b2, eps, wd)
        out = ops._mx_out(R, C, True, True, w.device)
        rc = lib.llmfp8_adamw_mx_step(w.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), R, C, lr, b1, b2, eps, wd, 1 - b1 ** step, 1 - b2 ** step,
                                      out.row.data_ptr(), out.sf_row.data_ptr(), out.col.data_ptr(), out.sf_col.data_ptr(), st)
        _lib.check(rc, "adamw_mx_step")
        torch.cuda.synchronize()
        _close_bf16(w, want_w, f"w step {step}")
        _close_bf16(m, want_m, f"exp_avg step {step}")
        _close_bf16(v, want_v, f"exp_avg_sq step {step}")
        fresh = ops.mx_quantize(w, True, True)                      # the copies are the block quantiser's bytes of the UPDATED weight
        assert torch.equal(out.row, fresh.row) and torch.equal(out.sf_row, fresh.sf_row)
        assert torch.equal(out.col, fresh.col) and torch.equal(out.sf_col, fresh.sf_col)
        # only one copy requested
        w2, m2, v2 = w.clone(), m.clone(), v.clone()
        only = ops._mx_out(R, C, True, False, w.device)
        rc = lib.llmfp8_adamw_mx_step(w2.data_ptr(), g.data_ptr(), m2.data_ptr(), v2.data_ptr(), R, C, lr, b1, b2, eps, wd, 1 - b1 ** (step + 1),
                                      1 - b2 ** (step + 1), only.row.data_ptr(), only.sf_row.data_ptr(), None, None, st)
        _lib.check(rc, "adamw_mx_step")
        torch.cuda.synchronize()
        assert torch.equal(only.row, ops.mx_quantize(w2, True, False).row)


def test_fused_adamw_end_to_end_and_forward_skips_its_weight_casts():
    import llm_fp8_b200 as te
    from llm_fp8_b200 import _lib
    from llm_fp8_b200.layer import LinearPathStack
    lib = _lib.load()
    te.reset_fp8_state()
    model = LinearPathStack("llama-3.2-1b", "mxfp8", 2, device="cuda", seed=5)
    twin = copy.deepcopy(model)
    opt = te.FusedAdamW(model, lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
    ref = torch.optim.AdamW(twin.parameters(), lr=1e-3, betas=(0.9, 0.95), weight_decay=0.1)
    g = torch.Generator().manual_seed(1)
    for step in range(3):
        x = torch.randn(2, 128, model.H, generator=g).bfloat16().cuda()
        opt.zero_grad()
        model(x).float().pow(2).mean().backward()
        # the SAME gradients for both optimizers: Adam normalises them, so two models one bf16 ulp apart drift chaotically (measured: 0.7 % of the
        # elements move in opposite directions by the third step) — what is under test is the update rule, not that feedback
        for p, q in zip(model.parameters(), twin.parameters()):
            q.grad = p.grad.detach().clone()
        opt.step()
        ref.step()
    for (n, p), (_, q) in zip(model.named_parameters(), twin.named_parameters()):
        d = (p.float() - q.float()).abs()
        spacing = q.float().abs().max().item() * 2 ** -7                # one bf16 step at the largest weight
        # torch's bf16 AdamW rounds after every elementwise op, ours once per tensor: a step or two apart on some elements, equal on average
        assert d.max().item() <= 4 * spacing + 1e-8, f"{n}: max diff {d.max().item():.3e} vs bf16 spacing {spacing:.3e}"
        assert d.mean().item() <= 0.25 * spacing + 1e-9, f"{n}: mean diff {d.mean().item():.3e} vs bf16 spacing {spacing:.3e}"
    # the forward after a step: no weight casts (4 GEMM weights per layer), identical bytes
    x = torch.randn(2, 128, model.H, generator=g).bfloat16().cuda()
    n0 = lib.llmfp8_launch_count()
    with torch.no_grad():
        y_cached = model(x)
    launches_cached = lib.llmfp8_launch_count() - n0
    for mod in model.modules():
        if hasattr(mod, "_wcache"):
            mod._wcache.clear()
    n0 = lib.llmfp8_launch_count()
    with torch.no_grad():
        y_recast = model(x)
    launches_recast = lib.llmfp8_launch_count() - n0
    assert torch.equal(y_cached, y_recast)
    assert launches_recast - launches_cached == 4 * model.n_layers, (launches_cached, launches_recast)
    # anything else touching a weight invalidates its copies (version counter)
    opt.zero_grad()
    model(x).float().pow(2).mean().backward()
    opt.step()
    with torch.no_grad():
        model.layers[0].proj.weight.mul_(0.5)
        y_after = model(x)
        for mod in model.modules():
            if hasattr(mod, "_wcache"):
                mod._wcache.clear()
        assert torch.equal(y_after, model(x))
