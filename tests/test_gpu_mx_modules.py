"""Module-level parity of the MXFP8 recipe (te_llama_mxfp8.py:28-29, 86-94): Linear / LayerNormLinear / LayerNormMLP forward + backward
against an oracle emulation of the SAME quantise -> GEMM sequence — which copy (1x32 rowwise or 32x1 colwise blocks) of which tensor
feeds fprop, dgrad and wgrad — at the north star's 1e-2 (relative to max |ref|), bias gradients included.  The emulation is
oracle/fp8_oracle.py's block quantiser + dequantise -> fp32 matmul; RMSNorm / SwiGLU values come from oracle/oracle_c.c."""
import pytest
import torch

from oracle import c_oracle as co
from oracle import fp8_oracle as fo

pytestmark = pytest.mark.gpu
TOL = 1e-2


def close(got, want, tol=TOL, what=""):
    got, want = got.float().cpu(), want.float().cpu()
    err = (got - want).abs().max().item() / (want.abs().max().item() + 1e-30)
    assert err <= tol, f"{what}: rel err {err:.3e}"
    return err


def rnd(shape, seed, std=1.0, heavy=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g) * std
    if heavy:                                       # heavy-tailed rows so neighbouring blocks get different scales
        x = x * torch.exp(torch.randn(shape[:-1] + (1,), generator=g))
    return x.bfloat16()


def mx_row(x):
    """dequantised 1x32-block copy (blocks along the last dim) of x [R, C]"""
    q, e = fo.mx_quantize_rowwise(x.float())
    return fo.mx_dequantize_rowwise(q, e)


def mx_col(x):
    """dequantised 32x1-block copy (blocks along dim 0)"""
    q, e = fo.mx_quantize_colwise(x.float())
    return fo.mx_dequantize_colwise(q, e)


def mm(a, b):
    return (a.double() @ b.double()).float()


def norm_bwd(dln_bf16, x2, gamma, rs, add=None):
    """fp32 RMSNorm backward on the bf16 gradient the dgrad GEMM hands over (K5's arithmetic, reduction order free)."""
    xh = x2.float() * rs[:, None]
    tg = dln_bf16.float() * gamma.float()
    dx = rs[:, None] * (tg - xh * (tg * xh).mean(-1, keepdim=True))
    if add is not None:
        dx = dx + add.float()
    return dx, (dln_bf16.float() * xh).sum(0)


def _recipe():
    import llm_fp8_b200 as te
    return te, te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)


@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("residual", [False, True])
def test_mx_linear_matches_oracle_sequence(bias, residual):
    te, recipe = _recipe()
    torch.manual_seed(0)
    K, N, B, S = 512, 768, 4, 64
    if residual:
        N = K
    lin = te.Linear(K, N, bias=bias, params_dtype=torch.bfloat16, device="cuda")
    if bias:
        with torch.no_grad():
            lin.bias.copy_(rnd((N,), 99, 0.1).cuda())
    w = lin.weight.detach().cpu()
    for step in range(2):
        x = rnd((B, S, K), 10 + step, 1.0 + step, heavy=True).cuda().requires_grad_(True)
        res = rnd((B, S, N), 40 + step).cuda().requires_grad_(True) if residual else None
        dy = rnd((B, S, N), 20 + step, 1e-2, heavy=True).cuda()
        lin.zero_grad()
        with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
            y = lin(x, residual=res) if residual else lin(x)
        y.backward(dy)
        x2, dy2 = x.detach().cpu().reshape(-1, K), dy.cpu().reshape(-1, N)
        want = mm(mx_row(x2), mx_row(w).t())
        if bias:
            want = want + lin.bias.detach().cpu().float()
        if residual:
            want = want + res.detach().cpu().float().reshape(-1, N)
        close(y.reshape(-1, N), want, what=f"y step {step}")
        close(x.grad.reshape(-1, K), mm(mx_row(dy2), mx_col(w)), what="dx")            # dY blocks along N, W blocks along N (colwise)
        close(lin.weight.grad, mm(mx_col(dy2).t(), mx_col(x2)), what="dw")              # both blocks along the tokens
        if bias:
            close(lin.bias.grad, dy2.float().sum(0), what="dbias")
        if residual:
            assert torch.equal(res.grad, dy)


def test_mx_layernorm_linear_matches_oracle_sequence():
    te, recipe = _recipe()
    torch.manual_seed(1)
    H, KV = 512, 128
    N = H + 2 * KV
    m = te.LayerNormLinear(H, N, eps=1e-5, bias=False, normalization="RMSNorm", parameters_split={"query": H, "key": KV, "value": KV},
                           params_dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        m.layer_norm_weight.copy_((1 + 0.1 * torch.randn(H)).bfloat16())
    gamma = m.layer_norm_weight.detach().cpu()
    x = rnd((2, 96, H), 3, heavy=True).cuda().requires_grad_(True)
    dy = rnd((2, 96, N), 4, 1e-2, heavy=True).cuda()
    with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
        y, x_alias = m(x, return_input=True)
    dres = rnd((2, 96, H), 5, 1e-2).cuda()
    torch.autograd.backward([y, x_alias], [dy, dres])
    w = torch.cat([m.query_weight, m.key_weight, m.value_weight]).detach().cpu()
    x2, dy2 = x.detach().cpu().reshape(-1, H), dy.cpu().reshape(-1, N)
    yln, rs = co.rmsnorm_y(x2, gamma, 1e-5)                           # fp32 normalised activations, the kernel's reduction order
    close(y.reshape(-1, N), mm(mx_row(yln), mx_row(w).t()), what="y")
    dw = mm(mx_col(dy2).t(), mx_col(yln))
    close(m.query_weight.grad, dw[:H], what="dq")
    close(m.key_weight.grad, dw[H:H + KV], what="dk")
    close(m.value_weight.grad, dw[H + KV:], what="dv")
    dln = mm(mx_row(dy2), mx_col(w)).bfloat16()
    dx, dgamma = norm_bwd(dln, x2, gamma, rs, add=dres.cpu().reshape(-1, H))
    close(x.grad.reshape(-1, H), dx, what="dx (norm backward + residual-branch gradient)")
    close(m.layer_norm_weight.grad, dgamma, what="dgamma")


@pytest.mark.parametrize("add_residual", [False, True])
def test_mx_layernorm_mlp_matches_oracle_sequence(add_residual):
    """K1m -> G1m(fc1)+b -> SwiGLU -> K2m -> G1m(fc2)+b (+x); backward K2m(dY) -> G2m(fc2) -> dSwiGLU(+dbias) -> K2m -> G3m(fc2),
    G2m(fc1), G3m(fc1) -> K5 — every operand the block-scaled copy the emulation says (MXFP8 variant)."""
    te, recipe = _recipe()
    torch.manual_seed(2)
    H, I = 512, 1024
    m = te.LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        m.layer_norm_weight.copy_((1 + 0.1 * torch.randn(H)).bfloat16())
        m.fc1_bias.copy_(rnd((2 * I,), 5, 0.05))
        m.fc2_bias.copy_(rnd((H,), 6, 0.05))
    x = rnd((2, 128, H), 7, heavy=True).cuda().requires_grad_(True)
    dy = rnd((2, 128, H), 8, 1e-1, heavy=True).cuda()
    with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
        y = m(x, add_residual=add_residual)
    y.backward(dy)
    p = {k: v.detach().cpu() for k, v in m.named_parameters()}
    x2, dy2 = x.detach().cpu().reshape(-1, H), dy.cpu().reshape(-1, H)
    yln, rs = co.rmsnorm_y(x2, p["layer_norm_weight"], 1e-5)
    fc1 = (mm(mx_row(yln), mx_row(p["fc1_weight"]).t()) + p["fc1_bias"].float()).bfloat16()
    a = co.swiglu_f32(fc1).bfloat16()
    want = mm(mx_row(a), mx_row(p["fc2_weight"]).t()) + p["fc2_bias"].float()
    if add_residual:
        want = want + x2.float()
    close(y.reshape(-1, H), want, what="y")
    da = mm(mx_row(dy2), mx_col(p["fc2_weight"])).bfloat16()                       # fc2 dgrad
    dfc1 = co.dswiglu_f32(da, fc1).bfloat16()
    close(m.fc2_weight.grad, mm(mx_col(dy2).t(), mx_col(a)), what="dW2")
    close(m.fc2_bias.grad, dy2.float().sum(0), what="db2")
    close(m.fc1_bias.grad, dfc1.float().sum(0), what="db1")
    close(m.fc1_weight.grad, mm(mx_col(dfc1).t(), mx_col(yln)), what="dW1")
    dln = mm(mx_row(dfc1), mx_col(p["fc1_weight"])).bfloat16()                     # fc1 dgrad
    dx, dgamma = norm_bwd(dln, x2, p["layer_norm_weight"], rs, add=dy2 if add_residual else None)
    close(x.grad.reshape(-1, H), dx, what="dx")
    close(m.layer_norm_weight.grad, dgamma, what="dgamma")


def test_mx_decoder_path_matches_oracle_layer():
    """One DecoderLinearPath block (the structure of te_llama_mxfp8.py:86-94 with the attention core = identity on the query slice),
    fused-residual form, against the plain composition of the three module emulations above — the residual wiring end to end."""
    from llm_fp8_b200.layer import DecoderLinearPath
    te, recipe = _recipe()
    torch.manual_seed(3)
    H, I, heads, kv, hd = 512, 1024, 8, 2, 64
    blk = DecoderLinearPath(H, I, heads, kv, hd, recipe, recipe, device="cuda")
    with torch.no_grad():
        blk.layernorm_qkv.layer_norm_weight.copy_((1 + 0.1 * torch.randn(H)).bfloat16())
        blk.layernorm_mlp.layer_norm_weight.copy_((1 + 0.1 * torch.randn(H)).bfloat16())
    x = rnd((2, 64, H), 11).cuda().requires_grad_(True)
    y = blk(x)
    p = {k: v.detach().cpu() for k, v in blk.named_parameters()}
    x2 = x.detach().cpu().reshape(-1, H)
    wqkv = torch.cat([p["layernorm_qkv.query_weight"], p["layernorm_qkv.key_weight"], p["layernorm_qkv.value_weight"]])
    yln, _ = co.rmsnorm_y(x2, p["layernorm_qkv.layer_norm_weight"], 1e-5)
    qkv = mm(mx_row(yln), mx_row(wqkv).t()).bfloat16()
    h1 = (mm(mx_row(qkv[:, :heads * hd]), mx_row(p["proj.weight"]).t()) + x2.float()).bfloat16()
    yln2, _ = co.rmsnorm_y(h1, p["layernorm_mlp.layer_norm_weight"], 1e-5)
    fc1 = (mm(mx_row(yln2), mx_row(p["layernorm_mlp.fc1_weight"]).t()) + p["layernorm_mlp.fc1_bias"].float()).bfloat16()
    a = co.swiglu_f32(fc1).bfloat16()
    want = mm(mx_row(a), mx_row(p["layernorm_mlp.fc2_weight"]).t()) + p["layernorm_mlp.fc2_bias"].float() + h1.float()
    close(y.reshape(-1, H), want, what="layer output")
    y.float().pow(2).mean().backward()
    assert all(torch.isfinite(v.grad.float()).all() for v in blk.parameters())


def test_first_backward_of_a_process_can_be_a_tma_kernel():
    """autograd's backward thread has no CUDA context bound until a runtime call runs on it, and cuTensorMapEncodeTiled (a driver entry
    point) does not bind one: the very first backward of a fresh process going straight into the MX quantiser must still work."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("import torch, llm_fp8_b200 as te\n"
            "lin = te.Linear(256, 256, bias=False, params_dtype=torch.bfloat16, device='cuda')\n"
            "x = torch.randn(64, 256, device='cuda').bfloat16().requires_grad_(True)\n"
            "with te.fp8_autocast(enabled=True, fp8_recipe=te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)):\n"
            "    y = lin(x)\n"
            "y.backward(torch.ones_like(y))\n"
            "torch.cuda.synchronize(); assert torch.isfinite(x.grad.float()).all(); print('ok')\n")
    out = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]
