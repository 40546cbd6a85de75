"""Module-level parity: Linear / LayerNormLinear / LayerNormMLP forward + backward against an oracle emulation of the same
quantise -> GEMM sequence, over several steps so the delayed-scaling state (history, scales) is exercised."""
import os

import pytest
import torch

from oracle import c_oracle as co
from oracle import fp8_oracle as fo

pytestmark = pytest.mark.gpu
E4M3, E5M2 = fo.E4M3, fo.E5M2
TOL = 1e-2     # BASELINE.json: GEMM outputs within 1e-2 (relative to max |ref|) on bf16


def close(got, want, tol=TOL, what=""):
    got, want = got.float().cpu(), want.float().cpu()
    err = (got - want).abs().max().item() / (want.abs().max().item() + 1e-30)
    assert err <= tol, f"{what}: rel err {err:.3e}"


def rnd(shape, seed, std=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * std).bfloat16()


class DelayedEmu:
    """Oracle-side delayed-scaling slots with the lazy 'update before next use' schedule."""

    def __init__(self, n, fmax, L=16):
        self.st = fo.DelayedScalingState(n, L, 0, torch.tensor(fmax))
        self.dirty = False

    def begin(self):
        if self.dirty:
            self.st.update()
            self.dirty = False

    def q(self, x, slot, dtype):
        q, am = fo.quantize(x, self.st.scale[slot], dtype)
        self.st.record(slot, am)
        return q.float() * self.st.scale_inv[slot]


@pytest.mark.parametrize("fmt_name", ["HYBRID", "E4M3"])
@pytest.mark.parametrize("bias", [False, True])
def test_linear_steps(fmt_name, bias):
    import llm_fp8_b200 as te
    te.reset_fp8_state()
    fmt = getattr(te.Format, fmt_name)
    bdt = E5M2 if fmt_name == "HYBRID" else E4M3
    recipe = te.DelayedScaling(fp8_format=fmt, amax_history_len=16, amax_compute_algo="max")
    torch.manual_seed(0)
    K, N, B, S = 512, 768, 4, 64
    lin = te.Linear(K, N, bias=bias, params_dtype=torch.bfloat16, device="cuda")
    if bias:
        with torch.no_grad():
            lin.bias.copy_(rnd((N,), 99, 0.1).cuda())
    w = lin.weight.detach().cpu()
    b = lin.bias.detach().cpu().float() if bias else None
    fw, bw = DelayedEmu(3, [448.0] * 3), DelayedEmu(2, [fmt.value.max_bwd] * 2)
    for step in range(4):
        x = rnd((B, S, K), 10 + step, 1.0 + step).cuda().requires_grad_(True)
        dy = rnd((B, S, N), 20 + step, 1e-2 * (1 + step)).cuda()
        with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
            y = lin(x)
        y.backward(dy)
        # oracle emulation
        fw.begin()
        xd, wd = fw.q(x.detach().cpu().reshape(-1, K), 0, E4M3), fw.q(w, 1, E4M3)
        fw.dirty = True
        want_y = xd @ wd.t() + (b if bias else 0)
        bw.begin()
        dyd = bw.q(dy.cpu().reshape(-1, N), 0, bdt)
        bw.dirty = True
        close(y.reshape(-1, N), want_y, what=f"y step {step}")
        close(x.grad.reshape(-1, K), dyd @ wd, what=f"dx step {step}")
        close(lin.weight.grad, dyd.t() @ xd, what=f"dw step {step}")
        if bias:
            close(lin.bias.grad, dy.float().cpu().reshape(-1, N).sum(0), what="dbias")
        lin.weight.grad = None
        if bias:
            lin.bias.grad = None
    # device state == oracle state after a final flush
    te.flush_all()
    fw.begin()
    bw.begin()
    s = lin._meta.get(recipe, torch.device("cuda", torch.cuda.current_device()))
    assert torch.equal(s["pf"].scale[s["f0"]:s["f0"] + 2].cpu(), fw.st.scale[:2])
    assert torch.equal(s["pf"].amax_history[:, s["f0"]:s["f0"] + 2].cpu(), fw.st.amax_history[:, :2])
    assert torch.equal(s["pb"].scale[s["b0"]:s["b0"] + 1].cpu(), bw.st.scale[:1])


def test_layernorm_linear_split_params():
    import llm_fp8_b200 as te
    te.reset_fp8_state()
    recipe = te.DelayedScaling(fp8_format=te.Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
    torch.manual_seed(1)
    H, KV = 512, 128
    m = te.LayerNormLinear(H, H + 2 * KV, eps=1e-5, bias=False, normalization="RMSNorm", parameters_split={"query": H, "key": KV, "value": KV},
                           params_dtype=torch.bfloat16, device="cuda")
    with torch.no_grad():
        m.layer_norm_weight.copy_((1 + 0.1 * torch.randn(H)).bfloat16())
    assert set(dict(m.named_parameters())) == {"layer_norm_weight", "query_weight", "key_weight", "value_weight"}
    x = rnd((2, 96, H), 3).cuda().requires_grad_(True)
    dy = rnd((2, 96, H + 2 * KV), 4, 1e-2).cuda()
    with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
        y = m(x)
    y.backward(dy)
    w = torch.cat([m.query_weight, m.key_weight, m.value_weight]).detach().cpu()
    x2 = x.detach().cpu().reshape(-1, H)
    q, rs, am, yln = co.rmsnorm_quant(x2, m.layer_norm_weight.detach().cpu(), 1e-5, 1.0, E4M3, want_y=True)
    xd = q.view(E4M3).float()
    wd = fo.sat_cast(w.float(), E4M3).float()
    close(y.reshape(-1, H + 2 * KV), xd @ wd.t(), what="y")
    dyd = fo.sat_cast(dy.cpu().float().reshape(-1, H + 2 * KV), E5M2).float()
    dw = dyd.t() @ xd
    close(m.query_weight.grad, dw[:H], what="dq")
    close(m.value_weight.grad, dw[H + KV:], what="dv")
    # dx through the norm: autograd of the float reference with the fp8 dgrad as upstream gradient
    xf = x2.float().requires_grad_(True)
    gf = m.layer_norm_weight.detach().cpu().float().requires_grad_(True)
    yr, _ = fo.rmsnorm_ref(xf, gf, 1e-5)
    yr.backward((dyd @ wd).bfloat16().float())
    close(x.grad.reshape(-1, H), xf.grad, tol=2e-2, what="dx")
    close(m.layer_norm_weight.grad, gf.grad, tol=2e-2, what="dgamma")


@pytest.mark.parametrize("recipe_name", ["E4M3", "HYBRID", "MXFP8"])
def test_layernorm_mlp_vs_float_reference(recipe_name):
    """End-to-end tolerance check of LayerNormMLP fwd+bwd against the unquantised fp32 computation (loose: FP8 noise)."""
    import llm_fp8_b200 as te
    te.reset_fp8_state()
    if recipe_name == "MXFP8":
        recipe = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)
    else:
        recipe = te.DelayedScaling(fp8_format=getattr(te.Format, recipe_name), amax_history_len=16, amax_compute_algo="max")
    torch.manual_seed(2)
    H, I = 512, 1024
    m = te.LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16, device="cuda")
    assert m.fc1_bias is not None and m.eps == 1e-5            # reference relies on these defaults (te_llama.py:58-63)
    with torch.no_grad():
        m.fc1_bias.copy_(rnd((2 * I,), 5, 0.05))
        m.fc2_bias.copy_(rnd((H,), 6, 0.05))
    x = rnd((2, 128, H), 7).cuda().requires_grad_(True)
    dy = rnd((2, 128, H), 8, 1e-1).cuda()
    for step in range(3):      # step 0 runs with scale 1 (delayed scaling warm-up); later steps use real scales
        x.grad = None
        m.zero_grad()
        with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
            y = m(x)
        y.backward(dy)
    xf = x.detach().cpu().float().requires_grad_(True)
    ps = {k: v.detach().cpu().float().requires_grad_(True) for k, v in m.named_parameters()}
    yln, _ = fo.rmsnorm_ref(xf, ps["layer_norm_weight"], 1e-5)
    f1 = yln @ ps["fc1_weight"].t() + ps["fc1_bias"]
    out = fo.swiglu_ref(f1) @ ps["fc2_weight"].t() + ps["fc2_bias"]
    out.backward(dy.cpu().float())
    tol = 1e-1       # two chained FP8 GEMMs — an FP8-noise band against the UNQUANTISED computation, not the parity bar
    close(y, out, tol, "y")
    close(x.grad, xf.grad, tol, "dx")
    for k, v in m.named_parameters():
        close(v.grad, ps[k].grad, tol, k)


def test_modules_refuse_outside_autocast():
    import llm_fp8_b200 as te
    lin = te.Linear(64, 64, params_dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="fp8_autocast"):
        lin(torch.zeros(2, 64, dtype=torch.bfloat16, device="cuda"))


def test_decoder_layer_swap_runs():
    """The reference's decoder-layer structure (te_llama.py:41-82) on the shim: tensor in, tensor out, grads flow, all recipes."""
    import transformer_engine as te
    from transformer_engine.common.recipe import DelayedScaling, Format, MXFP8BlockScaling
    from transformer_engine.pytorch.attention import RotaryPositionEmbedding
    import llm_fp8_b200
    llm_fp8_b200.reset_fp8_state()
    H, heads, kv, I = 512, 8, 2, 1024
    torch.manual_seed(3)

    class Layer(torch.nn.Module):
        def __init__(self, attn_recipe, mlp_recipe):
            super().__init__()
            te.pytorch.TransformerLayer
            self.self_attention = te.pytorch.MultiheadAttention(hidden_size=H, num_attention_heads=heads, bias=False, layernorm_epsilon=1e-5,
                                                                attention_dropout=0.0, fuse_qkv_params=False, normalization="RMSNorm",
                                                                num_gqa_groups=kv, qkv_format="bshd", input_layernorm=True)
            self.layernorm_mlp = te.pytorch.LayerNormMLP(hidden_size=H, ffn_hidden_size=I, normalization="RMSNorm", activation="swiglu")
            self.rope = RotaryPositionEmbedding(H // heads)(max_seq_len=256).cuda()
            self.ar, self.mr = attn_recipe, mlp_recipe

        def forward(self, h, attention_mask=None, **kw):
            with te.pytorch.fp8_autocast(enabled=True, fp8_recipe=self.ar):
                a = self.self_attention(h, attention_mask=attention_mask, rotary_pos_emb=self.rope)
            h = h + a
            with te.pytorch.fp8_autocast(enabled=True, fp8_recipe=self.mr):
                f = self.layernorm_mlp(h)
            return h + f

    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        for ar, mr in [(DelayedScaling(fp8_format=Format.HYBRID, amax_history_len=16, amax_compute_algo="max"),
                        DelayedScaling(fp8_format=Format.E4M3, amax_history_len=16, amax_compute_algo="max")),
                       (MXFP8BlockScaling(fp8_format=Format.E4M3),) * 2]:
            layer = Layer(ar, mr).cuda()
            x = rnd((2, 128, H), 9).cuda().requires_grad_(True)
            for _ in range(3):
                y = layer(x)
                assert isinstance(y, torch.Tensor) and y.shape == x.shape and y.dtype == torch.bfloat16
                y.float().pow(2).mean().backward()
            assert torch.isfinite(y.float()).all() and torch.isfinite(x.grad.float()).all()
            for n, p in layer.named_parameters():
                assert p.grad is not None and torch.isfinite(p.grad.float()).all(), n
    finally:
        torch.set_default_dtype(old)


@pytest.mark.parametrize("scenario", ["mxfp8", "default", "hybrid"])
def test_decoder_linear_path_fused_residual_matches_plain(scenario):
    """DecoderLinearPath with the residual adds fused (proj / fc2 GEMM epilogues, RMSNorm backward) against the plain
    `h = h + f(h)` form with identical weights.  The fused form rounds each sum once instead of twice, so h differs by a bf16 ulp in
    about half its elements, and every such difference can flip an FP8 rounding downstream: the two runs agree within an FP8-noise
    band (measured 2-6 % of max|ref| on the gradients), not bit for bit.  Exactness of each fused kernel is pinned separately
    (test_gpu_gemm: fused residual == plain + R after one rounding; test_gpu_quantize: rmsnorm_bwd + dres), the autograd wiring in
    tests/test_autograd_wiring_cpu.py.  The query slice reaches the proj quantiser as a strided view in both."""
    import llm_fp8_b200 as te
    from llm_fp8_b200.layer import DecoderLinearPath, recipes_for
    H, I, heads, kv, hd = 512, 1024, 8, 2, 64
    ar, mr = recipes_for(scenario)
    x0 = rnd((2, 128, H), 21).cuda()
    dy = rnd((2, 128, H), 22, 1e-1).cuda()
    res = []
    for fuse in (False, True):
        te.reset_fp8_state()
        torch.manual_seed(4)
        layer = DecoderLinearPath(H, I, heads, kv, hd, ar, mr, device="cuda", fuse_residual=fuse)
        with torch.no_grad():
            layer.layernorm_mlp.fc1_bias.copy_(rnd((2 * I,), 5, 0.05))
            layer.layernorm_mlp.fc2_bias.copy_(rnd((H,), 6, 0.05))
        for step in range(3):
            x = x0.clone().requires_grad_(True)
            layer.zero_grad()
            y = layer(x)
            y.backward(dy)
        res.append((y.detach(), x.grad, {n: p.grad for n, p in layer.named_parameters()}))
    tol = 1e-1
    close(res[1][0], res[0][0], 2e-2, "y")
    close(res[1][1], res[0][1], tol, "dx")
    for n in res[0][2]:
        close(res[1][2][n], res[0][2][n], tol, n)


@pytest.mark.parametrize("recipe_name", ["MXFP8", "HYBRID"])
def test_is_first_microbatch_keeps_weight_copies(recipe_name):
    """Gradient accumulation (TE's is_first_microbatch): micro-batch 2 reuses the fp8 weight copies of micro-batch 1.
    MXFP8 has no scaling state, so the reused copies give bit-identical results to casting again; with delayed scaling the kept copy
    carries the scale it was cast with (the pool's slot has moved on), and results stay inside the FP8 band.  Either way the
    quantiser launches for the weights disappear."""
    import llm_fp8_b200 as te
    from llm_fp8_b200 import _lib
    lib = _lib.load()
    te.reset_fp8_state()
    recipe = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3) if recipe_name == "MXFP8" else \
        te.DelayedScaling(fp8_format=te.Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
    torch.manual_seed(2)
    H, I = 512, 1024
    m = te.LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16, device="cuda")
    lin = te.Linear(H, 256, bias=False, params_dtype=torch.bfloat16, device="cuda")
    xs = [rnd((2, 128, H), 30 + i).cuda().requires_grad_(True) for i in range(2)]

    def step(flags):
        m.zero_grad(); lin.zero_grad()
        outs, n0 = [], lib.llmfp8_launch_count()
        for x, first in zip(xs, flags):
            x.grad = None
            with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
                y = lin(m(x, is_first_microbatch=first), is_first_microbatch=first)
            y.float().pow(2).mean().backward()
            outs.append((y.detach().clone(), x.grad.clone()))
        return outs, m.fc1_weight.grad.clone(), lin.weight.grad.clone(), lib.llmfp8_launch_count() - n0

    for _ in range(2):                       # settle the delayed-scaling history
        step((None, None))
    base = step((None, None))
    kept = step((True, False))
    assert kept[3] == base[3] - 3, (kept[3], base[3])            # fc1, fc2 and the Linear weight are cast once, not twice
    if recipe_name == "MXFP8":
        for (ya, ga), (yb, gb) in zip(base[0], kept[0]):
            assert torch.equal(ya, yb) and torch.equal(ga, gb)
        assert torch.equal(base[1], kept[1]) and torch.equal(base[2], kept[2])
    else:
        for (ya, ga), (yb, gb) in zip(base[0], kept[0]):
            close(yb, ya, 1e-1, "y")
            close(gb, ga, 1e-1, "dx")
        close(kept[1], base[1], 1e-1, "dW1")
        close(kept[2], base[2], 1e-1, "dW")


def test_wgrad_side_stream_matches_main_stream():
    """functional.WGRAD_SIDE_STREAM: the wgrad GEMMs of every linear on a side stream (joined before backward returns).  Stream
    placement changes no arithmetic: MXFP8 has no scaling state, so weight / input gradients must be bit-identical over several
    steps (a missing dependency would show up as a stale operand)."""
    import llm_fp8_b200 as te
    from llm_fp8_b200 import functional as fn
    te.reset_fp8_state()
    recipe = te.MXFP8BlockScaling(fp8_format=te.Format.E4M3)
    torch.manual_seed(3)
    H, I = 1024, 4096
    m = te.LayerNormMLP(H, I, normalization="RMSNorm", activation="swiglu", params_dtype=torch.bfloat16, device="cuda")
    lin = te.LayerNormLinear(H, 1536, bias=False, normalization="RMSNorm", params_dtype=torch.bfloat16, device="cuda")
    xs = [rnd((4, 512, H), 60 + i).cuda().requires_grad_(True) for i in range(3)]

    def run(flag):
        fn.WGRAD_SIDE_STREAM = flag
        out = []
        for x in xs:
            m.zero_grad(); lin.zero_grad(); x.grad = None
            with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
                y = lin(m(x))
            y.float().pow(2).mean().backward()
            out.append([t.clone() for t in (y.detach(), x.grad, m.fc1_weight.grad, m.fc2_weight.grad, lin.weight.grad, m.layer_norm_weight.grad)])
        torch.cuda.synchronize()
        return out

    saved = fn.WGRAD_SIDE_STREAM
    try:
        base, side = run(False), run(True)
    finally:
        fn.WGRAD_SIDE_STREAM = saved
    for a, b in zip(base, side):
        for ta, tb in zip(a, b):
            assert torch.equal(ta, tb)




def test_two_forwards_per_step_follow_the_oracle_state_machine():
    """Gradient accumulation WITHOUT is_first_microbatch (the paper's accum-4 runs, paper/conference_101719.tex:222; the reference never
    passes the argument): two forwards, then the two backwards in reverse order, per optimizer step.  TE updates the scales at every
    outermost autocast exit / after every module backward; here the update is deferred to the next use of the slots — which is the
    second forward / second backward of the same step, so the sequence of scales must equal fo.DelayedScalingState driven once per
    forward and once per backward.  Each backward de-scales the fp8 copies saved by ITS forward with the scale_inv of that forward."""
    import llm_fp8_b200 as te
    te.reset_fp8_state()
    recipe = te.DelayedScaling(fp8_format=te.Format.HYBRID, amax_history_len=16, amax_compute_algo="max")
    torch.manual_seed(0)
    K, N, T = 512, 768, 256
    lin = te.Linear(K, N, bias=False, params_dtype=torch.bfloat16, device="cuda")
    w = lin.weight.detach().cpu()
    fw, bw = DelayedEmu(3, [448.0] * 3), DelayedEmu(2, [57344.0] * 2)
    for step in range(4):
        xs = [rnd((T, K), 100 + 2 * step + i, (1.0 + step) * (1 + 3 * i)).cuda().requires_grad_(True) for i in range(2)]
        dys = [rnd((T, N), 200 + 2 * step + i, 1e-2 * (1 + step) * (1 + 2 * i)).cuda() for i in range(2)]
        ys, emu = [], []
        for x in xs:
            with te.fp8_autocast(enabled=True, fp8_recipe=recipe):
                ys.append(lin(x))
            fw.begin()
            xd, wd = fw.q(x.detach().cpu(), 0, E4M3), fw.q(w, 1, E4M3)
            fw.dirty = True
            emu.append((xd, wd))
        lin.weight.grad = None
        want_dw = 0
        for i in (1, 0):
            ys[i].backward(dys[i])
            bw.begin()
            dyd = bw.q(dys[i].cpu(), 0, E5M2)
            bw.dirty = True
            xd, wd = emu[i]
            close(ys[i], xd @ wd.t(), what=f"y[{i}] step {step}")
            close(xs[i].grad, dyd @ wd, what=f"dx[{i}] step {step}")
            want_dw = want_dw + dyd.t() @ xd
        close(lin.weight.grad, want_dw, what=f"accumulated dw step {step}")
    te.flush_all()
    fw.begin()
    bw.begin()
    s = lin._meta.get(recipe, torch.device("cuda", torch.cuda.current_device()))
    assert torch.equal(s["pf"].scale[s["f0"]:s["f0"] + 2].cpu(), fw.st.scale[:2])
    assert torch.equal(s["pf"].amax_history[:, s["f0"]:s["f0"] + 2].cpu(), fw.st.amax_history[:, :2])
    assert torch.equal(s["pb"].scale[s["b0"]:s["b0"] + 1].cpu(), bw.st.scale[:1])
    assert torch.equal(s["pb"].amax_history[:, s["b0"]:s["b0"] + 1].cpu(), bw.st.amax_history[:, :1])
