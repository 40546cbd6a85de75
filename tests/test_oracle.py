"""Pins the CPU oracle: reference fixtures (thesis known-answer encodings), and agreement of the two
independent restatements (torch float8 casts vs integer bit-level C) on every bf16 value."""
import numpy as np
import pytest
import torch

from oracle import c_oracle as co
from oracle import fp8_oracle as fo

E4M3, E5M2 = fo.E4M3, fo.E5M2


def all_bf16():
    return torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)


def same_bytes(a, b):
    """byte equality, with any-NaN == any-NaN of the same sign handled explicitly (both must be NaN)."""
    a = a.view(torch.uint8).reshape(-1)
    b = b.view(torch.uint8).reshape(-1)
    return torch.equal(a, b)


# --- reference fixtures: thesis/chapters/c2/c2_quantization.tex:100,102,103 -------------------
def test_thesis_known_answers():
    # E4M3 0.2979 -> 0.3125 (exp field 5, mantissa 2)
    q = fo.sat_cast(torch.tensor([0.2979]), E4M3)
    assert q.float().item() == 0.3125
    b = q.view(torch.uint8).item()
    assert (b >> 3) & 0xF == 5 and b & 7 == 2
    # E5M2 0.2979 -> 0.3125 (exp 13, mantissa 1)
    q = fo.sat_cast(torch.tensor([0.2979]), E5M2)
    b = q.view(torch.uint8).item()
    assert q.float().item() == 0.3125 and (b >> 2) & 0x1F == 13 and b & 3 == 1
    # E5M2 500 -> 512 (exp 24, mantissa 0)
    q = fo.sat_cast(torch.tensor([500.0]), E5M2)
    b = q.view(torch.uint8).item()
    assert q.float().item() == 512.0 and (b >> 2) & 0x1F == 24 and b & 3 == 0
    # line 101 (E4M3 500 -> 240) is wrong for e4m3fn; the saturating rule gives 448
    assert fo.sat_cast(torch.tensor([500.0]), E4M3).float().item() == 448.0
    for v, dt in [(0.2979, E4M3), (0.2979, E5M2), (500.0, E5M2), (500.0, E4M3)]:
        assert co.cast_f32(torch.tensor([v]), dt).item() == fo.sat_cast(torch.tensor([v]), dt).view(torch.uint8).item()


# --- two independent restatements agree everywhere --------------------------------------------
@pytest.mark.parametrize("dtype", [E4M3, E5M2])
@pytest.mark.parametrize("scale", [1.0, 0.37, 448.0 / 3.1, 57344.0 / 0.011, 2.0 ** -20, 3.0e38])
def test_c_vs_torch_all_bf16(dtype, scale):
    x = all_bf16()
    qt, at = fo.quantize(x, scale, dtype)
    qc, ac = co.quantize_bf16(x, scale, dtype)
    assert same_bytes(qt, qc)
    assert at.item() == ac          # NaN ignored, inf kept


@pytest.mark.parametrize("dtype", [E4M3, E5M2])
def test_decode_roundtrip_all_bytes(dtype):
    b = torch.arange(256, dtype=torch.int32).to(torch.uint8)
    dec_c = co.decode(b, dtype)
    dec_t = b.view(dtype).float()
    nan = torch.isnan(dec_t)
    assert torch.equal(torch.isnan(dec_c), nan)
    assert torch.equal(dec_c[~nan], dec_t[~nan])
    fin = torch.isfinite(dec_t)
    assert torch.equal(co.cast_f32(dec_t[fin], dtype), b[fin])        # idempotent on representables
    assert same_bytes(fo.sat_cast(dec_t[fin], dtype), b[fin])


@pytest.mark.parametrize("dtype", [E4M3, E5M2])
def test_ties_and_edges(dtype):
    # midpoints between every pair of adjacent representable magnitudes: RNE goes to the even byte
    b = torch.arange(0, 0x7E if dtype == E4M3 else 0x7B, dtype=torch.int32).to(torch.uint8)
    lo, hi = b.view(dtype).float(), (b + 1).view(dtype).float()
    mid = (lo.double() + hi.double()) / 2
    assert torch.equal(mid.float().double(), mid)                      # midpoints exact in fp32
    want = torch.where(b % 2 == 0, b, b + 1)
    for sign in (1.0, -1.0):
        got_c = co.cast_f32(sign * mid.float(), dtype)
        got_t = fo.sat_cast(sign * mid.float(), dtype).view(torch.uint8)
        w = want | (0x80 if sign < 0 else 0)
        assert torch.equal(got_c, w) and torch.equal(got_t, w)
    fmax = fo.FMAX[dtype]
    edge = torch.tensor([fmax, fmax * 1.0001, 1e30, float("inf"), -float("inf"), -fmax * 2, 0.0, -0.0])
    assert same_bytes(fo.sat_cast(edge, dtype), co.cast_f32(edge, dtype))
    assert fo.sat_cast(torch.tensor([float("inf")]), dtype).float().item() == fmax
    assert torch.isnan(fo.sat_cast(torch.tensor([float("nan")]), dtype).float()).all()
    assert co.cast_f32(torch.tensor([float("nan")]), dtype).item() == 0x7F


def test_random_fp32_casts():
    g = torch.Generator().manual_seed(7)
    x = torch.randn(1 << 18, generator=g) * torch.exp(4 * torch.randn(1 << 18, generator=g))
    for dt in (E4M3, E5M2):
        assert same_bytes(fo.sat_cast(x, dt), co.cast_f32(x, dt))


# --- E8M0 ----------------------------------------------------------------------------------------
def test_e8m0_roundup_rule():
    v = torch.tensor([0.0, 1.0, 1.0000001, 1.5, 2.0, 3.0, 2.0 ** -127, 2.0 ** -126, 1.5 * 2.0 ** -127,
                      2.0 ** 127, 1.5 * 2.0 ** 127, float("inf"), float("nan"), 2.0 ** -149, 0.75 * 2.0 ** -126])
    e = fo.e8m0_roundup(v)
    want = [0, 127, 128, 128, 128, 129, 0, 1, 1, 254, 254, 255, 255, 0, 1]
    assert e.tolist() == want
    for vi, wi in zip(v.tolist(), want):
        assert co.lib().orc_e8m0_roundup(vi) == wi
    # torch's own e8m0 cast rounds to nearest and must NOT be what we use
    assert torch.tensor([3.0]).to(torch.float8_e8m0fnu).float().item() == 4.0
    assert fo.e8m0_roundup(torch.tensor([2.5])).item() == 129       # up, not nearest (nearest would be 128)


def test_e8m0_scales_cover_block():
    g = torch.Generator().manual_seed(3)
    x = torch.randn(256, 256, generator=g) * torch.exp(3 * torch.randn(256, 1, generator=g))
    q, e = fo.mx_quantize_rowwise(x)
    # round-up => scaled block amax <= 448: nothing saturates
    inv = fo.e8m0_inv_scale(e)
    scaled = x.reshape(256, 8, 32) * inv.unsqueeze(-1)
    assert scaled.abs().max() <= 448.0
    assert (scaled.abs().amax(-1) > 224.0 - 1e-3).all()              # and the scale is the tightest power of two
    d = fo.mx_dequantize_rowwise(q, e)
    assert ((d - x).abs() <= x.reshape(256, 8, 32).abs().amax(-1, keepdim=True).expand(-1, -1, 32).reshape(256, 256) * 2 ** -3).all()


def test_mx_c_vs_torch_and_swizzle():
    g = torch.Generator().manual_seed(11)
    x = (torch.randn(256, 384, generator=g) * torch.exp(2 * torch.randn(256, 384, generator=g))).bfloat16().float()
    x[5, 32:64] = 0.0                      # all-zero block
    x[130, 7] = 3.0e38                     # huge
    x[:, 100] *= 1e-30                     # tiny column
    out = co.mx_quantize_f32(x)
    q, e = fo.mx_quantize_rowwise(x)
    assert same_bytes(q, out["q_row"])
    assert torch.equal(fo.swizzle_sf(e), out["sf_row"])
    assert e[5, 1].item() == 0 and (q[5, 32:64].view(torch.uint8) == 0).all()
    qc, ec = fo.mx_quantize_colwise(x)
    assert same_bytes(qc, out["q_col"])
    assert torch.equal(fo.swizzle_sf(ec), out["sf_col"])
    # swizzle is a bijection on the padded grid and matches the closed-form offset
    assert torch.equal(fo.unswizzle_sf(fo.swizzle_sf(e), *e.shape), e)
    kbp = (e.shape[1] + 3) // 4 * 4
    flat = fo.swizzle_sf(e)
    for r, kb in [(0, 0), (31, 3), (32, 0), (127, 11), (128, 4), (255, 11), (200, 7)]:
        assert flat[co.lib().orc_sf_offset(r, kb, kbp)].item() == e[r, kb].item()


# --- delayed scaling -------------------------------------------------------------------------------
def test_delayed_scaling_state_machine():
    T, L = 5, 16
    fmax = torch.tensor([448.0, 448.0, 57344.0, 448.0, 57344.0])
    st = fo.DelayedScalingState(T, L, 0, fmax)
    hc, sc, sic = st.amax_history.clone(), st.scale.clone(), st.scale_inv.clone()
    g = torch.Generator().manual_seed(5)
    seen = []
    for step in range(40):                                  # > 2 wraps of the 16-deep history
        am = torch.rand(T, generator=g) * 10 ** float(torch.randint(-3, 4, (1,), generator=g))
        if step == 7:
            am[1] = float("inf")                            # non-finite amax: scale must not change
        if step in (20, 21):
            am[3] = 0.0
        prev_scale = st.scale.clone()
        st.amax_history[0] = am
        hc[0] = am
        st.update()
        co.amax_scale_update(hc, sc, sic, fmax, 0)
        assert torch.equal(hc, st.amax_history) and torch.equal(sc, st.scale) and torch.equal(sic, st.scale_inv)
        seen.append(am)
        window = torch.stack(seen[-L:]).clone()
        if not torch.isfinite(window[:, 1]).all():
            assert st.scale[1] == prev_scale[1]
        fin = torch.isfinite(window).all(0)
        want = fmax / window.max(0).values
        assert torch.equal(st.scale[fin], want[fin])
        assert (st.amax_history[0] == 0).all()
    # all-zero history keeps the scale
    st2 = fo.DelayedScalingState(2, 4)
    st2.scale[:] = 3.0
    st2.update()
    assert (st2.scale == 3.0).all()


# --- pinned RMSNorm / SwiGLU restatements stay within float tolerance of the free-order torch refs --
def test_rmsnorm_c_close_to_torch():
    g = torch.Generator().manual_seed(1)
    for H in (2048, 3072, 4096, 5120, 136):
        x = torch.randn(9, H, generator=g).bfloat16()
        gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
        y, rs = co.rmsnorm_y(x, gamma, 1e-5)
        yr, rsr = fo.rmsnorm_ref(x, gamma, 1e-5)
        assert torch.allclose(rs, rsr, rtol=2e-6, atol=0)
        assert torch.allclose(y, yr, rtol=4e-6, atol=1e-7)
        q, rs2, am, _ = co.rmsnorm_quant(x, gamma, 1e-5, 17.0, E4M3)
        assert torch.equal(rs2, rs) and am == y.abs().max().item()
        assert same_bytes(q, fo.sat_cast(y * 17.0, E4M3))


def test_swiglu_c_close_to_torch():
    g = torch.Generator().manual_seed(2)
    fc1 = (torch.randn(64, 512, generator=g) * 4).bfloat16()
    fc1[0, :8] = torch.tensor([-100.0, 100.0, -88.0, 88.0, 0.0, -0.0, 30.0, -30.0]).bfloat16()
    a = co.swiglu_f32(fc1)
    ar = fo.swiglu_ref(fc1)
    assert torch.allclose(a, ar, rtol=1e-5, atol=1e-30)
    # the pinned exp(-g) (orc_exp_neg) against the double-precision exponential: full clamp range and the range that matters
    for lim, tol in ((80.0, 1.5e-6), (16.0, 5e-7)):
        xs = torch.linspace(-lim, lim, 20001)
        ex = torch.tensor([co.lib().orc_exp_neg(v) for v in xs.tolist()], dtype=torch.float64)
        rel = ((ex - torch.exp(-xs.double())) / torch.exp(-xs.double())).abs().max().item()
        assert rel < tol, (lim, rel)
    # pinned sigmoid over EVERY bf16 gate value in the clamp range against torch
    allg = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16).float()
    allg = allg[torch.isfinite(allg) & (allg.abs() <= 80)]
    sg = torch.tensor([co.lib().orc_sigmoid(v) for v in allg.tolist()], dtype=torch.float64)
    assert torch.allclose(sg, torch.sigmoid(allg.double()), rtol=2e-6, atol=0)
    q, am = co.swiglu_quant(fc1, 3.0, E4M3)
    assert same_bytes(q, fo.sat_cast(a * 3.0, E4M3)) and am == a.abs().max().item()
    # backward against autograd of the torch reference
    da = torch.randn(64, 256, generator=g).bfloat16()
    f = fc1.float().requires_grad_(True)
    (fo.swiglu_ref(f) * da.float()).sum().backward()
    d = co.dswiglu_f32(da, fc1)
    assert torch.allclose(d, f.grad, rtol=2e-5, atol=1e-6)
