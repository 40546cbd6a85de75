"""Bit-exact parity of the quantise kernels (through the C ABI) with the CPU oracle."""
import pytest
import torch

from oracle import c_oracle as co
from oracle import fp8_oracle as fo

pytestmark = pytest.mark.gpu
E4M3, E5M2 = fo.E4M3, fo.E5M2
FMT = {E4M3: 0, E5M2: 1}


def _ops():
    from llm_fp8_b200 import ops
    return ops


def dev(x):
    return x.cuda()


def gen(shape, seed, kind="normal"):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g)
    if kind == "heavy":
        x = x * torch.exp(torch.randn(shape, generator=g))
    if kind == "outlier":
        mask = torch.rand(shape, generator=g) < 1e-3
        x = torch.where(mask, x * 100, x)
    return x.bfloat16()


def assert_fp8_equal(got_u8, want_u8, dtype):
    """bytes equal; NaN positions only need to be NaN on both sides."""
    got_u8, want_u8 = got_u8.cpu().reshape(-1), want_u8.cpu().reshape(-1)
    nan_w = torch.isnan(want_u8.view(dtype).float())
    nan_g = torch.isnan(got_u8.view(dtype).float())
    assert torch.equal(nan_w, nan_g)
    bad = (got_u8 != want_u8) & ~nan_w
    assert not bad.any(), f"{int(bad.sum())} of {bad.numel()} bytes differ; first at {int(bad.nonzero()[0])}"


@pytest.mark.parametrize("dtype", [E4M3, E5M2])
def test_cast_all_bf16_values(dtype):
    ops = _ops()
    x = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    for scale in (1.0, 0.37, 448.0 / 3.1, 2.0 ** -20, 3.0e38):
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q = ops.quantize_raw(dev(x), FMT[dtype], s, amax)
        qo, ao = fo.quantize(x, scale, dtype)
        assert_fp8_equal(q, qo.view(torch.uint8), dtype)
        assert amax.item() == ao.item()


@pytest.mark.parametrize("kind", ["normal", "heavy", "outlier"])
@pytest.mark.parametrize("shape", [(6144, 4096), (48, 16), (1, 16), (333, 80)])
def test_cast_amax(shape, kind):
    ops = _ops()
    x = gen(shape, 1, kind)
    am0 = fo.amax_of(x).item()
    scale = 448.0 / am0
    for dtype in (E4M3, E5M2):
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q = ops.quantize_raw(dev(x), FMT[dtype], s, amax)
        qo, ao = fo.quantize(x, scale, dtype)
        assert_fp8_equal(q, qo.view(torch.uint8), dtype)
        assert amax.item() == ao.item()
    # amax accumulates (atomic max), never lowers
    amax = torch.full((1,), 1e30, dtype=torch.float32).cuda()
    ops.quantize_raw(dev(x), 0, s, amax)
    assert amax.item() == pytest.approx(1e30)


def test_cast_empty_and_contract():
    ops = _ops()
    x = torch.empty(0, 16, dtype=torch.bfloat16).cuda()
    assert ops.quantize_raw(x, 0, None, None).numel() == 0
    with pytest.raises(RuntimeError, match="multiple of 16"):
        ops.quantize_raw(torch.zeros(3, 8, dtype=torch.bfloat16).cuda(), 0, None, None)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.quantize_raw(torch.zeros(16, dtype=torch.bfloat16), 0, None, None)


@pytest.mark.parametrize("shape", [(8192, 4096), (100, 264), (7, 8), (513, 2048)])
def test_cast_bgrad(shape):
    ops = _ops()
    x = gen(shape, 2) * 1e-3
    x = x.bfloat16()
    scale = 57344.0 / fo.amax_of(x).item()
    s = torch.tensor([scale], dtype=torch.float32).cuda()
    amax = torch.zeros(1, dtype=torch.float32).cuda()
    q, bgrad = ops.quantize_bgrad_raw(dev(x), 1, s, amax)
    qo, ao = fo.quantize(x, scale, E5M2)
    assert_fp8_equal(q, qo.view(torch.uint8), E5M2)
    assert amax.item() == ao.item()
    want = co.colsum_bf16(x)
    tol = 1e-2 * want.abs().max().item() + 1e-6
    assert (bgrad.float().cpu() - want).abs().max().item() <= tol


@pytest.mark.parametrize("H", [2048, 3072, 4096, 5120, 136, 8192])
def test_rmsnorm_quant(H):
    ops = _ops()
    M = 6144 if H == 4096 else 67
    g = torch.Generator().manual_seed(H)
    x = gen((M, H), 3, "heavy" if H == 4096 else "normal")
    gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
    eps = 1e-5
    _, _, am, _ = co.rmsnorm_quant(x, gamma, eps, 1.0, E4M3)
    scale = 448.0 / am
    for dtype in (E4M3, E5M2):
        qo, rso, amo, _ = co.rmsnorm_quant(x, gamma, eps, scale, dtype)
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q, rs = ops.rmsnorm_quant_raw(dev(x), dev(gamma), eps, FMT[dtype], s, amax)
        assert torch.equal(rs.cpu(), rso), "rsigma must match the pinned reduction order bit for bit"
        assert_fp8_equal(q, qo, dtype)
        assert amax.item() == amo
    rs2 = ops.rmsnorm_rsigma(dev(x), eps)
    assert torch.equal(rs2.cpu(), rso)


def test_rmsnorm_bwd():
    ops = _ops()
    for M, H in ((300, 2048), (64, 4096), (33, 5120)):
        x, dy = gen((M, H), 4), gen((M, H), 5)
        g = torch.Generator().manual_seed(9)
        gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
        rs = ops.rmsnorm_rsigma(dev(x), 1e-5)
        dx, dgamma = ops.rmsnorm_bwd(dev(dy), dev(x), dev(gamma), rs)
        xf = x.float().requires_grad_(True)
        gf = gamma.float().requires_grad_(True)
        y, _ = fo.rmsnorm_ref(xf, gf, 1e-5)
        y.backward(dy.float())
        assert torch.allclose(dx.float().cpu(), xf.grad, rtol=2e-2, atol=2e-2 * xf.grad.abs().max().item())
        assert torch.allclose(dgamma.float().cpu(), gf.grad, rtol=2e-2, atol=2e-2 * gf.grad.abs().max().item())


@pytest.mark.parametrize("shape", [(6144, 2 * 1024), (50, 2 * 264), (3, 16)])
def test_swiglu_quant(shape):
    ops = _ops()
    fc1 = (gen(shape, 6) * 3).bfloat16()
    a = co.swiglu_f32(fc1)
    scale = 448.0 / a.abs().max().item()
    for dtype in (E4M3, E5M2):
        qo, amo = co.swiglu_quant(fc1, scale, dtype)
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q = ops.swiglu_quant_raw(dev(fc1), FMT[dtype], s, amax)
        assert_fp8_equal(q, qo, dtype)
        assert amax.item() == amo
    ab = ops.swiglu_bf16(dev(fc1))
    assert torch.equal(ab.cpu(), a.bfloat16())


def test_swiglu_every_bf16_gate_value():
    """The pinned sigmoid depends on the gate alone, and the gate is bf16: all 65536 gate bit patterns (x 4 up values) are checked
    bit for bit — exhaustive proof that MUFU.RCP + one Newton step equals the oracle's correctly rounded 1 / (1 + E)."""
    ops = _ops()
    allg = torch.arange(0, 65536, dtype=torch.int32).to(torch.int16).view(torch.bfloat16)
    ups = torch.tensor([1.0, -0.7421875, 3.140625, 1e-3]).bfloat16()
    gate = allg.repeat(4).reshape(4 * 64, 1024)                     # M = 256 rows of I = 1024
    up = ups.repeat_interleave(65536).reshape(4 * 64, 1024)
    fc1 = torch.cat([gate, up], dim=1).contiguous()
    a = co.swiglu_f32(fc1)
    ab = ops.swiglu_bf16(dev(fc1)).cpu()
    want = a.bfloat16()
    nan = torch.isnan(want.float())
    assert torch.equal(torch.isnan(ab.float()), nan)
    assert torch.equal(ab.view(torch.int16)[~nan], want.view(torch.int16)[~nan])
    da = torch.full((256, 1024), 0.5).bfloat16()
    d = co.dswiglu_f32(da, fc1)
    db, _ = ops.dswiglu_bf16(dev(da), dev(fc1))
    wantd = d.bfloat16()
    nand = torch.isnan(wantd.float())
    assert torch.equal(torch.isnan(db.cpu().float()), nand)
    assert torch.equal(db.cpu().view(torch.int16)[~nand], wantd.view(torch.int16)[~nand])


@pytest.mark.parametrize("shape", [(4096, 1024), (77, 264), (5, 8)])
def test_dswiglu_quant(shape):
    ops = _ops()
    M, I = shape
    fc1 = (gen((M, 2 * I), 7) * 3).bfloat16()
    da = (gen((M, I), 8) * 1e-2).bfloat16()
    d = co.dswiglu_f32(da, fc1)
    scale = 57344.0 / d.abs().max().item()
    for dtype in (E5M2, E4M3):
        qo = co.cast_f32(d * torch.tensor(scale, dtype=torch.float32), dtype)
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q, dbias = ops.dswiglu_quant_raw(dev(da), dev(fc1), FMT[dtype], s, amax)
        assert_fp8_equal(q, qo, dtype)
        assert amax.item() == d.abs().max().item()
        want = d.double().sum(0).float()
        assert (dbias.float().cpu() - want).abs().max().item() <= 1e-2 * want.abs().max().item() + 1e-6
    db, dbias2 = ops.dswiglu_bf16(dev(da), dev(fc1))
    assert torch.equal(db.cpu(), d.bfloat16())


def test_amax_scale_update_history_wrap():
    from llm_fp8_b200 import fp8
    T, L = 7, 16
    fmax = torch.tensor([448.0, 448.0, 57344.0, 448.0, 57344.0, 448.0, 448.0])
    hist = torch.zeros(L, T)
    scale, sinv = torch.ones(T), torch.ones(T)
    dh, ds, di, df = hist.cuda(), scale.cuda(), sinv.cuda(), fmax.cuda()
    g = torch.Generator().manual_seed(5)
    for step in range(40):
        am = torch.rand(T, generator=g) * 10 ** float(torch.randint(-3, 4, (1,), generator=g))
        if step == 7:
            am[1] = float("inf")
        if step in (20, 21):
            am[3] = 0.0
        hist[0] = am
        dh[0] = am.cuda()
        co.amax_scale_update(hist, scale, sinv, fmax, 0)
        fp8.amax_scale_update(dh, ds, di, df, T, 0)
        assert torch.equal(dh.cpu(), hist) and torch.equal(ds.cpu(), scale) and torch.equal(di.cpu(), sinv)


@pytest.mark.parametrize("shape", [(6144, 4096), (256, 384), (128, 128), (96, 160), (4096, 28672 // 8)])
def test_mx_quantize(shape):
    ops = _ops()
    R, C = shape
    x = gen(shape, 10, "heavy")
    x[5 % R, 32:64] = 0.0
    if R > 130:
        x[130, 7] = 3.0e38
    x[:, 100] = (x[:, 100].float() * 1e-30).bfloat16()
    want = co.mx_quantize_f32(x.float())
    out = ops.mx_quantize(dev(x), True, True)
    assert_fp8_equal(out.row, want["q_row"], E4M3)
    assert torch.equal(out.sf_row.cpu(), want["sf_row"])
    assert_fp8_equal(out.col, want["q_col"], E4M3)
    assert torch.equal(out.sf_col.cpu(), want["sf_col"])
    only_row = ops.mx_quantize(dev(x), True, False)
    assert only_row.col is None and torch.equal(only_row.row.cpu(), out.row.cpu())


@pytest.mark.parametrize("shape", [(6144, 4096), (256, 384), (96, 160), (992, 1056)])
def test_mx_quantize_bgrad(shape):
    """llmfp8_mx_quantize_bgrad: the block quantiser of a gradient dY with the bias gradient (column sums) taken from the tiles in shared
    memory — bytes and scales bit-identical to the oracle (as the plain quantiser), sums against the oracle's column sum."""
    ops = _ops()
    R, C = shape
    x = (gen(shape, 12, "heavy") * 1e-2).bfloat16()
    want = co.mx_quantize_f32(x.float())
    out, bgrad = ops.mx_quantize_bgrad(dev(x))
    assert_fp8_equal(out.row, want["q_row"], E4M3)
    assert torch.equal(out.sf_row.cpu(), want["sf_row"])
    assert_fp8_equal(out.col, want["q_col"], E4M3)
    assert torch.equal(out.sf_col.cpu(), want["sf_col"])
    ref = co.colsum_bf16(x)
    tol = 1e-2 * ref.abs().max().item() + 1e-6
    assert (bgrad.float().cpu() - ref).abs().max().item() <= tol


@pytest.mark.parametrize("shape", [(6144, 4096), (160, 2048), (64, 5120)])
def test_mx_rmsnorm_quantize(shape):
    ops = _ops()
    M, H = shape
    x = gen(shape, 11)
    g = torch.Generator().manual_seed(12)
    gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
    y, rso = co.rmsnorm_y(x, gamma, 1e-5)
    want = co.mx_quantize_f32(y)
    rs = ops.rmsnorm_rsigma(dev(x), 1e-5)
    assert torch.equal(rs.cpu(), rso)
    out = ops.mx_quantize(dev(x), True, True, rs, dev(gamma))
    assert_fp8_equal(out.row, want["q_row"], E4M3)
    assert torch.equal(out.sf_row.cpu(), want["sf_row"])
    assert_fp8_equal(out.col, want["q_col"], E4M3)
    assert torch.equal(out.sf_col.cpu(), want["sf_col"])


# ------------------------------------------------------------------------------------------------ column slices (row pitch > cols)
@pytest.mark.parametrize("R,C,W,off", [(6144, 4096, 6144, 0), (333, 80, 200, 16), (64, 2048, 2048 + 8, 8), (1, 16, 48, 32)])
def test_quantisers_read_column_slices_in_place(R, C, W, off):
    """x = wide[:, off:off+C] handed over as a view (the query slice of a fused qkv output): bytes, scales and amax must equal the
    oracle's on the dense copy — per-tensor cast and the MXFP8 block quantiser."""
    ops = _ops()
    wide = gen((R, W), 21, "heavy")
    x = wide[:, off:off + C].contiguous()
    view = dev(wide)[:, off:off + C]
    assert R == 1 or not view.is_contiguous()
    scale = 448.0 / fo.amax_of(x).item()
    for dtype in (E4M3, E5M2):
        s = torch.tensor([scale], dtype=torch.float32).cuda()
        amax = torch.zeros(1, dtype=torch.float32).cuda()
        q = ops.quantize_raw(view, FMT[dtype], s, amax)
        qo, ao = fo.quantize(x, scale, dtype)
        assert_fp8_equal(q, qo.view(torch.uint8), dtype)
        assert amax.item() == ao.item()
    if C % 32 == 0 and R % 32 == 0:
        want = co.mx_quantize_f32(x.float())
        out = ops.mx_quantize(view, True, True)
        assert_fp8_equal(out.row, want["q_row"], E4M3)
        assert torch.equal(out.sf_row.cpu(), want["sf_row"])
        assert_fp8_equal(out.col, want["q_col"], E4M3)
        assert torch.equal(out.sf_col.cpu(), want["sf_col"])


def test_rmsnorm_bwd_with_residual_gradient():
    """K5 + dres: equals the separate kernel result plus dres up to ONE bf16 rounding of the fp32 sum (checked against the fp32
    reference), and dgamma is untouched by the residual branch."""
    ops = _ops()
    for M, H in ((300, 2048), (6144, 4096), (33, 5120), (7, 136)):
        x, dy, dres = gen((M, H), 4), gen((M, H), 5), gen((M, H), 6)
        g = torch.Generator().manual_seed(9)
        gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
        rs = ops.rmsnorm_rsigma(dev(x), 1e-5)
        dx0, dg0 = ops.rmsnorm_bwd(dev(dy), dev(x), dev(gamma), rs)
        dx1, dg1 = ops.rmsnorm_bwd(dev(dy), dev(x), dev(gamma), rs, add=dev(dres))
        assert torch.equal(dg0, dg1)
        xf = x.float().requires_grad_(True)
        y, _ = fo.rmsnorm_ref(xf, gamma.float(), 1e-5)
        y.backward(dy.float())
        want = xf.grad + dres.float()
        assert torch.allclose(dx1.float().cpu(), want, rtol=2e-2, atol=2e-2 * want.abs().max().item())
        # against the unfused pair: differs by at most the double rounding
        two_step = (dx0.float() + dev(dres).float())
        assert (dx1.float() - two_step).abs().max().item() <= 2.0 ** -7 * two_step.abs().max().item()


@pytest.mark.parametrize("M,H", [(1, 16), (2, 4096), (5, 8192), (700, 4096), (1200, 3072)])
@pytest.mark.parametrize("with_dres", [False, True])
def test_rmsnorm_bwd_ring_edges(M, H, with_dres):
    """K5's shared-memory ring (rmsnorm.cu: rows streamed with cp.async.bulk, thread 0 refills a slot right after the row barrier): fewer
    rows than ring slots, more rows than resident CTAs x slots (several wraps of the mbarrier phases), the smallest and the largest hidden
    size, with and without the residual-branch gradient — against the fp32 autograd reference."""
    ops = _ops()
    x, dy, dres = gen((M, H), 14), gen((M, H), 15), gen((M, H), 16)
    g = torch.Generator().manual_seed(19)
    gamma = (1 + 0.1 * torch.randn(H, generator=g)).bfloat16()
    rs = ops.rmsnorm_rsigma(dev(x), 1e-5)
    dx, dgamma = ops.rmsnorm_bwd(dev(dy), dev(x), dev(gamma), rs, add=dev(dres) if with_dres else None)
    xf = x.float().requires_grad_(True)
    gf = gamma.float().requires_grad_(True)
    y, _ = fo.rmsnorm_ref(xf, gf, 1e-5)
    y.backward(dy.float())
    want = xf.grad + (dres.float() if with_dres else 0)
    assert torch.allclose(dx.float().cpu(), want, rtol=2e-2, atol=2e-2 * want.abs().max().item())
    assert torch.allclose(dgamma.float().cpu(), gf.grad, rtol=2e-2, atol=2e-2 * gf.grad.abs().max().item())


# ------------------------------------------------------------------------------------------------ activation + MX quantiser in one pass
import os  # noqa: E402

def _assert_mx_equal(got, want):
    for name in ("row", "sf_row", "col", "sf_col"):
        g, w = getattr(got, name), getattr(want, name)
        assert (g is None) == (w is None), name
        if g is not None:
            assert torch.equal(g, w), f"{name}: {int((g != w).sum())} of {g.numel()} bytes differ"


@pytest.mark.parametrize("M,I", [(128, 128), (256, 512), (1000 // 32 * 32, 1056), (96, 160), (6144, 14336)])
@pytest.mark.parametrize("kind", ["normal", "heavy", "outlier"])
def test_mx_swiglu_quantize_matches_two_kernel_sequence(M, I, kind):
    """llmfp8_mx_swiglu_quantize == llmfp8_swiglu_quant_fwd(bf16) -> llmfp8_mx_quantize (both pinned to the oracle above), byte for
    byte: data and swizzled E8M0 scales, rowwise and colwise, full and ragged tiles (I not a multiple of 128: the gate tile must
    zero-fill past column I, not read the up half)."""
    ops = _ops()
    if (M, I) == (6144, 14336) and kind != "normal":
        pytest.skip("full Llama-3.1-8B shape once")
    fc1 = dev(gen((M, 2 * I), 900 + M + I, kind) * 3)
    want = ops.mx_quantize(ops.swiglu_bf16(fc1), True, True)
    got = ops.mx_swiglu_quantize(fc1, True, True)
    torch.cuda.synchronize()
    _assert_mx_equal(got, want)
    row_only = ops.mx_swiglu_quantize(fc1, True, False)
    assert row_only.col is None and torch.equal(row_only.row, want.row) and torch.equal(row_only.sf_row, want.sf_row)


@pytest.mark.parametrize("M,I", [(128, 128), (256, 512), (992, 1152), (96, 384), (6144, 14336)])
@pytest.mark.parametrize("kind", ["normal", "heavy"])
def test_mx_dswiglu_quantize_matches_two_kernel_sequence(M, I, kind):
    """llmfp8_mx_dswiglu_quantize == llmfp8_dswiglu_quant_bwd(bf16) -> llmfp8_mx_quantize byte for byte; the bias gradient (fp32 column
    sums accumulated with atomics in both versions, in different orders) within bf16 rounding of an fp64 sum."""
    ops = _ops()
    if (M, I) == (6144, 14336) and kind != "normal":
        pytest.skip("full Llama-3.1-8B shape once")
    fc1 = dev(gen((M, 2 * I), 700 + M + I, kind) * 3)
    da = dev(gen((M, I), 701 + M + I, kind))
    d, db_want = ops.dswiglu_bf16(da, fc1, want_dbias=True)
    want = ops.mx_quantize(d, True, True)
    got, db = ops.mx_dswiglu_quantize(da, fc1, True, True, want_dbias=True)
    torch.cuda.synchronize()
    _assert_mx_equal(got, want)
    ref = db_want.float()
    assert torch.isfinite(db.float()).all()
    assert (db.float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item() + 1e-6
    got2, none = ops.mx_dswiglu_quantize(da, fc1, True, True, want_dbias=False)
    assert none is None
    _assert_mx_equal(got2, want)


def test_mx_act_quantize_contract_errors():
    ops = _ops()
    fc1 = torch.zeros(64, 2 * 48, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError):
        ops.mx_swiglu_quantize(fc1)                      # I % 32 != 0
    fc1 = torch.zeros(64, 2 * 96, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError):
        ops.mx_dswiglu_quantize(torch.zeros(64, 96, dtype=torch.bfloat16, device="cuda"), fc1)   # I % 128 != 0
    with pytest.raises(RuntimeError):
        ops.mx_swiglu_quantize(torch.zeros(40, 2 * 128, dtype=torch.bfloat16, device="cuda"), True, True)   # colwise needs M % 32 == 0
