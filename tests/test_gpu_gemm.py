"""FP8 / MXFP8 GEMMs through the C ABI against the oracle's dequantise -> fp32 matmul (BASELINE.json: rel. tol 1e-2 on bf16)."""
import os

import pytest
import torch

from oracle import fp8_oracle as fo

pytestmark = pytest.mark.gpu
E4M3, E5M2 = fo.E4M3, fo.E5M2
FMT = {E4M3: 0, E5M2: 1}
REL_TOL = 1e-2          # BASELINE.json north star: <= 1e-2 on bf16 outputs (we observe ~2^-9 from the bf16 rounding)


def _ops():
    from llm_fp8_b200 import ops
    return ops


def rand_fp8(shape, dtype, seed, spread=1.0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g) * spread
    return fo.sat_cast(x, dtype)


def ref_matmul(a_f32, b_f32):
    """[M,K] x [N,K]^T in fp64 on the GPU (exact products, ~exact accumulation)."""
    return (a_f32.double().cuda() @ b_f32.double().cuda().t()).float().cpu()


def check(got, want, tol=REL_TOL):
    got = got.float().cpu()
    denom = want.abs().max().item() + 1e-30
    err = (got - want).abs().max().item() / denom
    assert err <= tol, f"max abs err / max|ref| = {err:.3e}"
    return err


def qt(ops, t_fp8, dtype, scale_inv):
    return ops.QTensor(t_fp8.view(torch.uint8).cuda().contiguous(), FMT[dtype], torch.tensor([scale_inv], dtype=torch.float32).cuda(), 0)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (256, 512, 512), (384, 768, 4096), (200, 264, 160), (16, 8, 16), (1000, 3000, 2064)])
def test_fp8_gemm_layouts(M, N, K, a_mn, b_mn):
    ops = _ops()
    if (a_mn and M % 16) or (b_mn and N % 16):
        pytest.skip("MN-major operands need a 16 B pitch")
    a = rand_fp8((M, K), E4M3, 1)
    b = rand_fp8((N, K), E4M3, 2)
    sa, sb = 0.37, 1.9
    want = ref_matmul(a.float() * sa, b.float() * sb)
    A = qt(ops, a.t().contiguous() if a_mn else a, E4M3, sa)
    B = qt(ops, b.t().contiguous() if b_mn else b, E4M3, sb)
    got = ops.fp8_gemm(A, a_mn, B, b_mn, M, N, K)
    check(got, want)


@pytest.mark.parametrize("fa,fb", [(E5M2, E4M3), (E4M3, E5M2), (E5M2, E5M2)])
def test_fp8_gemm_mixed_formats(fa, fb):
    ops = _ops()
    M, N, K = 256, 512, 384
    a, b = rand_fp8((M, K), fa, 3, 4.0), rand_fp8((N, K), fb, 4, 0.1)
    want = ref_matmul(a.float(), b.float())
    got = ops.fp8_gemm(qt(ops, a, fa, 1.0), False, qt(ops, b, fb, 1.0), False, M, N, K)
    check(got, want)
    # dgrad-style: e5m2 gradient x e4m3 weight read MN-major
    got = ops.fp8_gemm(qt(ops, a, fa, 1.0), False, qt(ops, b.t().contiguous(), fb, 1.0), True, M, N, K)
    check(got, want)


def test_fp8_gemm_bias_accumulate_fp32():
    ops = _ops()
    M, N, K = 320, 520, 256
    a, b = rand_fp8((M, K), E4M3, 5), rand_fp8((N, K), E4M3, 6)
    g = torch.Generator().manual_seed(7)
    bias = torch.randn(N, generator=g).bfloat16()
    base = ref_matmul(a.float() * 0.5, b.float() * 0.25)
    A, B = qt(ops, a, E4M3, 0.5), qt(ops, b, E4M3, 0.25)
    got = ops.fp8_gemm(A, False, B, False, M, N, K, bias=bias.cuda())
    check(got, base + bias.float())
    out32 = torch.full((M, N), 3.0, dtype=torch.float32).cuda()
    ops.fp8_gemm(A, False, B, False, M, N, K, out=out32, accumulate=True)
    check(out32, base + 3.0, tol=1e-5)
    out16 = torch.full((M, N), 2.0, dtype=torch.bfloat16).cuda()
    ops.fp8_gemm(A, False, B, False, M, N, K, out=out16, accumulate=True)
    check(out16, base + 2.0)
    out32b = ops.fp8_gemm(A, False, B, False, M, N, K, out_dtype=torch.float32)
    check(out32b, base, tol=1e-5)


def test_fp8_gemm_persistence_and_linearity():
    """More tiles than CTAs (ring + TMEM double-buffer wrap) and a size-independent property at a full Llama-8B shape."""
    ops = _ops()
    M, N, K = 1024, 2048, 1024
    a, b = rand_fp8((M, K), E4M3, 8), rand_fp8((N, K), E4M3, 9)
    want = ref_matmul(a.float(), b.float())
    for max_ctas in (1, 3, 0):
        got = ops.fp8_gemm(qt(ops, a, E4M3, 1.0), False, qt(ops, b, E4M3, 1.0), False, M, N, K, max_ctas=max_ctas)
        check(got, want)
    # full size: qkv fprop at Llama-3.1-8B (M=6144, N=6144, K=4096): alpha-linearity and agreement with fp32 torch on a row sample
    M, N, K = 6144, 6144, 4096
    a, b = rand_fp8((M, K), E4M3, 10), rand_fp8((N, K), E4M3, 11)
    A1, B1 = qt(ops, a, E4M3, 1.0), qt(ops, b, E4M3, 1.0)
    A2 = qt(ops, a, E4M3, 0.25)
    y1 = ops.fp8_gemm(A1, False, B1, False, M, N, K, out_dtype=torch.float32)
    y2 = ops.fp8_gemm(A2, False, B1, False, M, N, K, out_dtype=torch.float32)
    assert torch.equal(y2, y1 * 0.25)
    rows = torch.tensor([0, 1, 127, 128, 3000, 6143])
    want = ref_matmul(a[rows].float(), b.float())
    check(y1[rows.cuda()], want, tol=1e-5)


# ------------------------------------------------------------------------------------------------ MXFP8
def mx_operands(rows, K, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, K, generator=g) * torch.exp(torch.randn(rows, 1, generator=g))).bfloat16()
    return x


@pytest.mark.parametrize("M,N,K", [(128, 256, 128), (256, 512, 512), (384, 768, 4096), (200, 264, 160), (1000, 3000, 2080)])
def test_mxfp8_gemm_fprop(M, N, K):
    ops = _ops()
    x, w = mx_operands(M, K, 1), mx_operands(N, K, 2)
    xq, xe = fo.mx_quantize_rowwise(x)
    wq, we = fo.mx_quantize_rowwise(w)
    want = ref_matmul(fo.mx_dequantize_rowwise(xq, xe), fo.mx_dequantize_rowwise(wq, we))
    got = ops.mxfp8_gemm(xq.view(torch.uint8).cuda(), fo.swizzle_sf(xe).cuda(), False, wq.view(torch.uint8).cuda(), fo.swizzle_sf(we).cuda(), False,
                         M, N, K)
    check(got, want)


@pytest.mark.parametrize("M,N,K", [(256, 512, 512), (384, 768, 1024), (160, 272, 96)])
def test_mxfp8_gemm_dgrad_wgrad_layouts(M, N, K):
    """dgrad: A rowwise K-major, B colwise MN-major.  wgrad: both colwise MN-major."""
    ops = _ops()
    a, b = mx_operands(M, K, 3), mx_operands(N, K, 4)
    # B stored [K, N] (contraction K is the row dim): its colwise quantisation has blocks along K
    bt = b.t().contiguous()                                   # [K, N]
    bq, be = fo.mx_quantize_colwise(bt)                       # data [K, N], scales [N, K/32]
    aq, ae = fo.mx_quantize_rowwise(a)
    want = ref_matmul(fo.mx_dequantize_rowwise(aq, ae), fo.mx_dequantize_colwise(bq, be).t())
    got = ops.mxfp8_gemm(aq.view(torch.uint8).cuda(), fo.swizzle_sf(ae).cuda(), False, bq.view(torch.uint8).cuda(), fo.swizzle_sf(be).cuda(), True,
                         M, N, K)
    check(got, want)
    at = a.t().contiguous()                                   # [K, M]
    aq2, ae2 = fo.mx_quantize_colwise(at)
    want = ref_matmul(fo.mx_dequantize_colwise(aq2, ae2).t(), fo.mx_dequantize_colwise(bq, be).t())
    got = ops.mxfp8_gemm(aq2.view(torch.uint8).cuda(), fo.swizzle_sf(ae2).cuda(), True, bq.view(torch.uint8).cuda(), fo.swizzle_sf(be).cuda(), True,
                         M, N, K)
    check(got, want)


def test_mxfp8_quantize_then_gemm_roundtrip():
    """Kernel-produced MX tensors feed the kernel GEMM: fprop + dgrad + wgrad of one Linear vs the fp32 reference on dequantised data."""
    ops = _ops()
    T, N, K = 512, 768, 1024
    x, w, dy = mx_operands(T, K, 5), mx_operands(N, K, 6), mx_operands(T, N, 7)
    xq, wq, dyq = ops.mx_quantize(x.cuda()), ops.mx_quantize(w.cuda()), ops.mx_quantize(dy.cuda())
    def deq_row(t, R, C):
        return fo.mx_dequantize_rowwise(t.row.cpu().view(E4M3), fo.unswizzle_sf(t.sf_row.cpu(), R, C // 32))
    def deq_col(t, R, C):
        return fo.mx_dequantize_colwise(t.col.cpu().view(E4M3), fo.unswizzle_sf(t.sf_col.cpu(), C, R // 32))
    y = ops.mxfp8_gemm(xq.row, xq.sf_row, False, wq.row, wq.sf_row, False, T, N, K)
    check(y, ref_matmul(deq_row(xq, T, K), deq_row(wq, N, K)))
    dx = ops.mxfp8_gemm(dyq.row, dyq.sf_row, False, wq.col, wq.sf_col, True, T, K, N)
    check(dx, ref_matmul(deq_row(dyq, T, N), deq_col(wq, N, K).t()))
    dw = ops.mxfp8_gemm(dyq.col, dyq.sf_col, True, xq.col, xq.sf_col, True, N, K, T)
    check(dw, ref_matmul(deq_col(dyq, T, N).t(), deq_col(xq, T, K).t()))


# ------------------------------------------------------------------------------------------------ tile-width / CTA-group variants
@pytest.fixture()
def force():
    from llm_fp8_b200 import _lib
    lib = _lib.load()

    def f(bn, cg):
        lib.llmfp8_gemm_force_bn(bn)
        lib.llmfp8_gemm_force_cg(cg)
    yield f
    lib.llmfp8_gemm_force_bn(0)
    lib.llmfp8_gemm_force_cg(0)


VARIANTS = [(1, 128), (1, 192), (1, 256), (2, 128), (2, 256)]


@pytest.mark.parametrize("cg,bn", VARIANTS + [(2, 192)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_fp8_gemm_forced_variant(cg, bn, a_mn, b_mn, force):
    """Every (CTA group, tile width) the host heuristic may pick, on a shape with partial tiles in M and N and > 148 tiles."""
    ops = _ops()
    if cg == 2 and bn == 192 and b_mn:
        pytest.skip("not built: 96 MN-major columns per CTA")
    force(bn, cg)
    M, N, K = 2000, 3280, 640
    a, b = rand_fp8((M, K), E4M3, 31), rand_fp8((N, K), E5M2, 32, 0.5)
    want = ref_matmul(a.float(), b.float())
    A = qt(ops, a.t().contiguous() if a_mn else a, E4M3, 1.0)
    B = qt(ops, b.t().contiguous() if b_mn else b, E5M2, 1.0)
    check(ops.fp8_gemm(A, a_mn, B, b_mn, M, N, K), want)
    g = torch.Generator().manual_seed(5)
    bias = torch.randn(N, generator=g).bfloat16()
    got = ops.fp8_gemm(A, a_mn, B, b_mn, M, N, K, bias=bias.cuda(), out_dtype=torch.float32)
    check(got, want + bias.float(), tol=1e-5)


@pytest.mark.parametrize("cg,bn", VARIANTS + [(2, 192)])
def test_mxfp8_gemm_forced_variant(cg, bn, force):
    """MX with every variant: 192 starts every other tile 64 rows into a 128-row scale tile (TMEM column offset 2)."""
    ops = _ops()
    force(bn, cg)
    M, N, K = 640, 1600, 1024
    x, w = mx_operands(M, K, 41), mx_operands(N, K, 42)
    xq, xe = fo.mx_quantize_rowwise(x)
    wq, we = fo.mx_quantize_rowwise(w)
    want = ref_matmul(fo.mx_dequantize_rowwise(xq, xe), fo.mx_dequantize_rowwise(wq, we))
    got = ops.mxfp8_gemm(xq.view(torch.uint8).cuda(), fo.swizzle_sf(xe).cuda(), False, wq.view(torch.uint8).cuda(), fo.swizzle_sf(we).cuda(), False, M, N, K)
    check(got, want)
    if cg == 2 and bn == 192:
        return
    # dgrad / wgrad layouts at the same width
    bt = w.t().contiguous()
    bq, be = fo.mx_quantize_colwise(bt)
    want = ref_matmul(fo.mx_dequantize_rowwise(xq, xe), fo.mx_dequantize_colwise(bq, be).t())
    got = ops.mxfp8_gemm(xq.view(torch.uint8).cuda(), fo.swizzle_sf(xe).cuda(), False, bq.view(torch.uint8).cuda(), fo.swizzle_sf(be).cuda(), True, M, N, K)
    check(got, want)
    at = x.t().contiguous()
    aq2, ae2 = fo.mx_quantize_colwise(at)
    want = ref_matmul(fo.mx_dequantize_colwise(aq2, ae2).t(), fo.mx_dequantize_colwise(bq, be).t())
    got = ops.mxfp8_gemm(aq2.view(torch.uint8).cuda(), fo.swizzle_sf(ae2).cuda(), True, bq.view(torch.uint8).cuda(), fo.swizzle_sf(be).cuda(), True, M, N, K)
    check(got, want)


@pytest.mark.parametrize("cg", [1, 2])
def test_gemm_full_shape_checksum(cg, force):
    """Full Llama-3.1-8B fc2 fprop shape through both kernels: identical results (same fp32 accumulation order per output) and a
    column-sum identity against an fp64 reference: sum_m Y[m, :] == (sum_m A[m, :]) B^T."""
    ops = _ops()
    force(0, cg)
    M, N, K = 6144, 4096, 14336
    a, b = rand_fp8((M, K), E4M3, 51), rand_fp8((N, K), E4M3, 52)
    y = ops.fp8_gemm(qt(ops, a, E4M3, 1.0), False, qt(ops, b, E4M3, 1.0), False, M, N, K, out_dtype=torch.float32)
    colsum = y.double().sum(0).cpu()
    want = (a.double().cuda().sum(0, keepdim=True) @ b.double().cuda().t()).cpu()[0]
    err = (colsum - want).abs().max().item() / want.abs().max().item()
    assert err < 1e-5, err


@pytest.mark.parametrize("M,N,K", [(6144, 28672, 4096), (992, 1344, 992), (512, 768, 4096)])
@pytest.mark.parametrize("layout", ["fprop", "dgrad", "wgrad"])
def test_mxfp8_gemm_overlapped_accumulators_identical(M, N, K, layout, force):
    """MX at 256 columns runs two TMEM accumulators that share 32 columns (gemm.cu: OVL; 64 columns with debug bit 8); debug bit 4 selects the single-accumulator
    kernel.  Both issue the same MMAs in the same order, so bf16 and fp32 outputs must be bit-identical — on the Llama-3.1-8B fc1 shape
    (36 tiles per CTA pair), a ragged shape and a single-CTA shape.  Same calls as tools/gemm_ovl_check.py
    (profiles/r01c_gemm_ovl_check.txt)."""
    ops = _ops()
    from llm_fp8_b200 import _lib
    lib = _lib.load()
    force(256, 0)
    dev = "cuda"

    def run(dt):
        g = torch.Generator(device=dev).manual_seed(M + 3 * N + 7 * K)
        r = lambda *s: torch.randint(0, 120, s, device=dev, dtype=torch.uint8, generator=g)
        sf = lambda rows, k: torch.randint(120, 134, (ops.mx_scale_bytes(rows, k),), device=dev, dtype=torch.uint8, generator=g)
        x, w, dy = r(M, K), r(N, K), r(M, N)
        if layout == "fprop":
            return ops.mxfp8_gemm(x, sf(M, K), False, w, sf(N, K), False, M, N, K, out_dtype=dt)
        if layout == "dgrad":
            return ops.mxfp8_gemm(dy, sf(M, N), False, w, sf(K, N), True, M, K, N, out_dtype=dt)
        return ops.mxfp8_gemm(dy, sf(N, M), True, x, sf(K, M), True, N, K, M, out_dtype=dt)

    try:
        for dt in (torch.bfloat16, torch.float32):
            lib.llmfp8_gemm_debug(16)
            single = run(dt)
            lib.llmfp8_gemm_debug(0)
            ovl = run(dt)
            lib.llmfp8_gemm_debug(256)                    # accumulators 192 apart: 64 shared columns, four scale buffers (the first form)
            narrow = run(dt)
            torch.cuda.synchronize()
            assert torch.isfinite(ovl.float()).all()
            assert torch.equal(single, ovl), f"{layout} {dt}"
            assert torch.equal(single, narrow), f"{layout} {dt} (64 shared columns)"
    finally:
        lib.llmfp8_gemm_debug(0)


@pytest.mark.parametrize("M,N,K", [(6144, 4096, 14336), (1000, 1544, 640), (512, 768, 4096), (300, 264, 160)])
def test_mxfp8_gemm_residual_overlapped_accumulators_identical(M, N, K, force):
    """The fused-residual MX epilogue on the overlapped accumulators (gemm.cu: RES + OVL, the default) against the single-accumulator
    residual kernel (debug bit 5).  Same MMAs, same fp32 sum, same single bf16 rounding -> bit-identical, ragged edges included
    (the downward chunk walk meets the out-of-range chunks first)."""
    ops = _ops()
    from llm_fp8_b200 import _lib
    lib = _lib.load()
    force(256, 0)
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(M + N + K)
    x = torch.randint(0, 120, (M, K), device=dev, dtype=torch.uint8, generator=g)
    w = torch.randint(0, 120, (N, K), device=dev, dtype=torch.uint8, generator=g)
    sx = torch.randint(120, 134, (ops.mx_scale_bytes(M, K),), device=dev, dtype=torch.uint8, generator=g)
    sw = torch.randint(120, 134, (ops.mx_scale_bytes(N, K),), device=dev, dtype=torch.uint8, generator=g)
    res = torch.randn(M, N, device=dev, generator=g).bfloat16() * 100
    bias = torch.randn(N, device=dev, generator=g).bfloat16()
    try:
        lib.llmfp8_gemm_debug(32)
        want = ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K, bias=bias, residual=res)
        lib.llmfp8_gemm_debug(0)
        got = ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K, bias=bias, residual=res)
        lib.llmfp8_gemm_debug(256)                        # accumulators 192 apart (64 shared columns: the first form)
        narrow = ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K, bias=bias, residual=res)
        torch.cuda.synchronize()
        assert torch.equal(want, got)
        assert torch.equal(want, narrow)
    finally:
        lib.llmfp8_gemm_debug(0)


# ------------------------------------------------------------------------------------------------ fused residual epilogue
@pytest.mark.parametrize("cg,bn", [(0, 0), (1, 128), (1, 256), (2, 192), (2, 256)])
def test_fp8_gemm_fused_residual(cg, bn, force):
    """D = A B^T (+bias) + R with R a bf16 matrix read through its own leading dimension: partial tiles in M and N, R both dense and
    as a column slice of a wider matrix.  One rounding (fp32 sum -> bf16), so the tolerance is the plain GEMM's."""
    ops = _ops()
    force(bn, cg)
    M, N, K = 1000, 1544, 640
    a, b = rand_fp8((M, K), E4M3, 61), rand_fp8((N, K), E4M3, 62)
    g = torch.Generator().manual_seed(63)
    bias = torch.randn(N, generator=g).bfloat16()
    wide = (torch.randn(M, N + 72, generator=g) * 20).bfloat16()
    base = ref_matmul(a.float() * 0.5, b.float() * 0.25)
    A, B = qt(ops, a, E4M3, 0.5), qt(ops, b, E4M3, 0.25)
    r_dense = wide[:, :N].contiguous()
    got = ops.fp8_gemm(A, False, B, False, M, N, K, residual=r_dense.cuda())
    check(got, base + r_dense.float())
    r_view = wide.cuda()[:, 8:8 + N]                       # pitch N + 72, 16 B aligned start
    got = ops.fp8_gemm(A, False, B, False, M, N, K, bias=bias.cuda(), residual=r_view)
    check(got, base + bias.float() + wide[:, 8:8 + N].float())
    # no residual given: the plain kernel still runs through the same wrapper
    check(ops.fp8_gemm(A, False, B, False, M, N, K), base)


def test_mxfp8_gemm_fused_residual_full_shape():
    """Llama-3.1-8B proj fprop (M=6144, N=K=4096) with the residual: equals the unfused result + R up to one bf16 rounding."""
    ops = _ops()
    M, N, K = 6144, 4096, 4096
    x, w = mx_operands(M, K, 71), mx_operands(N, K, 72)
    g = torch.Generator().manual_seed(73)
    r = torch.randn(M, N, generator=g).bfloat16().cuda()
    xq, wq = ops.mx_quantize(x.cuda(), True, False), ops.mx_quantize(w.cuda(), True, False)
    plain = ops.mxfp8_gemm(xq.row, xq.sf_row, False, wq.row, wq.sf_row, False, M, N, K, out_dtype=torch.float32)
    fused = ops.mxfp8_gemm(xq.row, xq.sf_row, False, wq.row, wq.sf_row, False, M, N, K, residual=r)
    want = (plain + r.float()).bfloat16()
    assert torch.equal(fused, want), f"{(fused.float() - want.float()).abs().max().item()}"


def test_fused_residual_contract_errors():
    ops = _ops()
    M, N, K = 128, 128, 128
    a, b = rand_fp8((M, K), E4M3, 1), rand_fp8((N, K), E4M3, 2)
    A, B = qt(ops, a, E4M3, 1.0), qt(ops, b, E4M3, 1.0)
    r = torch.zeros(M, N, dtype=torch.bfloat16).cuda()
    with pytest.raises(RuntimeError):
        ops.fp8_gemm(A, False, B, False, M, N, K, residual=r, out_dtype=torch.float32)
    with pytest.raises(RuntimeError):
        ops.fp8_gemm(A, False, qt(ops, b.t().contiguous(), E4M3, 1.0), True, M, N, K, residual=r)     # dgrad layout: not built


# ------------------------------------------------------------------------------------------------ MXFP8 dgrad / wgrad at Llama-3.1-8B shapes
@pytest.mark.parametrize("name,T,N,K", [("qkv", 6144, 6144, 4096), ("fc2", 6144, 4096, 14336)])
def test_mxfp8_dgrad_wgrad_at_8b_shapes_vs_oracle_sample(name, T, N, K):
    """BASELINE.json configs[3] sizes (M = 12 x 512 tokens): operands block-quantised by the ORACLE (oracle_c.c), the block-scaled GEMM in
    its dgrad (A rowwise K-major, B colwise MN-major) and wgrad (both colwise, MN-major) forms against dequantise -> fp64 matmul on a
    sample of output rows that touches the first / last tile, a tile boundary and the CTA-pair halves."""
    from oracle import c_oracle as co
    ops = _ops()
    x, w, dy = mx_operands(T, K, 51), mx_operands(N, K, 52), mx_operands(T, N, 53)
    xo = co.mx_quantize_f32(x.float(), rowwise=False, colwise=True)
    wo = co.mx_quantize_f32(w.float(), rowwise=False, colwise=True)
    dyo = co.mx_quantize_f32(dy.float(), rowwise=True, colwise=True)

    def deq_row(o, R, C):
        return fo.mx_dequantize_rowwise(o["q_row"].view(E4M3), fo.unswizzle_sf(o["sf_row"], R, C // 32))

    def deq_col(o, R, C):
        return fo.mx_dequantize_colwise(o["q_col"].view(E4M3), fo.unswizzle_sf(o["sf_col"], C, R // 32))

    c = lambda t: t.cuda()      # noqa: E731
    rows = torch.tensor([0, 1, 127, 128, 255, 256, 3071, 3072, T - 129, T - 1])
    dx = ops.mxfp8_gemm(c(dyo["q_row"]), c(dyo["sf_row"]), False, c(wo["q_col"]), c(wo["sf_col"]), True, T, K, N, out_dtype=torch.float32)
    check(dx[rows.cuda()], ref_matmul(deq_row(dyo, T, N)[rows], deq_col(wo, N, K).t()), tol=1e-5)
    nrows = torch.tensor([0, 1, 127, 128, 255, 256, N // 2 - 1, N // 2, N - 129, N - 1])
    dw = ops.mxfp8_gemm(c(dyo["q_col"]), c(dyo["sf_col"]), True, c(xo["q_col"]), c(xo["sf_col"]), True, N, K, T, out_dtype=torch.float32)
    check(dw[nrows.cuda()], ref_matmul(deq_col(dyo, T, N).t()[nrows], deq_col(xo, T, K).t()), tol=1e-5)
    dwb = ops.mxfp8_gemm(c(dyo["q_col"]), c(dyo["sf_col"]), True, c(xo["q_col"]), c(xo["sf_col"]), True, N, K, T)      # bf16, as the modules ask
    check(dwb[nrows.cuda()], ref_matmul(deq_col(dyo, T, N).t()[nrows], deq_col(xo, T, K).t()))


@pytest.mark.parametrize("cg", [1, 2])
def test_mxfp8_gemm_224_single_column_tile_matches_256(cg, force):
    """224-column tiles (two whole TMEM accumulators beside the scale factors) work — bit-identical to 256 — only while every tile starts on a
    scale-tile boundary, i.e. for N <= 224: any further tile starts 96 / 64 / 32 rows into a scale tile and the block-scaled MMA rejects the odd
    SFB column (gemm.cu: pick_bn).  This keeps the code under it (third SFB tile per stage, re-storing last epilogue chunk) exercised."""
    ops = _ops()
    M, N, K = 512, 224, 512
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randint(0, 120, (M, K), device="cuda", dtype=torch.uint8, generator=g)
    w = torch.randint(0, 120, (N, K), device="cuda", dtype=torch.uint8, generator=g)
    sx = torch.randint(120, 134, (ops.mx_scale_bytes(M, K),), device="cuda", dtype=torch.uint8, generator=g)
    sw = torch.randint(120, 134, (ops.mx_scale_bytes(N, K),), device="cuda", dtype=torch.uint8, generator=g)
    force(256, cg)
    want = ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K)
    force(224, cg)
    got = ops.mxfp8_gemm(x, sx, False, w, sw, False, M, N, K)
    torch.cuda.synchronize()
    assert torch.equal(want, got)
