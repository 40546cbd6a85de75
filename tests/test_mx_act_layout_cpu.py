"""Index model of csrc/mx_act_quantize.cu, checked on CPU against the oracle's layouts.

The arithmetic of the fused SwiGLU / dSwiGLU + MXFP8 kernels is shared, line for line, with kernels pinned on hardware (swiglu.cu,
mx_quantize.cu); what is their own is index logic — which tile lands where, which bytes of which 512 B scale tile it owns, when the bias
sums are flushed — and that part is replayed here in numpy-style torch:
each function below walks the tiles exactly as the CUDA code does (same formulas, cited) and must reproduce
oracle.fp8_oracle's full-matrix results.  No product code runs here; the GPU parity tests are in tests/test_gpu_quantize.py."""
import pytest
import torch

from oracle import fp8_oracle as fo

AT = 128


def _tile(x, r0, c0):
    """TMA load of a 128 x 128 box at (r0, c0): zero fill outside the tensor."""
    t = torch.zeros(AT, AT, dtype=x.dtype)
    r1, c1 = min(r0 + AT, x.shape[0]), min(c0 + AT, x.shape[1])
    if r1 > r0 and c1 > c0:
        t[: r1 - r0, : c1 - c0] = x[r0:r1, c0:c1]
    return t


def _store(dst, tile, r0, c0):
    """TMA store: clipped at the tensor edge."""
    r1, c1 = min(r0 + AT, dst.shape[0]), min(c0 + AT, dst.shape[1])
    if r1 > r0 and c1 > c0:
        dst[r0:r1, c0:c1] = tile[: r1 - r0, : c1 - c0]


def _sf_tile_row(e_tile):
    """act_row_pass: sf_row[(r & 31) * 16 + (r >> 5) * 4 + blk] = e(r, blk) for r < 128, blk < 4."""
    out = torch.zeros(512, dtype=torch.uint8)
    for r in range(AT):
        for blk in range(4):
            out[(r & 31) * 16 + (r >> 5) * 4 + blk] = e_tile[r, blk]
    return out


def _sf_tile_col(e_tile):
    """act_col_pass: sf_col[(c & 31) * 16 + (c >> 5) * 4 + rb] = e(c, rb) for c < 128, rb < 4."""
    return _sf_tile_row(e_tile)          # same formula with (c, rb) in place of (r, blk)


def _quantise_tile(a_tile):
    """Both block passes over one bf16 tile -> (q_row, e_row [128, 4], q_col, e_col [128, 4])."""
    q_row, e_row = fo.mx_quantize_rowwise(a_tile)
    q_col, e_col = fo.mx_quantize_colwise(a_tile)
    return q_row.view(torch.uint8), e_row, q_col.view(torch.uint8), e_col


def _tile_rows(x, r0, c0, rows):
    """TMA load of a rows x 128 box at (r0, c0): zero fill outside the tensor."""
    t = torch.zeros(rows, AT, dtype=x.dtype)
    r1, c1 = min(r0 + rows, x.shape[0]), min(c0 + AT, x.shape[1])
    if r1 > r0 and c1 > c0:
        t[: r1 - r0, : c1 - c0] = x[r0:r1, c0:c1]
    return t


def _store_rows(dst, tile, r0, c0):
    r1, c1 = min(r0 + tile.shape[0], dst.shape[0]), min(c0 + AT, dst.shape[1])
    if r1 > r0 and c1 > c0:
        dst[r0:r1, c0:c1] = tile[: r1 - r0, : c1 - c0]


def model_fwd(fc1, M, I, grid=3):
    """mx_swiglu_quantize2_kernel (mx_act_quantize2.cu): 64-row tiles over [M, I], row-tile count rounded up to even so every 128-row scale
    tile is written whole; gate view = fc1[:, :I], up view = fc1[:, I:] (two tensor maps, each bounded to I); each CTA owns a contiguous
    range of tiles; a tile fills 8 consecutive bytes per (row & 31) of its rowwise scale tile (staging sf_row[32][8]) and 2 per column of
    its colwise one (sf_col[128][2])."""
    R = 64
    gate, up = fc1[:, :I], fc1[:, I:]
    tiles_c, tiles_r128 = (I + AT - 1) // AT, (M + AT - 1) // AT
    tiles_r = 2 * tiles_r128
    ntiles = tiles_c * tiles_r
    q_row = torch.zeros(M, I, dtype=torch.uint8)
    q_col = torch.zeros(M, I, dtype=torch.uint8)
    sf_row = torch.full((tiles_r128 * tiles_c * 512,), 0xAA, dtype=torch.uint8)           # poison: every byte must be written
    sf_col = torch.full((tiles_c * tiles_r128 * 512,), 0xAA, dtype=torch.uint8)
    seen = 0
    for cta in range(grid):
        for t in range(ntiles * cta // grid, ntiles * (cta + 1) // grid):
            seen += 1
            tr, tc = t // tiles_c, t % tiles_c
            row0, col0 = tr * R, tc * AT
            g, u = _tile_rows(gate, row0, col0, R), _tile_rows(up, row0, col0, R)
            a = fo.swiglu_ref(torch.cat((g, u), 1)).bfloat16()
            q_r, e_r = fo.mx_quantize_rowwise(a)                      # e_r [64, 4]
            q_c, e_c = fo.mx_quantize_colwise(a)                      # e_c [128, 2]
            _store_rows(q_row, q_r.view(torch.uint8), row0, col0)
            _store_rows(q_col, q_c.view(torch.uint8), row0, col0)
            upper = 1 if (row0 & 64) else 0
            stage = torch.zeros(32, 8, dtype=torch.uint8)             # s.sf_row[r & 31][(r >> 5) * 4 + blk]
            for r in range(R):
                for blk in range(4):
                    stage[r & 31, (r >> 5) * 4 + blk] = e_r[r, blk]
            base = ((row0 >> 7) * tiles_c + tc) * 512
            for q in range(32):
                sf_row[base + q * 16 + upper * 8: base + q * 16 + upper * 8 + 8] = stage[q]
            base = (tc * tiles_r128 + (row0 >> 7)) * 512
            for c in range(AT):                                       # s.sf_col[c][rb]
                for rb in range(2):
                    sf_col[base + (c & 31) * 16 + (c >> 5) * 4 + upper * 2 + rb] = e_c[c, rb]
    assert seen == ntiles
    return q_row, sf_row, q_col, sf_col


@pytest.mark.parametrize("M,I", [(128, 128), (96, 160), (256, 288)])
def test_fwd_tile_walk_reproduces_full_matrix_layout(M, I):
    g = torch.Generator().manual_seed(M + I)
    fc1 = (torch.randn(M, 2 * I, generator=g) * 2).bfloat16()
    a = fo.swiglu_ref(fc1).bfloat16()
    want_q, want_e = fo.mx_quantize_rowwise(a)
    want_qc, want_ec = fo.mx_quantize_colwise(a)
    q_row, sf_row, q_col, sf_col = model_fwd(fc1, M, I)
    assert torch.equal(q_row, want_q.view(torch.uint8)) and torch.equal(q_col, want_qc.view(torch.uint8))
    assert torch.equal(sf_row, fo.swizzle_sf(want_e)) and torch.equal(sf_col, fo.swizzle_sf(want_ec))


def _dswiglu(da, g, u):
    sig = torch.sigmoid(g.float())
    dg = (da.float() * u.float()) * (sig * (1 + g.float() * (1 - sig)))
    du = da.float() * (g.float() * sig)
    return dg, du


def model_bwd(da, fc1, M, I, grid=3):
    """mx_dswiglu_quantize2_kernel (mx_act_quantize2.cu): 32-row tiles over [M, I], row-tile count rounded up to a multiple of four so every
    128-row scale tile is written whole; each CTA owns a contiguous range of tiles in COLUMN-major order; output [M, 2I] = (dgate | dup)
    with the dup half at tile column tiles_c + tc; a 32-row tile fills 4 consecutive bytes per (row & 31) of its rowwise scale tile and
    one per column of its colwise one (the staging arrays sf_row[2][32][4] / sf_col[2][128]); bias sums flushed when the tile column
    changes."""
    R = 32
    gate, up = fc1[:, :I], fc1[:, I:]
    tiles_c, tiles_r128 = I // AT, (M + AT - 1) // AT
    tiles_r = 4 * tiles_r128
    ntiles = tiles_c * tiles_r
    q_row = torch.zeros(M, 2 * I, dtype=torch.uint8)
    q_col = torch.zeros(M, 2 * I, dtype=torch.uint8)
    sf_row = torch.full((tiles_r128 * 2 * tiles_c * 512,), 0xAA, dtype=torch.uint8)      # poison: every byte must be written
    sf_col = torch.full((2 * tiles_c * tiles_r128 * 512,), 0xAA, dtype=torch.uint8)
    dbias = torch.zeros(2 * I, dtype=torch.float64)
    flushes = 0

    for cta in range(grid):
        t0, t1 = ntiles * cta // grid, ntiles * (cta + 1) // grid
        acc, cur_tc = torch.zeros(2 * AT, dtype=torch.float64), -1
        for t in range(t0, t1):
            tc, tr = t // tiles_r, t % tiles_r
            row0, col0 = tr * R, tc * AT
            if tc != cur_tc:
                if cur_tc >= 0:
                    dbias[cur_tc * AT:cur_tc * AT + AT] += acc[:AT]
                    dbias[I + cur_tc * AT:I + cur_tc * AT + AT] += acc[AT:]
                    flushes += 1
                acc, cur_tc = torch.zeros(2 * AT, dtype=torch.float64), tc
            dg, du = _dswiglu(_tile_rows(da, row0, col0, R), _tile_rows(gate, row0, col0, R), _tile_rows(up, row0, col0, R))
            acc[:AT] += dg.double().sum(0)
            acc[AT:] += du.double().sum(0)
            V = torch.cat((dg, du), 0).bfloat16()                       # the virtual 64 x 128 tile: rows 0..31 dgate, 32..63 dup
            q_r, e_r = fo.mx_quantize_rowwise(V)                        # e_r [64, 4]
            q_c, e_c = fo.mx_quantize_colwise(V)                        # e_c [128, 2]
            qr, qc = q_r.view(torch.uint8), q_c.view(torch.uint8)
            quarter = (row0 >> 5) & 3
            for h in range(2):
                r1 = min(row0 + R, M)
                if r1 > row0:
                    q_row[row0:r1, h * I + col0:h * I + col0 + AT] = qr[h * R:h * R + (r1 - row0)]
                    q_col[row0:r1, h * I + col0:h * I + col0 + AT] = qc[h * R:h * R + (r1 - row0)]
                # rowwise scale staging: s.sf_row[r >> 5][r & 31][blk], written out as 4 bytes at q * 16 + quarter * 4
                base = ((row0 >> 7) * (2 * tiles_c) + h * tiles_c + tc) * 512
                for r in range(h * R, h * R + R):
                    for blk in range(4):
                        sf_row[base + (r & 31) * 16 + quarter * 4 + blk] = e_r[r, blk]
                # colwise: s.sf_col[rb][c], one byte at (c & 31) * 16 + (c >> 5) * 4 + quarter
                base = ((h * tiles_c + tc) * tiles_r128 + (row0 >> 7)) * 512
                for c in range(AT):
                    sf_col[base + (c & 31) * 16 + (c >> 5) * 4 + quarter] = e_c[c, h]
        if cur_tc >= 0:
            dbias[cur_tc * AT:cur_tc * AT + AT] += acc[:AT]
            dbias[I + cur_tc * AT:I + cur_tc * AT + AT] += acc[AT:]
            flushes += 1
    assert flushes <= grid + tiles_c          # a few atomics per CTA, not one set per tile
    return q_row, sf_row, q_col, sf_col, dbias


@pytest.mark.parametrize("M,I", [(128, 128), (96, 256), (160, 384)])
def test_bwd_tile_walk_reproduces_full_matrix_layout(M, I):
    g = torch.Generator().manual_seed(7 * M + I)
    fc1 = (torch.randn(M, 2 * I, generator=g) * 2).bfloat16()
    da = torch.randn(M, I, generator=g).bfloat16()
    dg, du = _dswiglu(da, fc1[:, :I], fc1[:, I:])
    d = torch.cat((dg, du), 1)
    want_q, want_e = fo.mx_quantize_rowwise(d.bfloat16())
    want_qc, want_ec = fo.mx_quantize_colwise(d.bfloat16())
    q_row, sf_row, q_col, sf_col, dbias = model_bwd(da, fc1, M, I)
    assert torch.equal(q_row, want_q.view(torch.uint8)) and torch.equal(q_col, want_qc.view(torch.uint8))
    assert torch.equal(sf_row, fo.swizzle_sf(want_e)) and torch.equal(sf_col, fo.swizzle_sf(want_ec))
    assert torch.allclose(dbias, d.double().sum(0), rtol=1e-9, atol=1e-9)
