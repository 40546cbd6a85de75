// MXFP8 block quantiser for sm_100a (K2m / K1m).
//   in : bf16 [R, C]   (optionally y = (x * rsigma[r]) * gamma[c] applied on the fly: the RMSNorm variant)
//   out: rowwise copy  e4m3 [R, C] + E8M0 scales per 1x32 block,   scale matrix (r, c/32)
//        colwise copy  e4m3 [R, C] + E8M0 scales per 32x1 block,   scale matrix (c, r/32)
//        both scale matrices in the 128x4-tile swizzled layout that tcgen05.cp (and cuBLASLt) consume, so the
//        GEMM can bulk-copy 512 B tiles straight into shared memory and on into TMEM.
// Persistent CTAs walk 128x128 tiles: TMA load (double-buffered, mbarrier) -> registers -> fp8 tiles staged in
// shared memory -> TMA store.  A 128x128 tile is exactly one 512 B scale tile per direction.
// Replaces TE's mxfp8 quantize (cast_mxfp8_2D) reached under te_llama_mxfp8.py:86,93.
// Arithmetic = oracle_c.c: orc_mx_quantize_f32 (amax/448 with a true division, E8M0 round-up, x * 2^(127-e), satfinite RNE).
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

constexpr int MXT = 128;                    // tile edge
constexpr int MX_IN_BYTES = MXT * MXT * 2;  // 32 KB
constexpr int MX_OUT_BYTES = MXT * MXT;     // 16 KB

struct MxSmem {
    alignas(128) uint8_t in[2][MX_IN_BYTES];
    alignas(128) uint8_t out_row[MX_OUT_BYTES];
    alignas(128) uint8_t out_col[MX_OUT_BYTES];
    alignas(16) uint8_t sf_row[512];
    alignas(16) uint8_t sf_col[512];
    float rs[MXT];
    float colsum[4][MXT];                   // BG: per row-group column sums of the tile
    alignas(8) uint64_t full[2];
};

// BG (plain mode, colwise copy requested): also accumulates the column sums of x into bgrad_ws (fp32 [C], zeroed by the host) — the bias
// gradient of a Linear whose dY is being quantised, taken from the tile while it is in shared memory instead of a second pass over dY.
template <int MODE /*0 plain, 1 rmsnorm*/, bool BG = false>
__global__ void __launch_bounds__(256) mx_quantize_kernel(const __grid_constant__ CUtensorMap map_in, const __grid_constant__ CUtensorMap map_row,
                                                          const __grid_constant__ CUtensorMap map_col, uint8_t* __restrict__ sf_row,
                                                          uint8_t* __restrict__ sf_col, const float* __restrict__ rsigma,
                                                          const uint16_t* __restrict__ gamma, int R, int C, int do_row, int do_col,
                                                          float* __restrict__ bgrad_ws) {
    extern __shared__ uint8_t smem_raw[];
    // pointer arithmetic (not an integer round trip) so the compiler keeps the shared address space: LDS / STS, not generic LD / ST
    MxSmem& s = *reinterpret_cast<MxSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x;
    const int tiles_c = (C + MXT - 1) / MXT, tiles_r = (R + MXT - 1) / MXT;
    const int ntiles = tiles_c * tiles_r;

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_in);
        ptx::mbar_init(&s.full[0], 1);
        ptx::mbar_init(&s.full[1], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();
    if (tid == 0 && static_cast<int>(blockIdx.x) < ntiles) {
        const int t = blockIdx.x;
        ptx::mbar_arrive_expect_tx(&s.full[0], MX_IN_BYTES);
        ptx::tma_load_2d(&map_in, &s.full[0], s.in[0], (t % tiles_c) * MXT, (t / tiles_c) * MXT);
    }

    int it = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, it++) {
        const int stage = it & 1, phase = (it >> 1) & 1;
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * MXT, col0 = tc * MXT;
        if (tid == 0) {
            const int tn = t + gridDim.x;
            if (tn < ntiles) {
                ptx::mbar_arrive_expect_tx(&s.full[stage ^ 1], MX_IN_BYTES);
                ptx::tma_load_2d(&map_in, &s.full[stage ^ 1], s.in[stage ^ 1], (tn % tiles_c) * MXT, (tn / tiles_c) * MXT);
            }
            ptx::tma_store_wait_read<0>();   // previous tile's stores no longer read out_row/out_col
        }
        if (MODE == 1 && tid < MXT) s.rs[tid] = (row0 + tid < R) ? rsigma[row0 + tid] : 0.f;
        __syncthreads();
        ptx::mbar_wait(&s.full[stage], phase);
        const uint8_t* in = s.in[stage];

        if (do_row) {
            // thread: one whole 1x32 block (four 16 B segments of a row, read in a lane-rotated order so the 64 B lane stride stays
            // bank-conflict free); the block amax and its scale need no cross-lane traffic
#pragma unroll 1
            for (int i = 0; i < (MXT * MXT / 32) / 256; i++) {
                const int idx = i * 256 + tid;
                const int r = idx >> 2, blk = idx & 3;
                const int rot = (tid >> 1) & 3;
                const uint8_t* src = in + r * (MXT * 2) + blk * 64;
                float rs = 1.f;
                if (MODE == 1) rs = s.rs[r];
                f32x2 f[4][4];
                float am = 0.f;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int seg = (j + rot) & 3;
                    const uint4 v = *reinterpret_cast<const uint4*>(src + seg * 16);
                    float t[8];
                    unpack8(v, t);
                    if (MODE == 1) {
                        float gf[8];
                        uint4 gv = make_uint4(0, 0, 0, 0);
                        if (col0 + blk * 32 + seg * 8 < C) gv = __ldg(reinterpret_cast<const uint4*>(gamma + col0 + blk * 32 + seg * 8));
                        unpack8(gv, gf);
#pragma unroll
                        for (int k = 0; k < 8; k++) t[k] = __fmul_rn(__fmul_rn(t[k], rs), gf[k]);
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        am = fmaxf(am, fmaxf(fabsf(t[2 * k]), fabsf(t[2 * k + 1])));
                        f[j][k] = pk2(t[2 * k], t[2 * k + 1]);
                    }
                }
                const uint32_t e = e8m0_from_amax(am);
                const float inv = e8m0_inv_scale(e);
                const f32x2 inv2 = pk2(inv, inv);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int seg = (j + rot) & 3;
                    uint32_t h[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        float lo, hi;
                        upk2(mul2(f[j][k], inv2), lo, hi);
                        h[k] = cvt_fp8x2<FMT_E4M3>(lo, hi);
                    }
                    *reinterpret_cast<uint2*>(s.out_row + r * MXT + blk * 32 + seg * 8) = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
                }
                s.sf_row[(r & 31) * 16 + (r >> 5) * 4 + blk] = static_cast<uint8_t>(e);
            }
        }
        if (do_col) {
            // thread: 2 adjacent columns x 32 rows = two 32x1 blocks
            const int cp = tid & 63, rb = tid >> 6;
            uint32_t pk[32];
#pragma unroll
            for (int r = 0; r < 32; r++) pk[r] = *reinterpret_cast<const uint32_t*>(in + (rb * 32 + r) * (MXT * 2) + cp * 4);
            float g0 = 1.f, g1 = 1.f;
            if (MODE == 1) {
                uint32_t gv = 0;
                if (col0 + cp * 2 < C) gv = __ldg(reinterpret_cast<const uint32_t*>(gamma + col0 + cp * 2));
                g0 = bf16lo(gv);
                g1 = bf16hi(gv);
            }
            float am0 = 0.f, am1 = 0.f;
            if (MODE == 1) {
                // normalised values are fp32: compute them once, keep them packed
                f32x2 y[32];
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    const float rs = s.rs[rb * 32 + r];
                    const float a = __fmul_rn(__fmul_rn(bf16lo(pk[r]), rs), g0), b = __fmul_rn(__fmul_rn(bf16hi(pk[r]), rs), g1);
                    am0 = fmaxf(am0, fabsf(a));
                    am1 = fmaxf(am1, fabsf(b));
                    y[r] = pk2(a, b);
                }
                const uint32_t e0 = e8m0_from_amax(am0), e1 = e8m0_from_amax(am1);
                const f32x2 inv2 = pk2(e8m0_inv_scale(e0), e8m0_inv_scale(e1));
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    float lo, hi;
                    upk2(mul2(y[r], inv2), lo, hi);
                    *reinterpret_cast<uint16_t*>(s.out_col + (rb * 32 + r) * MXT + cp * 2) = static_cast<uint16_t>(cvt_fp8x2<FMT_E4M3>(lo, hi));
                }
                const int c0 = cp * 2, c1 = cp * 2 + 1;
                s.sf_col[(c0 & 31) * 16 + (c0 >> 5) * 4 + rb] = static_cast<uint8_t>(e0);
                s.sf_col[(c1 & 31) * 16 + (c1 >> 5) * 4 + rb] = static_cast<uint8_t>(e1);
            } else {
                // plain bf16 input: |x| of both columns at once as u16x2 integers (non-negative bf16 order like their bit patterns)
                uint32_t mx = 0;
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    uint32_t t;
                    asm("max.u16x2 %0, %1, %2;" : "=r"(t) : "r"(mx), "r"(pk[r] & 0x7fff7fffu));
                    mx = t;
                }
                // NaN patterns (> 0x7f80) must not win: fmaxf in the oracle drops NaN.  Rare: redo that column the slow way.
                am0 = bf16lo(mx);
                am1 = bf16hi(mx);
                if (((mx & 0xffffu) > 0x7f80u) || ((mx >> 16) > 0x7f80u)) {
                    am0 = am1 = 0.f;
#pragma unroll
                    for (int r = 0; r < 32; r++) {
                        am0 = fmaxf(am0, fabsf(bf16lo(pk[r])));
                        am1 = fmaxf(am1, fabsf(bf16hi(pk[r])));
                    }
                }
                if constexpr (BG) {
                    float s0 = 0.f, s1 = 0.f;           // rows past R come zero-filled from TMA
#pragma unroll
                    for (int r = 0; r < 32; r++) {
                        s0 += bf16lo(pk[r]);
                        s1 += bf16hi(pk[r]);
                    }
                    s.colsum[rb][cp * 2] = s0;
                    s.colsum[rb][cp * 2 + 1] = s1;
                }
                const uint32_t e0 = e8m0_from_amax(am0), e1 = e8m0_from_amax(am1);
                const f32x2 inv2 = pk2(e8m0_inv_scale(e0), e8m0_inv_scale(e1));
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    float lo, hi;
                    upk2(mul2(pk2(bf16lo(pk[r]), bf16hi(pk[r])), inv2), lo, hi);
                    *reinterpret_cast<uint16_t*>(s.out_col + (rb * 32 + r) * MXT + cp * 2) = static_cast<uint16_t>(cvt_fp8x2<FMT_E4M3>(lo, hi));
                }
                const int c0 = cp * 2, c1 = cp * 2 + 1;
                s.sf_col[(c0 & 31) * 16 + (c0 >> 5) * 4 + rb] = static_cast<uint8_t>(e0);
                s.sf_col[(c1 & 31) * 16 + (c1 >> 5) * 4 + rb] = static_cast<uint8_t>(e1);
            }
        }
        ptx::fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            if (do_row) ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
            if (do_col) ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
            ptx::tma_store_commit();
        }
        if (do_row && tid >= 32 && tid < 64) {
            const int64_t off = (static_cast<int64_t>(tr) * tiles_c + tc) * 512;
            reinterpret_cast<uint4*>(sf_row + off)[tid - 32] = reinterpret_cast<const uint4*>(s.sf_row)[tid - 32];
        }
        if (do_col && tid >= 64 && tid < 96) {
            const int64_t off = (static_cast<int64_t>(tc) * tiles_r + tr) * 512;
            reinterpret_cast<uint4*>(sf_col + off)[tid - 64] = reinterpret_cast<const uint4*>(s.sf_col)[tid - 64];
        }
        if constexpr (BG) {
            if (tid >= 128 && col0 + (tid - 128) < C) {
                const int c = tid - 128;
                atomicAdd(bgrad_ws + col0 + c, (s.colsum[0][c] + s.colsum[1][c]) + (s.colsum[2][c] + s.colsum[3][c]));
            }
        }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// sf_row: u8 [ceil(R/128)*128 * ceil(C/128)*4]; sf_col: u8 [ceil(C/128)*128 * ceil(R/128)*4]
// ldx: row pitch of x in elements (>= C, multiple of 8): a column slice of a wider matrix is quantised in place, without a copy
int launch_f32_to_bf16(const float* in, void* out, int64_t n, cudaStream_t stream);

// bgrad_ws (fp32 [C]) / bgrad_bf16 ([C]) together or both null: column sums of x (the bias gradient when x is a Linear's dY); needs the
// colwise copy (the sums ride in its pass) and the plain mode.
int launch_mx_quantize(const void* x, int64_t ldx, int64_t R, int64_t C, void* q_row, void* sf_row, void* q_col, void* sf_col, const float* rsigma,
                       const void* gamma, float* bgrad_ws, void* bgrad_bf16, cudaStream_t stream) {
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    LLMFP8_REQUIRE(do_row || do_col, "mx_quantize: nothing to produce");
    LLMFP8_REQUIRE(!do_row || (C % 32 == 0 && sf_row), "mx_quantize: rowwise needs cols %% 32 == 0 (got %lld) and a scale buffer", (long long)C);
    LLMFP8_REQUIRE(!do_col || (R % 32 == 0 && sf_col), "mx_quantize: colwise needs rows %% 32 == 0 (got %lld) and a scale buffer", (long long)R);
    LLMFP8_REQUIRE(C % 16 == 0, "mx_quantize: cols %lld must be a multiple of 16", (long long)C);
    LLMFP8_REQUIRE(ldx >= C && ldx % 8 == 0 && reinterpret_cast<uintptr_t>(x) % 16 == 0, "mx_quantize: row pitch %lld must be >= cols, a multiple of 8, x 16 B aligned", (long long)ldx);
    LLMFP8_REQUIRE((rsigma == nullptr) == (gamma == nullptr), "mx_quantize: rsigma and gamma go together");
    LLMFP8_REQUIRE(R < (1ll << 31) && C < (1ll << 31), "mx_quantize: dims too large");
    LLMFP8_REQUIRE((bgrad_ws == nullptr) == (bgrad_bf16 == nullptr), "mx_quantize: bgrad workspace and output go together");
    LLMFP8_REQUIRE(!bgrad_ws || (do_col && !rsigma), "mx_quantize: the bias gradient rides in the colwise pass of the plain mode");
    if (R == 0 || C == 0) return 0;
    if (bgrad_ws) {
        cudaError_t e0 = cudaMemsetAsync(bgrad_ws, 0, sizeof(float) * C, stream);
        if (e0 != cudaSuccess) { set_error("mx_quantize: memset: %s", cudaGetErrorString(e0)); return (int)e0; }
    }
    CUtensorMap m_in, m_row, m_col;
    int rc = make_tensor_map_2d(&m_in, x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, C, R, ldx * 2, MXT, MXT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, C, R, C, MXT, MXT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, C, R, C, MXT, MXT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int ntiles = static_cast<int>(((R + MXT - 1) / MXT) * ((C + MXT - 1) / MXT));
    const int smem = sizeof(MxSmem) + 128;
    const int cap = sm_count() * 2;
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e;
    if (rsigma) {
        e = cudaFuncSetAttribute(mx_quantize_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("mx_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
        LLMFP8_LAUNCH((mx_quantize_kernel<1>), grid, 256, smem, stream, m_in, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, rsigma, (const uint16_t*)gamma,
                                                           (int)R, (int)C, do_row, do_col, nullptr);
    } else if (bgrad_ws) {
        e = cudaFuncSetAttribute(mx_quantize_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("mx_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
        LLMFP8_LAUNCH((mx_quantize_kernel<0, true>), grid, 256, smem, stream, m_in, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, nullptr, nullptr, (int)R, (int)C,
                                                                 do_row, do_col, bgrad_ws);
        LLMFP8_LAUNCH_CHECK("mx_quantize");
        return launch_f32_to_bf16(bgrad_ws, bgrad_bf16, C, stream);
    } else {
        e = cudaFuncSetAttribute(mx_quantize_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) { set_error("mx_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
        LLMFP8_LAUNCH((mx_quantize_kernel<0>), grid, 256, smem, stream, m_in, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, nullptr, nullptr, (int)R, (int)C,
                                                           do_row, do_col, nullptr);
    }
    LLMFP8_LAUNCH_CHECK("mx_quantize");
    return 0;
}

}  // namespace llmfp8
