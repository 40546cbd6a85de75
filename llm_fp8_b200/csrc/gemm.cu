// FP8 GEMMs for sm_100a: TMA -> shared-memory ring -> tcgen05.mma (accumulators in TMEM) -> fused epilogue.
//   D[M,N] = alpha * sum_k A[m,k] * B[n,k]  (+ bias[n]) (+ D)      A, B fp8 (e4m3 | e5m2 each), fp32 accumulate
//   per-tensor scaling : kind::f8f6f4,  alpha = a_scale_inv * b_scale_inv read from device memory   (G1/G2/G3)
//   MXFP8              : kind::mxf8f6f4.block_scale, E8M0 scale per 32 along k, swizzled scale tiles staged
//                        global -> smem (bulk copy) -> TMEM (tcgen05.cp)                              (G1m/G2m/G3m)
// Operand storage ("major"):  K-major  = k is the contiguous dimension of the stored matrix ([rows=m or n, k]),
//                             MN-major = m (or n) is contiguous ([k, m or n]) — lets dgrad/wgrad read the SAME fp8
//                             bytes the forward pass produced, with no transposed copies:
//   fprop  Y = X W^T      : A = X  [M,K]  K-major,   B = W  [N,K]  K-major
//   dgrad  dX = dY W      : A = dY [M,N]  K-major,   B = W  [N,K]  MN-major (contraction over N)
//   wgrad  dW = dY^T X    : A = dY [T,N]  MN-major,  B = X  [T,K]  MN-major (contraction over tokens T)
// This is what TE does through cuBLASLt (tex.generic_gemm) for every Linear reached from te_llama.py:45-63.
//
// CTA = 6 warps: warp 0 TMA producer, warp 1 MMA issuer (+TMEM owner), warps 2..5 epilogue (one TMEM lane quadrant each).
// Persistent over 128 x BN output tiles, grouped rasterisation so a wave's A/B panels stay L2-resident.
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

constexpr int BM = 128;          // tile rows  (UMMA M, cta_group::1)
constexpr int BK = 128;          // k elements per pipeline stage (= one 128 B swizzle row when K-major)
constexpr int UMMA_K = 32;       // k per tcgen05.mma for 8-bit operands
constexpr int GEMM_THREADS = 192;
constexpr int GROUP_M = 16;

struct GemmParams {
    void* d;                 // output, row-major [M, N], leading dimension ldd (elements)
    int64_t ldd;
    int M, N, K;
    const float* a_scale_inv;   // per-tensor only
    const float* b_scale_inv;
    const void* bias;           // bf16 [N] or null
    int accumulate;             // D += result
    int out_fp32;               // 0: bf16 output, 1: fp32 output
    const uint8_t* sfa;         // MX only: swizzled E8M0 tiles, [ceil(M/128)][ceil(K/128)][512]
    const uint8_t* sfb;         //          [ceil(N/128)][ceil(K/128)][512]
    uint32_t idesc;
};

template <int BN, int STAGES, bool MX>
struct GemmSmem {
    static constexpr int A_BYTES = BM * BK;
    static constexpr int B_BYTES = BN * BK;
    static constexpr int SF_BYTES = MX ? (512 + (BN / 128) * 512) : 0;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * (STAGE_BYTES + SF_BYTES);
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;   // + barriers + alignment slack
};

template <int BN, int STAGES, int ACC_STAGES, bool A_MN, bool B_MN, bool MX>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
fp8_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const GemmParams p) {
    using L = GemmSmem<BN, STAGES, MX>;
    static_assert(BN % 128 == 0 && BN <= 256, "BN is 128 or 256");
    constexpr int ACC_COLS = BN;
    constexpr int SFA_COL = ACC_STAGES * ACC_COLS;
    constexpr int SFB_COL = SFA_COL + 4;
    constexpr int TMEM_COLS = 512;
    static_assert(!MX || SFB_COL + BN / 32 <= TMEM_COLS, "TMEM overflow");

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sf_base = smem + STAGES * L::STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint64_t* tmem_empty = tmem_full + ACC_STAGES;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_empty + ACC_STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_m = (p.M + BM - 1) / BM, num_n = (p.N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_kb = (p.K + BK - 1) / BK;
    const int kb_tiles = num_kb;   // scale tiles along k (one 512 B tile per 128 k)
    const int n_sf_tiles = (p.N + 127) / 128;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_b);
        for (int s = 0; s < STAGES; s++) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < ACC_STAGES; s++) {
            ptx::mbar_init(&tmem_full[s], 1);
            ptx::mbar_init(&tmem_empty[s], 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) {
        ptx::tmem_alloc(tmem_ptr_smem, TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    auto tile_coords = [&](int tile, int& m_blk, int& n_blk) {
        const int per_group = GROUP_M * num_n;
        const int group = tile / per_group;
        const int first_m = group * GROUP_M;
        const int gsize = (num_m - first_m) < GROUP_M ? (num_m - first_m) : GROUP_M;
        const int in_group = tile - group * per_group;
        m_blk = first_m + in_group % gsize;
        n_blk = in_group / gsize;
    };

    if (warp == 0) {
        // ================================ TMA producer
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int m_blk, n_blk;
                tile_coords(tile, m_blk, n_blk);
                const int m0 = m_blk * BM, n0 = n_blk * BN;
                for (int kb = 0; kb < num_kb; kb++) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * L::STAGE_BYTES;
                    uint8_t* sb = sa + L::A_BYTES;
                    // scale tiles past the last 128-row tile of B do not exist in memory: skip them (their columns are never stored)
                    int sfb_tiles = BN / 128;
                    if (MX && (n_blk * (BN / 128) + sfb_tiles) > n_sf_tiles) sfb_tiles = n_sf_tiles - n_blk * (BN / 128);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], L::STAGE_BYTES + (MX ? 512 + 512 * sfb_tiles : 0));
                    const int k0 = kb * BK;
                    if (!A_MN) ptx::tma_load_2d(&map_a, &full_bar[stage], sa, k0, m0);
                    else ptx::tma_load_2d(&map_a, &full_bar[stage], sa, m0, k0);
                    if (!B_MN) {
                        ptx::tma_load_2d(&map_b, &full_bar[stage], sb, k0, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 128; j++) ptx::tma_load_2d(&map_b, &full_bar[stage], sb + j * (128 * BK), n0 + j * 128, k0);
                    }
                    if (MX) {
                        uint8_t* ssf = sf_base + stage * L::SF_BYTES;
                        ptx::bulk_load_1d(ssf, p.sfa + (static_cast<int64_t>(m_blk) * kb_tiles + kb) * 512, 512, &full_bar[stage]);
#pragma unroll
                        for (int j = 0; j < BN / 128; j++)
                            if (j < sfb_tiles)
                                ptx::bulk_load_1d(ssf + 512 + j * 512,
                                                  p.sfb + (static_cast<int64_t>(n_blk * (BN / 128) + j) * kb_tiles + kb) * 512, 512, &full_bar[stage]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (one thread)
        if (lane == 0) {
            int stage = 0, phase = 0, acc = 0, acc_phase = 0;
            constexpr uint32_t A_LBO = A_MN ? (BK * 128) : 16, B_LBO = B_MN ? (BK * 128) : 16;
            constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K * 128) : UMMA_K, B_KSTEP = B_MN ? (UMMA_K * 128) : UMMA_K;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * ACC_COLS;
                for (int kb = 0; kb < num_kb; kb++) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem + stage * L::STAGE_BYTES);
                    const uint32_t sb = sa + L::A_BYTES;
                    if (MX) {
                        const uint32_t ssf = ptx::smem_u32(sf_base + stage * L::SF_BYTES);
                        ptx::tc_cp_32x128b_warpx4(tmem_base + SFA_COL, ptx::make_smem_desc(ssf, 0, 128, 0));
#pragma unroll
                        for (int j = 0; j < BN / 128; j++)
                            ptx::tc_cp_32x128b_warpx4(tmem_base + SFB_COL + 4 * j, ptx::make_smem_desc(ssf + 512 + j * 512, 0, 128, 0));
                    }
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; k++) {
                        const uint64_t da = ptx::make_smem_desc(sa + k * A_KSTEP, A_LBO, 1024, 2);
                        const uint64_t db = ptx::make_smem_desc(sb + k * B_KSTEP, B_LBO, 1024, 2);
                        const uint32_t accum = (kb | k) != 0;
                        if (MX) {
                            const uint32_t idesc = p.idesc | (static_cast<uint32_t>(k) << 4) | (static_cast<uint32_t>(k) << 29);
                            ptx::mma_mxf8f6f4(tmem_d, da, db, idesc, accum, tmem_base + SFA_COL, tmem_base + SFB_COL);
                        } else {
                            ptx::mma_f8f6f4(tmem_d, da, db, p.idesc, accum);
                        }
                    }
                    ptx::tc_commit(&empty_bar[stage]);          // smem slot reusable once these MMAs have read it
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::tc_commit(&tmem_full[acc]);                // accumulator complete -> epilogue
                if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ================================ epilogue: TMEM -> registers -> (alpha, bias, accumulate) -> global
        const int quad = warp & 3;
        int acc = 0, acc_phase = 0;
        float alpha = 1.0f;
        if (!MX) alpha = __fmul_rn(__ldg(p.a_scale_inv), __ldg(p.b_scale_inv));
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            int m_blk, n_blk;
            tile_coords(tile, m_blk, n_blk);
            const int row = m_blk * BM + quad * 32 + lane;
            const int n0 = n_blk * BN;
            ptx::mbar_wait(&tmem_full[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * ACC_COLS;
#pragma unroll 1
            for (int c = 0; c < BN / 32; c++) {
                uint32_t r[32];
                ptx::tmem_ld_32x32b_x32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                const int col = n0 + c * 32;
                if (row < p.M && col < p.N) {
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; j++) v[j] = __uint_as_float(r[j]) * alpha;
                    if (p.bias) {
                        const uint4* bp = reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.bias) + col);
#pragma unroll
                        for (int g = 0; g < 4; g++) {
                            if (col + g * 8 < p.N) {
                                float bf[8];
                                unpack8(__ldg(bp + g), bf);
#pragma unroll
                                for (int j = 0; j < 8; j++) v[g * 8 + j] += bf[j];
                            }
                        }
                    }
                    if (p.out_fp32) {
                        float* dp = static_cast<float*>(p.d) + static_cast<int64_t>(row) * p.ldd + col;
#pragma unroll
                        for (int g = 0; g < 8; g++) {
                            if (col + g * 4 < p.N) {
                                float4 o = make_float4(v[g * 4], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
                                if (p.accumulate) {
                                    const float4 old = *reinterpret_cast<const float4*>(dp + g * 4);
                                    o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                                }
                                *reinterpret_cast<float4*>(dp + g * 4) = o;
                            }
                        }
                    } else {
                        uint16_t* dp = static_cast<uint16_t*>(p.d) + static_cast<int64_t>(row) * p.ldd + col;
#pragma unroll
                        for (int g = 0; g < 4; g++) {
                            if (col + g * 8 < p.N) {
                                if (p.accumulate) {
                                    float of[8];
                                    unpack8(*reinterpret_cast<const uint4*>(dp + g * 8), of);
#pragma unroll
                                    for (int j = 0; j < 8; j++) v[g * 8 + j] += of[j];
                                }
                                uint4 o;
                                o.x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
                                o.y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
                                o.z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
                                o.w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
                                *reinterpret_cast<uint4*>(dp + g * 8) = o;
                            }
                        }
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty[acc]);
            if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------ host side
template <int BN, int STAGES, int ACC_STAGES, bool A_MN, bool B_MN, bool MX>
static int launch_variant(const CUtensorMap& ma, const CUtensorMap& mb, const GemmParams& p, int grid, cudaStream_t stream) {
    using L = GemmSmem<BN, STAGES, MX>;
    auto kern = fp8_gemm_kernel<BN, STAGES, ACC_STAGES, A_MN, B_MN, MX>;
    static bool configured = false;   // benign race: the attribute call is idempotent
    if (!configured) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL);
        if (e != cudaSuccess) { set_error("fp8_gemm: smem attribute (%d B): %s", L::TOTAL, cudaGetErrorString(e)); return (int)e; }
        configured = true;
    }
    kern<<<grid, GEMM_THREADS, L::TOTAL, stream>>>(ma, mb, p);
    LLMFP8_LAUNCH_CHECK("fp8_gemm");
    return 0;
}

// a: fp8 bytes. !a_mn: stored [M, K] (lda = row pitch in bytes >= K);  a_mn: stored [K, M] (lda >= M).  Same for b with N.
int launch_fp8_gemm(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd,
                    int out_fp32, int M, int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* sfa, const void* sfb,
                    const void* bias, int accumulate, int max_ctas, cudaStream_t stream) {
    const bool mx = sfa != nullptr || sfb != nullptr;
    LLMFP8_REQUIRE(M > 0 && N > 0 && K > 0, "fp8_gemm: empty problem M=%d N=%d K=%d", M, N, K);
    LLMFP8_REQUIRE((a_fmt == FMT_E4M3 || a_fmt == FMT_E5M2) && (b_fmt == FMT_E4M3 || b_fmt == FMT_E5M2), "fp8_gemm: bad operand formats");
    LLMFP8_REQUIRE(mx ? (sfa && sfb) : (a_scale_inv && b_scale_inv), "fp8_gemm: %s", mx ? "MX needs both scale tensors" : "per-tensor needs both scale_inv pointers");
    LLMFP8_REQUIRE(lda % 16 == 0 && ldb % 16 == 0, "fp8_gemm: operand pitches must be multiples of 16 B (lda=%lld ldb=%lld)", (long long)lda, (long long)ldb);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(d)) % 16 == 0, "fp8_gemm: pointers must be 16 B aligned");
    LLMFP8_REQUIRE(N % 8 == 0 && ldd % 8 == 0, "fp8_gemm: N and ldd must be multiples of 8 (N=%d ldd=%lld)", N, (long long)ldd);
    LLMFP8_REQUIRE(!mx || K % 32 == 0, "fp8_gemm: MXFP8 needs K %% 32 == 0 (K=%d)", K);
    LLMFP8_REQUIRE(!mx || (reinterpret_cast<uintptr_t>(sfa) | reinterpret_cast<uintptr_t>(sfb)) % 16 == 0, "fp8_gemm: scale tensors must be 16 B aligned");
    LLMFP8_REQUIRE(!(mx && (a_mn != b_mn) && a_mn), "fp8_gemm: operand layout combination (A MN-major, B K-major) is not built");

    CUtensorMap ma, mb;
    int rc;
    // K-major operand [rows, K]: inner dim = K, box = {128 k, tile rows}.   MN-major operand [K, rows]: inner = rows, box = {128 rows, 128 k}.
    constexpr int BN = 256;
    if (!a_mn) rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, K, M, lda, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, K, lda, 128, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn) rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, K, N, ldb, BK, BN, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, N, K, ldb, 128, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;

    GemmParams p;
    p.d = d; p.ldd = ldd; p.M = M; p.N = N; p.K = K;
    p.a_scale_inv = a_scale_inv; p.b_scale_inv = b_scale_inv; p.bias = bias; p.accumulate = accumulate; p.out_fp32 = out_fp32;
    p.sfa = static_cast<const uint8_t*>(sfa); p.sfb = static_cast<const uint8_t*>(sfb);
    p.idesc = mx ? ptx::make_idesc_mxf8(a_fmt, b_fmt, a_mn, b_mn, BM, BN, 0, 0) : ptx::make_idesc_f8(a_fmt, b_fmt, a_mn, b_mn, BM, BN);

    const int num_tiles = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
    int grid = sm_count();
    if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
    if (num_tiles < grid) grid = num_tiles;

    constexpr int ST = 4;
    if (!mx) {
        if (!a_mn && !b_mn) return launch_variant<BN, ST, 2, false, false, false>(ma, mb, p, grid, stream);
        if (!a_mn && b_mn) return launch_variant<BN, ST, 2, false, true, false>(ma, mb, p, grid, stream);
        if (a_mn && b_mn) return launch_variant<BN, ST, 2, true, true, false>(ma, mb, p, grid, stream);
        return launch_variant<BN, ST, 2, true, false, false>(ma, mb, p, grid, stream);
    }
    if (!a_mn && !b_mn) return launch_variant<BN, ST, 1, false, false, true>(ma, mb, p, grid, stream);
    if (!a_mn && b_mn) return launch_variant<BN, ST, 1, false, true, true>(ma, mb, p, grid, stream);
    return launch_variant<BN, ST, 1, true, true, true>(ma, mb, p, grid, stream);
}

}  // namespace llmfp8
