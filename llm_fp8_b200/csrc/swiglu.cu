// SwiGLU kernels for sm_100a (per-tensor scaling).
//   K3f swiglu_quant_fwd   : fc1_out bf16 [M,2I] (gate | up) -> fp8 [M,I], amax                 (5*M*I bytes)
//   K3b dswiglu_quant_bwd  : da bf16 [M,I], fc1_out bf16 [M,2I] -> fp8 [M,2I] (dgate | dup), amax, dbias[2I]  (8*M*I bytes)
// Replaces TE's swiglu / dswiglu (+cast, +dbias) inside LayerNormMLP(activation="swiglu") (te_llama.py:58-63);
// gate rows come first (te_llama.py:225-233).  exp/sigmoid are the pinned oracle versions (oracle_c.c: orc_expf).
#include <cstdlib>
#include "common.cuh"

namespace llmfp8 {

constexpr int FMT_BF16 = 2;   // unquantised bf16 output (feeds the MXFP8 block quantiser)

// 8 packed results -> 8 fp8 bytes (scale multiply separately rounded, as the oracle does) / amax / bf16
template <int FMT>
__device__ __forceinline__ uint2 quant8_p(const f32x2 (&a)[4], f32x2 scale2) {
    uint32_t h[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float lo, hi;
        upk2(mul2(a[j], scale2), lo, hi);
        h[j] = cvt_fp8x2<FMT>(lo, hi);
    }
    return make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
}
__device__ __forceinline__ float amax8_p(const f32x2 (&a)[4], float m) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float lo, hi;
        upk2(a[j], lo, hi);
        m = fmaxf(m, fmaxf(fabsf(lo), fabsf(hi)));
    }
    return m;
}
__device__ __forceinline__ uint4 bf16x8_p(const f32x2 (&a)[4]) {
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float lo, hi;
        upk2(a[j], lo, hi);
        o[j] = pack_bf16x2(lo, hi);
    }
    return make_uint4(o[0], o[1], o[2], o[3]);
}

// One thread: UNROLL units of (8 gate + 8 up values -> 8 outputs), all loads of the batch issued before the first use.
template <int FMT, int UNROLL>
__global__ void __launch_bounds__(256) swiglu_quant_kernel(const uint4* __restrict__ fc1, uint2* __restrict__ q, int64_t M, int64_t ivec /* I/8 */,
                                                           const float* __restrict__ scale_ptr, float* __restrict__ amax) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const f32x2 scale2 = pk2(scale, scale);
    const uint32_t total = static_cast<uint32_t>(M * ivec), iv = static_cast<uint32_t>(ivec);    // host guarantees < 2^31
    const uint32_t base = blockIdx.x * (256u * UNROLL) + threadIdx.x;
    float m = 0.f;
    uint4 vg[UNROLL], vu[UNROLL];
    {
        uint32_t row = base / iv, c = base - row * iv;       // one division per thread; later units advance (row, c) by 256 columns
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (base + u * 256u < total) {
                const uint4* p = fc1 + static_cast<int64_t>(row) * 2 * iv + c;
                vg[u] = ldg_stream(p);
                vu[u] = ldg_stream(p + iv);
            }
            c += 256u;
            while (c >= iv) { c -= iv; row++; }
        }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
        const uint32_t i = base + u * 256u;
        if (i < total) {
            float g[8], up[8];
            unpack8(vg[u], g);
            unpack8(vu[u], up);
            f32x2 a[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const f32x2 sig = sigmoid2_pinned(g[2 * j], g[2 * j + 1]);
                a[j] = mul2(mul2(pk2(g[2 * j], g[2 * j + 1]), sig), pk2(up[2 * j], up[2 * j + 1]));
            }
            if constexpr (FMT == FMT_BF16) {
                reinterpret_cast<uint4*>(q)[i] = bf16x8_p(a);
            } else {
                m = amax8_p(a, m);
                q[i] = quant8_p<FMT>(a, scale2);
            }
        }
    }
    if (FMT != FMT_BF16 && amax) block_amax_to_global(m, red, amax);
}

int launch_swiglu_quant_fwd(const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax, cudaStream_t stream) {
    LLMFP8_REQUIRE(I % 8 == 0, "swiglu_quant_fwd: ffn size %lld must be a multiple of 8", (long long)I);
    LLMFP8_REQUIRE(reinterpret_cast<uintptr_t>(fc1_out) % 16 == 0 && reinterpret_cast<uintptr_t>(q) % 8 == 0, "swiglu_quant_fwd: misaligned pointers");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2 || fmt == FMT_BF16, "swiglu_quant_fwd: bad output format %d", fmt);
    LLMFP8_REQUIRE(fmt != FMT_BF16 || reinterpret_cast<uintptr_t>(q) % 16 == 0, "swiglu_quant_fwd: misaligned bf16 output");
    if (M == 0 || I == 0) return 0;
    constexpr int UNROLL = 4;
    const int64_t total = M * (I / 8);
    const int64_t want = (total + 256 * UNROLL - 1) / (256 * UNROLL);
    LLMFP8_REQUIRE(total < (int64_t(1) << 31), "swiglu_quant_fwd: problem too large (%lld x %lld)", (long long)M, (long long)I);
    const int grid = static_cast<int>(want);
    if (fmt == FMT_E4M3)
        LLMFP8_LAUNCH((swiglu_quant_kernel<FMT_E4M3, UNROLL>), grid, 256, 0, stream, (const uint4*)fc1_out, (uint2*)q, M, I / 8, scale, amax);
    else if (fmt == FMT_E5M2)
        LLMFP8_LAUNCH((swiglu_quant_kernel<FMT_E5M2, UNROLL>), grid, 256, 0, stream, (const uint4*)fc1_out, (uint2*)q, M, I / 8, scale, amax);
    else
        LLMFP8_LAUNCH((swiglu_quant_kernel<FMT_BF16, UNROLL>), grid, 256, 0, stream, (const uint4*)fc1_out, (uint2*)q, M, I / 8, nullptr, nullptr);
    LLMFP8_LAUNCH_CHECK("swiglu_quant_fwd");
    return 0;
}

// Backward.  CTA = 32 column-vectors x 8 row lanes over a 256-column strip of I (so 256 dgate + 256 dup columns).
constexpr int DS_CV = 32, DS_RL = 8, DS_COLS = DS_CV * 8;
template <int FMT, int UN, int MINB>
__global__ void __launch_bounds__(256, MINB) dswiglu_quant_kernel(const uint4* __restrict__ da, const uint4* __restrict__ fc1, uint2* __restrict__ q,
                                                            int64_t M, int64_t I, int64_t rows_per_cta, const float* __restrict__ scale_ptr,
                                                            float* __restrict__ amax, float* __restrict__ dbias_ws /* [2I] or null */) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[8];
    __shared__ float part[DS_RL][2 * DS_COLS + 8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const int cv = threadIdx.x % DS_CV, rl = threadIdx.x / DS_CV;
    const int64_t col0 = static_cast<int64_t>(blockIdx.x) * DS_COLS + cv * 8;
    const int64_t ivec = I / 8;
    const int64_t r_begin = static_cast<int64_t>(blockIdx.y) * rows_per_cta;
    const int64_t r_end = r_begin + rows_per_cta < M ? r_begin + rows_per_cta : M;
    float m = 0.f;
    const f32x2 one2 = pk2(1.0f, 1.0f), scale2 = pk2(scale, scale);
    f32x2 sg[4], su[4];
#pragma unroll
    for (int j = 0; j < 4; j++) sg[j] = su[j] = pk2(0.f, 0.f);
    if (col0 < I) {
        const uint4* pda = da + r_begin * ivec + col0 / 8;
        const uint4* pfc = fc1 + r_begin * 2 * ivec + col0 / 8;
        uint2* pq = q + r_begin * 2 * ivec + col0 / 8;
        const int nrows = static_cast<int>(r_end - r_begin);
        for (int r0 = rl; r0 < nrows; r0 += DS_RL * UN) {
            uint4 vda[UN], vg[UN], vu[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int r = r0 + u * DS_RL;
                if (r < nrows) {
                    vda[u] = ldg_stream(pda + r * ivec);
                    vg[u] = ldg_stream(pfc + r * 2 * ivec);
                    vu[u] = ldg_stream(pfc + r * 2 * ivec + ivec);
                }
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int r = r0 + u * DS_RL;
                if (r < nrows) {
                    float dy[8], g[8], up[8];
                    unpack8(vda[u], dy);
                    unpack8(vg[u], g);
                    unpack8(vu[u], up);
                    f32x2 dg[4], du[4];
#pragma unroll
                    for (int j = 0; j < 4; j++) {
                        const f32x2 g2 = pk2(g[2 * j], g[2 * j + 1]), u2 = pk2(up[2 * j], up[2 * j + 1]), dy2 = pk2(dy[2 * j], dy[2 * j + 1]);
                        const f32x2 sig = sigmoid2_pinned(g[2 * j], g[2 * j + 1]);
                        const f32x2 tt = fma2(g2, sub2(one2, sig), one2);
                        const f32x2 dsilu = mul2(sig, tt);
                        dg[j] = mul2(mul2(dy2, u2), dsilu);
                        du[j] = mul2(dy2, mul2(g2, sig));
                        sg[j] = add2(dg[j], sg[j]);
                        su[j] = add2(du[j], su[j]);
                    }
                    if constexpr (FMT == FMT_BF16) {
                        reinterpret_cast<uint4*>(q)[(r_begin + r) * 2 * ivec + col0 / 8] = bf16x8_p(dg);
                        reinterpret_cast<uint4*>(q)[(r_begin + r) * 2 * ivec + ivec + col0 / 8] = bf16x8_p(du);
                    } else {
                        m = amax8_p(dg, m);
                        m = amax8_p(du, m);
                        pq[r * 2 * ivec] = quant8_p<FMT>(dg, scale2);
                        pq[r * 2 * ivec + ivec] = quant8_p<FMT>(du, scale2);
                    }
                }
            }
        }
    }
    if (dbias_ws) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            float lo, hi;
            upk2(sg[j], lo, hi);
            part[rl][cv * 8 + 2 * j] = lo;
            part[rl][cv * 8 + 2 * j + 1] = hi;
            upk2(su[j], lo, hi);
            part[rl][DS_COLS + cv * 8 + 2 * j] = lo;
            part[rl][DS_COLS + cv * 8 + 2 * j + 1] = hi;
        }
    }
    if (FMT != FMT_BF16 && amax) block_amax_to_global(m, red, amax);
    else __syncthreads();
    if (dbias_ws) {
        const int c = threadIdx.x;
        float s0 = 0.f, s1 = 0.f;
#pragma unroll
        for (int j = 0; j < DS_RL; j++) { s0 += part[j][c]; s1 += part[j][DS_COLS + c]; }
        const int64_t col = static_cast<int64_t>(blockIdx.x) * DS_COLS + c;
        if (col < I) {
            atomicAdd(dbias_ws + col, s0);
            atomicAdd(dbias_ws + I + col, s1);
        }
    }
}

int launch_f32_to_bf16(const float* in, void* out, int64_t n, cudaStream_t stream);

// dbias_ws: fp32 [2I] workspace, dbias_bf16: [2I] output; both may be null (no bias grad wanted)
int launch_dswiglu_quant_bwd(const void* da, const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax,
                             float* dbias_ws, void* dbias_bf16, cudaStream_t stream) {
    LLMFP8_REQUIRE(I % 8 == 0, "dswiglu_quant_bwd: ffn size %lld must be a multiple of 8", (long long)I);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(da) | reinterpret_cast<uintptr_t>(fc1_out)) % 16 == 0 && reinterpret_cast<uintptr_t>(q) % 8 == 0,
                   "dswiglu_quant_bwd: misaligned pointers");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2 || fmt == FMT_BF16, "dswiglu_quant_bwd: bad output format %d", fmt);
    LLMFP8_REQUIRE(fmt != FMT_BF16 || reinterpret_cast<uintptr_t>(q) % 16 == 0, "dswiglu_quant_bwd: misaligned bf16 output");
    LLMFP8_REQUIRE((dbias_ws == nullptr) == (dbias_bf16 == nullptr), "dswiglu_quant_bwd: dbias workspace and output go together");
    if (M == 0 || I == 0) return 0;
    if (dbias_ws) {
        cudaError_t e = cudaMemsetAsync(dbias_ws, 0, sizeof(float) * 2 * I, stream);
        if (e != cudaSuccess) { set_error("dswiglu_quant_bwd: memset: %s", cudaGetErrorString(e)); return (int)e; }
    }
    const int strips = static_cast<int>((I + DS_COLS - 1) / DS_COLS);
    // many short CTAs (64 rows x 256 columns) rather than ~1.4 waves of long ones: the tail of the last wave stays a few percent
    // (the long-CTA version measured 67 % of HBM bandwidth with 616 CTAs on 444 resident slots)
    int64_t rows_per_cta = 64;
    while ((M + rows_per_cta - 1) / rows_per_cta > 65535) rows_per_cta *= 2;
    const int64_t ychunks = (M + rows_per_cta - 1) / rows_per_cta;
    dim3 grid(strips, static_cast<unsigned>(ychunks));
    static const int variant = [] { const char* e = getenv("LLMFP8_TUNE_DSWIGLU"); return e ? atoi(e) : 0; }();
#define DSW_LAUNCH(UN, MINB)                                                                                                                  \
    do {                                                                                                                                      \
        if (fmt == FMT_E4M3)                                                                                                                  \
            LLMFP8_LAUNCH((dswiglu_quant_kernel<FMT_E4M3, UN, MINB>), grid, 256, 0, stream, (const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, scale, amax, dbias_ws); \
        else if (fmt == FMT_E5M2)                                                                                                             \
            LLMFP8_LAUNCH((dswiglu_quant_kernel<FMT_E5M2, UN, MINB>), grid, 256, 0, stream, (const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, scale, amax, dbias_ws); \
        else                                                                                                                                  \
            LLMFP8_LAUNCH((dswiglu_quant_kernel<FMT_BF16, UN, MINB>), grid, 256, 0, stream, (const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, nullptr, nullptr, dbias_ws); \
    } while (0)
    if (variant == 1) DSW_LAUNCH(2, 4);
    else if (variant == 2) DSW_LAUNCH(1, 4);
    else if (variant == 3) DSW_LAUNCH(4, 2);
    else DSW_LAUNCH(2, 3);
#undef DSW_LAUNCH
    LLMFP8_LAUNCH_CHECK("dswiglu_quant_bwd");
    if (dbias_ws) return launch_f32_to_bf16(dbias_ws, dbias_bf16, 2 * I, stream);
    return 0;
}

}  // namespace llmfp8
