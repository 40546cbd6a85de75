// SwiGLU kernels for sm_100a (per-tensor scaling).
//   K3f swiglu_quant_fwd   : fc1_out bf16 [M,2I] (gate | up) -> fp8 [M,I], amax                 (5*M*I bytes)
//   K3b dswiglu_quant_bwd  : da bf16 [M,I], fc1_out bf16 [M,2I] -> fp8 [M,2I] (dgate | dup), amax, dbias[2I]  (8*M*I bytes)
// Replaces TE's swiglu / dswiglu (+cast, +dbias) inside LayerNormMLP(activation="swiglu") (te_llama.py:58-63);
// gate rows come first (te_llama.py:225-233).  exp/sigmoid are the pinned oracle versions (oracle_c.c: orc_expf).
#include "common.cuh"

namespace llmfp8 {

constexpr int FMT_BF16 = 2;   // unquantised bf16 output (feeds the MXFP8 block quantiser)

// One thread: 8 gate + 8 up values -> 8 fp8 bytes.
template <int FMT, int UNROLL>
__global__ void __launch_bounds__(256) swiglu_quant_kernel(const uint4* __restrict__ fc1, uint2* __restrict__ q, int64_t M, int64_t ivec /* I/8 */,
                                                           const float* __restrict__ scale_ptr, float* __restrict__ amax) {
    __shared__ float red[8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const int64_t total = M * ivec;
    const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
    float m = 0.f;
    for (int64_t i0 = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < total; i0 += stride * UNROLL) {
        uint4 vg[UNROLL], vu[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int64_t i = i0 + u * stride;
            if (i < total) {
                const int64_t row = i / ivec, c = i - row * ivec;
                vg[u] = ldg_stream(fc1 + row * 2 * ivec + c);
                vu[u] = ldg_stream(fc1 + row * 2 * ivec + ivec + c);
            }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            const int64_t i = i0 + u * stride;
            if (i < total) {
                float g[8], up[8], a[8];
                unpack8(vg[u], g);
                unpack8(vu[u], up);
#pragma unroll
                for (int j = 0; j < 8; j++) a[j] = __fmul_rn(__fmul_rn(g[j], sigmoid_pinned(g[j])), up[j]);
                if constexpr (FMT == FMT_BF16) {
                    uint4 o;
                    o.x = pack_bf16x2(a[0], a[1]); o.y = pack_bf16x2(a[2], a[3]);
                    o.z = pack_bf16x2(a[4], a[5]); o.w = pack_bf16x2(a[6], a[7]);
                    reinterpret_cast<uint4*>(q)[i] = o;
                } else {
                    m = amax8(a, m);
                    q[i] = quant8<FMT>(a, scale);
                }
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
    constexpr int UNROLL = 2;
    const int64_t total = M * (I / 8);
    int64_t want = (total + 256 * UNROLL - 1) / (256 * UNROLL);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(want < cap ? want : cap);
    if (fmt == FMT_E4M3)
        swiglu_quant_kernel<FMT_E4M3, UNROLL><<<grid, 256, 0, stream>>>((const uint4*)fc1_out, (uint2*)q, M, I / 8, scale, amax);
    else if (fmt == FMT_E5M2)
        swiglu_quant_kernel<FMT_E5M2, UNROLL><<<grid, 256, 0, stream>>>((const uint4*)fc1_out, (uint2*)q, M, I / 8, scale, amax);
    else
        swiglu_quant_kernel<FMT_BF16, UNROLL><<<grid, 256, 0, stream>>>((const uint4*)fc1_out, (uint2*)q, M, I / 8, nullptr, nullptr);
    LLMFP8_LAUNCH_CHECK("swiglu_quant_fwd");
    return 0;
}

// Backward.  CTA = 32 column-vectors x 8 row lanes over a 256-column strip of I (so 256 dgate + 256 dup columns).
constexpr int DS_CV = 32, DS_RL = 8, DS_COLS = DS_CV * 8;
template <int FMT>
__global__ void __launch_bounds__(256) dswiglu_quant_kernel(const uint4* __restrict__ da, const uint4* __restrict__ fc1, uint2* __restrict__ q,
                                                            int64_t M, int64_t I, int64_t rows_per_cta, const float* __restrict__ scale_ptr,
                                                            float* __restrict__ amax, float* __restrict__ dbias_ws /* [2I] or null */) {
    __shared__ float red[8];
    __shared__ float part[DS_RL][2 * DS_COLS + 8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const int cv = threadIdx.x % DS_CV, rl = threadIdx.x / DS_CV;
    const int64_t col0 = static_cast<int64_t>(blockIdx.x) * DS_COLS + cv * 8;
    const int64_t ivec = I / 8;
    const int64_t r_begin = static_cast<int64_t>(blockIdx.y) * rows_per_cta;
    const int64_t r_end = r_begin + rows_per_cta < M ? r_begin + rows_per_cta : M;
    float m = 0.f;
    float sg[8], su[8];
#pragma unroll
    for (int j = 0; j < 8; j++) sg[j] = su[j] = 0.f;
    if (col0 < I) {
        constexpr int UN = 2;
        for (int64_t r0 = r_begin + rl; r0 < r_end; r0 += DS_RL * UN) {
            uint4 vda[UN], vg[UN], vu[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int64_t r = r0 + u * DS_RL;
                if (r < r_end) {
                    vda[u] = ldg_stream(da + r * ivec + col0 / 8);
                    vg[u] = ldg_stream(fc1 + r * 2 * ivec + col0 / 8);
                    vu[u] = ldg_stream(fc1 + r * 2 * ivec + ivec + col0 / 8);
                }
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int64_t r = r0 + u * DS_RL;
                if (r < r_end) {
                    float dy[8], g[8], up[8], dg[8], du[8];
                    unpack8(vda[u], dy);
                    unpack8(vg[u], g);
                    unpack8(vu[u], up);
#pragma unroll
                    for (int j = 0; j < 8; j++) {
                        const float sig = sigmoid_pinned(g[j]);
                        const float tt = __fmaf_rn(g[j], __fsub_rn(1.0f, sig), 1.0f);
                        const float dsilu = __fmul_rn(sig, tt);
                        dg[j] = __fmul_rn(__fmul_rn(dy[j], up[j]), dsilu);
                        du[j] = __fmul_rn(dy[j], __fmul_rn(g[j], sig));
                        sg[j] += dg[j];
                        su[j] += du[j];
                    }
                    if constexpr (FMT == FMT_BF16) {
                        uint4 o;
                        o.x = pack_bf16x2(dg[0], dg[1]); o.y = pack_bf16x2(dg[2], dg[3]);
                        o.z = pack_bf16x2(dg[4], dg[5]); o.w = pack_bf16x2(dg[6], dg[7]);
                        reinterpret_cast<uint4*>(q)[r * 2 * ivec + col0 / 8] = o;
                        o.x = pack_bf16x2(du[0], du[1]); o.y = pack_bf16x2(du[2], du[3]);
                        o.z = pack_bf16x2(du[4], du[5]); o.w = pack_bf16x2(du[6], du[7]);
                        reinterpret_cast<uint4*>(q)[r * 2 * ivec + ivec + col0 / 8] = o;
                    } else {
                        m = amax8(dg, m);
                        m = amax8(du, m);
                        q[r * 2 * ivec + col0 / 8] = quant8<FMT>(dg, scale);
                        q[r * 2 * ivec + ivec + col0 / 8] = quant8<FMT>(du, scale);
                    }
                }
            }
        }
    }
    if (dbias_ws) {
#pragma unroll
        for (int j = 0; j < 8; j++) {
            part[rl][cv * 8 + j] = sg[j];
            part[rl][DS_COLS + cv * 8 + j] = su[j];
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
    int64_t ychunks = (static_cast<int64_t>(sm_count()) * 4 + strips - 1) / strips;
    int64_t rows_per_cta = (M + ychunks - 1) / ychunks;
    rows_per_cta = (rows_per_cta + 15) / 16 * 16;
    ychunks = (M + rows_per_cta - 1) / rows_per_cta;
    dim3 grid(strips, static_cast<unsigned>(ychunks));
    if (fmt == FMT_E4M3)
        dswiglu_quant_kernel<FMT_E4M3><<<grid, 256, 0, stream>>>((const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, scale, amax, dbias_ws);
    else if (fmt == FMT_E5M2)
        dswiglu_quant_kernel<FMT_E5M2><<<grid, 256, 0, stream>>>((const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, scale, amax, dbias_ws);
    else
        dswiglu_quant_kernel<FMT_BF16><<<grid, 256, 0, stream>>>((const uint4*)da, (const uint4*)fc1_out, (uint2*)q, M, I, rows_per_cta, nullptr, nullptr, dbias_ws);
    LLMFP8_LAUNCH_CHECK("dswiglu_quant_bwd");
    if (dbias_ws) return launch_f32_to_bf16(dbias_ws, dbias_bf16, 2 * I, stream);
    return 0;
}

}  // namespace llmfp8
