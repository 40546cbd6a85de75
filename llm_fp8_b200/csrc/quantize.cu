// Per-tensor (delayed-scaling) quantise kernels for sm_100a.
//   K2  cast_amax        bf16 [n]   -> fp8 [n], amax                     (3 B/element of HBM traffic)
//   K2b cast_amax_bgrad  bf16 [R,C] -> fp8 [R,C], amax, column sums      (dY cast + bias grad in one pass)
//   K4  amax_scale_update  amax history roll + scale / scale_inv update  (one launch for all fp8 tensors)
// Replaces TE's cast/cast_transpose_dbias + amax_and_scale_update ops reached from te_llama.py:76-80.
// Arithmetic is the oracle's (oracle/oracle_c.c orc_quantize_bf16 / orc_amax_scale_update), bit for bit.
#include "common.cuh"

namespace llmfp8 {

// ------------------------------------------------------------------------------------------- K2
// Each thread owns UNROLL units of 8 elements (one 16 B load, one 8 B store), a warp's units contiguous in memory so every request
// covers whole 32 B sectors exactly once; all UNROLL loads are issued before the first use.
template <int FMT, int UNROLL>
__global__ void __launch_bounds__(256) cast_amax_kernel(const uint4* __restrict__ x, uint2* __restrict__ q, int64_t nunits,
                                                        const float* __restrict__ scale_ptr, float* __restrict__ amax) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const f32x2 scale2 = pk2(scale, scale);
    float m = 0.f;
    const int64_t base = static_cast<int64_t>(blockIdx.x) * (256 * UNROLL) + threadIdx.x;
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++)
        if (base + u * 256 < nunits) v[u] = ldg_stream(x + base + u * 256);
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
        if (base + u * 256 < nunits) {
            float f[8];
            unpack8(v[u], f);
            uint32_t h[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                m = fmaxf(m, fmaxf(fabsf(f[2 * k]), fabsf(f[2 * k + 1])));
                float lo, hi;
                upk2(mul2(pk2(f[2 * k], f[2 * k + 1]), scale2), lo, hi);
                h[k] = cvt_fp8x2<FMT>(lo, hi);
            }
            q[base + u * 256] = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
        }
    }
    if (amax) block_amax_to_global(m, red, amax);
}

int launch_cast_amax(const void* x, void* q, int64_t n, int fmt, const float* scale, float* amax, cudaStream_t stream) {
    LLMFP8_REQUIRE(n % 16 == 0, "cast_amax: element count %lld must be a multiple of 16", (long long)n);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) % 16 == 0, "cast_amax: pointers must be 16 B aligned");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2, "cast_amax: bad fp8 format %d", fmt);
    if (n == 0) return 0;
    const int64_t nunits = n / 8;
    constexpr int UNROLL = 8;
    const int64_t want = (nunits + 256 * UNROLL - 1) / (256 * UNROLL);
    LLMFP8_REQUIRE(want < (int64_t(1) << 31), "cast_amax: tensor too large (%lld elements)", (long long)n);
    const int grid = static_cast<int>(want);
    if (fmt == FMT_E4M3)
        LLMFP8_LAUNCH((cast_amax_kernel<FMT_E4M3, UNROLL>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), nunits, scale, amax);
    else
        LLMFP8_LAUNCH((cast_amax_kernel<FMT_E5M2, UNROLL>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), nunits, scale, amax);
    LLMFP8_LAUNCH_CHECK("cast_amax");
    return 0;
}

// K2 on a column slice: x is [R, C] with row pitch ldx elements (a view into a wider matrix), q is dense [R, C].
// Same unit mapping as the flat kernel over the DENSE index space; only the load address carries the pitch.
template <int FMT, int UNROLL>
__global__ void __launch_bounds__(256) cast_amax_2d_kernel(const uint4* __restrict__ x, uint2* __restrict__ q, uint32_t nunits, uint32_t cvec,
                                                           int64_t ldvec, const float* __restrict__ scale_ptr, float* __restrict__ amax) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const f32x2 scale2 = pk2(scale, scale);
    float m = 0.f;
    const uint32_t base = blockIdx.x * (256u * UNROLL) + threadIdx.x;
    uint4 v[UNROLL];
    {
        uint32_t row = base / cvec, c = base - row * cvec;
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (base + u * 256u < nunits) v[u] = ldg_stream(x + static_cast<int64_t>(row) * ldvec + c);
            c += 256u;
            while (c >= cvec) { c -= cvec; row++; }
        }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; u++) {
        if (base + u * 256u < nunits) {
            float f[8];
            unpack8(v[u], f);
            uint32_t h[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                m = fmaxf(m, fmaxf(fabsf(f[2 * k]), fabsf(f[2 * k + 1])));
                float lo, hi;
                upk2(mul2(pk2(f[2 * k], f[2 * k + 1]), scale2), lo, hi);
                h[k] = cvt_fp8x2<FMT>(lo, hi);
            }
            q[base + u * 256u] = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
        }
    }
    if (amax) block_amax_to_global(m, red, amax);
}

int launch_cast_amax_2d(const void* x, int64_t ldx, void* q, int64_t R, int64_t C, int fmt, const float* scale, float* amax, cudaStream_t stream) {
    if (ldx == C) return launch_cast_amax(x, q, R * C, fmt, scale, amax, stream);
    LLMFP8_REQUIRE(C % 16 == 0 && ldx % 8 == 0 && ldx > C, "cast_amax_2d: cols %lld must be a multiple of 16, pitch %lld a multiple of 8 and >= cols", (long long)C, (long long)ldx);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) % 16 == 0, "cast_amax_2d: pointers must be 16 B aligned");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2, "cast_amax_2d: bad fp8 format %d", fmt);
    if (R == 0 || C == 0) return 0;
    const int64_t nunits = R * (C / 8);
    LLMFP8_REQUIRE(nunits < (int64_t(1) << 31), "cast_amax_2d: tensor too large (%lld x %lld)", (long long)R, (long long)C);
    constexpr int UNROLL = 8;
    const int grid = static_cast<int>((nunits + 256 * UNROLL - 1) / (256 * UNROLL));
    if (fmt == FMT_E4M3)
        LLMFP8_LAUNCH((cast_amax_2d_kernel<FMT_E4M3, UNROLL>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), (uint32_t)nunits, (uint32_t)(C / 8), ldx / 8, scale, amax);
    else
        LLMFP8_LAUNCH((cast_amax_2d_kernel<FMT_E5M2, UNROLL>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), (uint32_t)nunits, (uint32_t)(C / 8), ldx / 8, scale, amax);
    LLMFP8_LAUNCH_CHECK("cast_amax_2d");
    return 0;
}

// ------------------------------------------------------------------------------------------- K2b
// CTA = 256 threads = 32 column-vectors (8 bf16 each: 256 columns, 512 B contiguous per row) x 8 row lanes.
// A CTA walks `rows_per_cta` rows of its 256-column strip, keeping 8 fp32 column sums per thread, then folds the
// 8 row lanes through shared memory and adds the strip's sums into the fp32 workspace with atomics.
constexpr int BG_CV = 32, BG_RL = 8, BG_COLS = BG_CV * 8;
template <int FMT>
__global__ void __launch_bounds__(256) cast_amax_bgrad_kernel(const uint4* __restrict__ x, uint2* __restrict__ q, int64_t R, int64_t C,
                                                              int64_t rows_per_cta, const float* __restrict__ scale_ptr,
                                                              float* __restrict__ amax, float* __restrict__ colsum_ws) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[8];
    __shared__ float part[BG_RL][BG_COLS + 8];
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const int cv = threadIdx.x % BG_CV, rl = threadIdx.x / BG_CV;
    const int64_t col0 = static_cast<int64_t>(blockIdx.x) * BG_COLS + cv * 8;
    const int64_t r_begin = static_cast<int64_t>(blockIdx.y) * rows_per_cta;
    const int64_t r_end = r_begin + rows_per_cta < R ? r_begin + rows_per_cta : R;
    const bool col_ok = col0 < C;
    const int64_t rowvec = C / 8;
    float m = 0.f;
    float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (col_ok) {
        constexpr int UN = 4;
        for (int64_t r0 = r_begin + rl; r0 < r_end; r0 += BG_RL * UN) {
            uint4 v[UN];
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int64_t r = r0 + u * BG_RL;
                if (r < r_end) v[u] = ldg_stream(x + r * rowvec + col0 / 8);
            }
#pragma unroll
            for (int u = 0; u < UN; u++) {
                const int64_t r = r0 + u * BG_RL;
                if (r < r_end) {
                    float f[8];
                    unpack8(v[u], f);
                    m = amax8(f, m);
#pragma unroll
                    for (int i = 0; i < 8; i++) cs[i] += f[i];
                    q[r * rowvec + col0 / 8] = quant8<FMT>(f, scale);
                }
            }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) part[rl][cv * 8 + i] = cs[i];
    if (amax) block_amax_to_global(m, red, amax);   // contains a __syncthreads
    else __syncthreads();
    {
        const int c = threadIdx.x;   // 256 threads <-> 256 columns
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < BG_RL; j++) s += part[j][c];
        const int64_t col = static_cast<int64_t>(blockIdx.x) * BG_COLS + c;
        if (col < C) atomicAdd(colsum_ws + col, s);
    }
}

__global__ void f32_to_bf16_kernel(const float* __restrict__ in, uint16_t* __restrict__ out, int64_t n) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) out[i] = static_cast<uint16_t>(pack_bf16x2(in[i], 0.f) & 0xffffu);
}

// colsum_ws: fp32 [C] workspace (zeroed here); bgrad_bf16: [C] output
int launch_cast_amax_bgrad(const void* x, void* q, int64_t R, int64_t C, int fmt, const float* scale, float* amax, float* colsum_ws,
                           void* bgrad_bf16, cudaStream_t stream) {
    LLMFP8_REQUIRE(C % 8 == 0, "cast_amax_bgrad: cols %lld must be a multiple of 8", (long long)C);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(x) % 16 == 0) && (reinterpret_cast<uintptr_t>(q) % 8 == 0), "cast_amax_bgrad: misaligned pointers");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2, "cast_amax_bgrad: bad fp8 format %d", fmt);
    LLMFP8_REQUIRE(colsum_ws && bgrad_bf16, "cast_amax_bgrad: workspace and output required");
    if (R == 0 || C == 0) return 0;
    cudaError_t e = cudaMemsetAsync(colsum_ws, 0, sizeof(float) * C, stream);
    if (e != cudaSuccess) { set_error("cast_amax_bgrad: memset: %s", cudaGetErrorString(e)); return (int)e; }
    const int strips = static_cast<int>((C + BG_COLS - 1) / BG_COLS);
    int64_t ychunks = (static_cast<int64_t>(sm_count()) * 6 + strips - 1) / strips;
    int64_t rows_per_cta = (R + ychunks - 1) / ychunks;
    rows_per_cta = (rows_per_cta + 31) / 32 * 32;
    ychunks = (R + rows_per_cta - 1) / rows_per_cta;
    dim3 grid(strips, static_cast<unsigned>(ychunks));
    if (fmt == FMT_E4M3)
        LLMFP8_LAUNCH((cast_amax_bgrad_kernel<FMT_E4M3>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), R, C, rows_per_cta, scale, amax, colsum_ws);
    else
        LLMFP8_LAUNCH((cast_amax_bgrad_kernel<FMT_E5M2>), grid, 256, 0, stream, static_cast<const uint4*>(x), static_cast<uint2*>(q), R, C, rows_per_cta, scale, amax, colsum_ws);
    LLMFP8_LAUNCH_CHECK("cast_amax_bgrad");
    LLMFP8_LAUNCH((f32_to_bf16_kernel), static_cast<unsigned>((C + 255) / 256), 256, 0, stream, colsum_ws, static_cast<uint16_t*>(bgrad_bf16), C);
    LLMFP8_LAUNCH_CHECK("cast_amax_bgrad.finalize");
    return 0;
}

int launch_f32_to_bf16(const float* in, void* out, int64_t n, cudaStream_t stream) {
    if (n == 0) return 0;
    LLMFP8_LAUNCH((f32_to_bf16_kernel), static_cast<unsigned>((n + 255) / 256), 256, 0, stream, in, static_cast<uint16_t*>(out), n);
    LLMFP8_LAUNCH_CHECK("f32_to_bf16");
    return 0;
}

// ------------------------------------------------------------------------------------------- K4
// One thread per fp8 tensor slot; hist is [L, T] row-major.  Oracle: orc_amax_scale_update.
__global__ void amax_scale_update_kernel(float* __restrict__ hist, float* __restrict__ scale, float* __restrict__ scale_inv,
                                         const float* __restrict__ fmax, int L, int T, float margin_div) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    float am = 0.f;
    bool nonfinite = false;
    const float first = hist[t];
    for (int l = 0; l < L; l++) {
        const float h = hist[l * T + t];
        nonfinite |= !isfinite(h);
        am = (h > am) ? h : am;
        if (l + 1 < L) hist[l * T + t] = hist[(l + 1) * T + t];   // roll(-1): row l takes row l+1
    }
    hist[(L - 1) * T + t] = first;
    hist[t] = 0.f;
    const float old = scale[t];
    float sf = __fdiv_rn(__fdiv_rn(fmax[t], am), margin_div);
    if (!(am > 0.f)) sf = old;
    if (nonfinite) sf = old;
    if (isinf(sf)) sf = 3.402823466e38f;
    scale[t] = sf;
    scale_inv[t] = __fdiv_rn(1.0f, sf);
}

int launch_amax_scale_update(float* hist, float* scale, float* scale_inv, const float* fmax, int L, int T, int margin, cudaStream_t stream) {
    LLMFP8_REQUIRE(L >= 1 && T >= 0, "amax_scale_update: bad sizes L=%d T=%d", L, T);
    LLMFP8_REQUIRE(margin >= 0 && margin < 120, "amax_scale_update: bad margin %d", margin);
    if (T == 0) return 0;
    LLMFP8_LAUNCH((amax_scale_update_kernel), (T + 127) / 128, 128, 0, stream, hist, scale, scale_inv, fmax, L, T, ldexpf(1.0f, margin));
    LLMFP8_LAUNCH_CHECK("amax_scale_update");
    return 0;
}

}  // namespace llmfp8
