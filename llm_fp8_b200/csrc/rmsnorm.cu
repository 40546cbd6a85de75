// RMSNorm kernels for sm_100a (row per 128-thread CTA, persistent over rows, 16 B loads).
//   K1   rmsnorm_quant_fwd : x bf16 [M,H], gamma -> fp8 [M,H] (per-tensor scale), rsigma[M], amax   (~3*M*H bytes)
//   K1r  rmsnorm_rsigma    : x -> rsigma[M] only (first pass of the MXFP8 variant, see mx_quantize.cu)
//   K5   rmsnorm_bwd       : dy, x, gamma, rsigma -> dx bf16, dgamma bf16 (rows through a cp.async.bulk ring; two-stage column reduction)
// K1 / K1r stay register-staged: a ring form of them (row pairs per 256-thread CTA) measured SLOWER on hardware (20.5 -> 22.5 us at
// [6144, 4096]: 16 resident CTAs of 4 warps hide the row's serial chain better than 2 CTAs of 8) and was removed (profiles/r02/NOTES.md, call 24).
// Replaces TE's rmsnorm_fwd(+fp8 cast)/rmsnorm_bwd reached from LayerNormLinear / LayerNormMLP
// (te_llama.py:45-63).  The sum-of-squares reduction ORDER is the one oracle_c.c:orc_rms_rsigma fixes:
// chunk c (8 elements) -> thread c % 128, in-thread fma chain, xor butterfly, ((w0+w1)+w2)+w3.
#include <mutex>
#include "common.cuh"
#include "ptx.cuh"

namespace llmfp8 {

constexpr int RMS_T = 128;

// Persistent row kernels run exactly one wave: grid = resident CTAs per SM (occupancy query) x SMs.  A grid of 1.1-1.4 waves
// costs a whole extra wave (measured: rmsnorm_bwd 592 CTAs on 444 slots ran at 36-40 % of HBM bandwidth).
template <typename K>
static int resident_grid(K kern, int threads, int64_t M) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, 0) != cudaSuccess || per_sm < 1) per_sm = 1;
    const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;
    return static_cast<int>(M < cap ? M : cap);
}

template <int NIT>
__device__ __forceinline__ float rms_row_rsigma(const uint4 (&v)[NIT], int t, int nchunk, int H, float eps, float* red /*[4]*/) {
    float acc = 0.f;
#pragma unroll
    for (int it = 0; it < NIT; it++) {
        if (it * RMS_T + t < nchunk) {
            float f[8];
            unpack8(v[it], f);
#pragma unroll
            for (int j = 0; j < 8; j++) acc = __fmaf_rn(f[j], f[j], acc);
        }
    }
    acc = warp_sum_pinned(acc);
    if ((t & 31) == 0) red[t >> 5] = acc;
    __syncthreads();
    const float total = __fadd_rn(__fadd_rn(__fadd_rn(red[0], red[1]), red[2]), red[3]);
    const float mean = __fdiv_rn(total, static_cast<float>(H));
    return __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(mean, eps)));
}

// OUT_MODE: 0 = fp8 quantised output (+amax), 1 = rsigma only
template <int FMT, int NIT, int OUT_MODE>
__global__ void __launch_bounds__(RMS_T) rmsnorm_fwd_kernel(const uint4* __restrict__ x, const uint4* __restrict__ gamma, uint2* __restrict__ q,
                                                            float* __restrict__ rsigma, int64_t M, int H, float eps,
                                                            const float* __restrict__ scale_ptr, float* __restrict__ amax) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[2][4];
    __shared__ float red_amax[4];
    const int t = threadIdx.x;
    const int nchunk = H / 8;
    const float scale = (OUT_MODE == 0 && scale_ptr) ? __ldg(scale_ptr) : 1.0f;
    uint4 g[NIT];
    if (OUT_MODE == 0) {
#pragma unroll
        for (int it = 0; it < NIT; it++)
            if (it * RMS_T + t < nchunk) g[it] = __ldg(gamma + it * RMS_T + t);
    }
    float m = 0.f;
    int par = 0;
    for (int64_t row = blockIdx.x; row < M; row += gridDim.x, par ^= 1) {
        uint4 v[NIT];
        const uint4* xr = x + row * nchunk;
#pragma unroll
        for (int it = 0; it < NIT; it++)
            if (it * RMS_T + t < nchunk) v[it] = ldg_stream(xr + it * RMS_T + t);
        const float rs = rms_row_rsigma<NIT>(v, t, nchunk, H, eps, red[par]);
        if (t == 0) rsigma[row] = rs;
        if (OUT_MODE == 0) {
            uint2* qr = q + row * nchunk;
#pragma unroll
            for (int it = 0; it < NIT; it++) {
                if (it * RMS_T + t < nchunk) {
                    float f[8], gf[8];
                    unpack8(v[it], f);
                    unpack8(g[it], gf);
#pragma unroll
                    for (int j = 0; j < 8; j++) f[j] = __fmul_rn(__fmul_rn(f[j], rs), gf[j]);
                    m = amax8(f, m);
                    qr[it * RMS_T + t] = quant8<FMT>(f, scale);
                }
            }
        }
    }
    if (OUT_MODE == 0 && amax) block_amax_to_global(m, red_amax, amax);
}

template <int FMT, int OUT_MODE>
static int dispatch_rms_fwd(int nit, int grid, cudaStream_t s, const uint4* x, const uint4* g, uint2* q, float* rs, int64_t M, int H, float eps,
                            const float* scale, float* amax) {
#define RMS_CASE(N)                                                                                              \
    case N:                                                                                                      \
        grid = resident_grid(rmsnorm_fwd_kernel<FMT, N, OUT_MODE>, RMS_T, M);                                    \
        LLMFP8_LAUNCH((rmsnorm_fwd_kernel<FMT, N, OUT_MODE>), grid, RMS_T, 0, s, x, g, q, rs, M, H, eps, scale, amax);        \
        break;
    switch (nit) {
        RMS_CASE(1) RMS_CASE(2) RMS_CASE(3) RMS_CASE(4) RMS_CASE(5) RMS_CASE(6) RMS_CASE(7) RMS_CASE(8)
        default: set_error("rmsnorm: hidden size %d too large (max 8192)", H); return -1;
    }
#undef RMS_CASE
    return 0;
}

int launch_rmsnorm_quant_fwd(const void* x, const void* gamma, void* q, float* rsigma, int64_t M, int64_t H, float eps, int fmt,
                             const float* scale, float* amax, cudaStream_t stream) {
    LLMFP8_REQUIRE(H % 8 == 0 && H > 0 && H <= 8192, "rmsnorm_quant_fwd: hidden size %lld must be a multiple of 8 and <= 8192", (long long)H);
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2, "rmsnorm_quant_fwd: bad fp8 format %d", fmt);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gamma)) % 16 == 0 && reinterpret_cast<uintptr_t>(q) % 8 == 0,
                   "rmsnorm_quant_fwd: misaligned pointers");
    if (M == 0) return 0;
    const int nit = static_cast<int>((H / 8 + RMS_T - 1) / RMS_T);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(M < cap ? M : cap);
    int rc = fmt == FMT_E4M3
                 ? dispatch_rms_fwd<FMT_E4M3, 0>(nit, grid, stream, (const uint4*)x, (const uint4*)gamma, (uint2*)q, rsigma, M, (int)H, eps, scale, amax)
                 : dispatch_rms_fwd<FMT_E5M2, 0>(nit, grid, stream, (const uint4*)x, (const uint4*)gamma, (uint2*)q, rsigma, M, (int)H, eps, scale, amax);
    if (rc) return rc;
    LLMFP8_LAUNCH_CHECK("rmsnorm_quant_fwd");
    return 0;
}

int launch_rmsnorm_rsigma(const void* x, float* rsigma, int64_t M, int64_t H, float eps, cudaStream_t stream) {
    LLMFP8_REQUIRE(H % 8 == 0 && H > 0 && H <= 8192, "rmsnorm_rsigma: hidden size %lld must be a multiple of 8 and <= 8192", (long long)H);
    LLMFP8_REQUIRE(reinterpret_cast<uintptr_t>(x) % 16 == 0, "rmsnorm_rsigma: misaligned pointer");
    if (M == 0) return 0;
    const int nit = static_cast<int>((H / 8 + RMS_T - 1) / RMS_T);
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(M < cap ? M : cap);
    int rc = dispatch_rms_fwd<FMT_E4M3, 1>(nit, grid, stream, (const uint4*)x, nullptr, nullptr, rsigma, M, (int)H, eps, nullptr, nullptr);
    if (rc) return rc;
    LLMFP8_LAUNCH_CHECK("rmsnorm_rsigma");
    return 0;
}

// ------------------------------------------------------------------------------------------- K5
// dx = rs * (dy*g - xhat * mean(dy*g*xhat)),  dgamma = sum_rows dy * xhat   (xhat = x * rs), fp32 math.
// 256 threads per row (two 16 B chunks per thread at H = 4096).
constexpr int RB_T = 256;
// ADD: dx += dres (the gradient arriving on the residual branch around the norm), so autograd needs no separate accumulation pass.
// Rows of x, dy (and dres) stream into a ring of shared-memory stages with 1-D bulk copies (cp.async.bulk, mbarrier completion) and the
// 256 threads take their 16 B chunks from the ring.  Thread 0 is the producer: the block barrier of the row reduction already proves
// that every thread holds its chunks of the current stage in registers, so right after it the slot is refilled with the row `stages`
// iterations ahead — no second barrier, no producer warp.  (Round 1's form prefetched ONE row per CTA into registers: 0.59 / 0.71 of
// the measured HBM peak without / with dres at [6144, 4096]; the ring: 0.65 / 0.82, profiles/r02/NOTES.md calls 23-24.  What is left
// without dres is the serial chain of one row per CTA and step — wait, reduce, barrier, store — not bytes in flight: 4 -> 6 ring
// stages changed nothing.)
template <int NIT, bool ADD>
__global__ void __launch_bounds__(RB_T, (NIT <= 2 ? 2 : 1)) rmsnorm_bwd_kernel(const uint8_t* __restrict__ dy, const uint8_t* __restrict__ x,
                                                                                   const uint4* __restrict__ gamma, const float* __restrict__ rsigma,
                                                                                   const uint8_t* __restrict__ dres, uint4* __restrict__ dx,
                                                                                   float* __restrict__ dgamma_part, int64_t M, int H, int stages) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    const uint32_t row_bytes = static_cast<uint32_t>(H) * 2u;
    const uint32_t stage_bytes = ((ADD ? 3u : 2u) * row_bytes + 127u) & ~127u;
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + stages * stage_bytes);
    float* red = reinterpret_cast<float*>(full + stages);           // [2][RB_T / 32]
    const int t = threadIdx.x;
    auto load_row = [&](int64_t row, int stage) {                   // thread 0 only
        uint8_t* dst = ring + stage * stage_bytes;
        ptx::mbar_arrive_expect_tx(&full[stage], (ADD ? 3u : 2u) * row_bytes);
        ptx::bulk_load_1d(dst, x + row * row_bytes, row_bytes, &full[stage]);
        ptx::bulk_load_1d(dst + row_bytes, dy + row * row_bytes, row_bytes, &full[stage]);
        if constexpr (ADD) ptx::bulk_load_1d(dst + 2 * row_bytes, dres + row * row_bytes, row_bytes, &full[stage]);
    };
    if (t == 0) {
        for (int i = 0; i < stages; i++) ptx::mbar_init(&full[i], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    if (t == 0) {
        int64_t row = blockIdx.x;
        for (int i = 0; i < stages && row < M; i++, row += gridDim.x) load_row(row, i);
    }

    const int nchunk = H / 8;
    f32x2 dg[NIT][4], g2[NIT][4];
#pragma unroll
    for (int it = 0; it < NIT; it++) {
        float gf[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (it * RB_T + t < nchunk) unpack8(__ldg(gamma + it * RB_T + t), gf);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            dg[it][k] = pk2(0.f, 0.f);
            g2[it][k] = pk2(gf[2 * k], gf[2 * k + 1]);
        }
    }
    const float inv_h = 1.0f / static_cast<float>(H);
    const int64_t ahead = static_cast<int64_t>(stages) * gridDim.x;
    int stage = 0, phase = 0, par = 0;
    int64_t row = blockIdx.x;
    float nrs = row < M ? rsigma[row] : 0.f;
    for (; row < M; row += gridDim.x, par ^= 1) {
        const float rs = nrs;
        if (row + gridDim.x < M) nrs = rsigma[row + gridDim.x];
        const f32x2 rs2 = pk2(rs, rs);
        ptx::mbar_wait(&full[stage], phase);
        const uint4* sx = reinterpret_cast<const uint4*>(ring + stage * stage_bytes);
        const uint4* sdy = reinterpret_cast<const uint4*>(ring + stage * stage_bytes + row_bytes);
        const uint4* sres = reinterpret_cast<const uint4*>(ring + stage * stage_bytes + 2 * row_bytes);
        f32x2 xh[NIT][4], tg[NIT][4];       // xhat = x * rs ; tg = dy * gamma
        f32x2 fr[ADD ? NIT : 1][4];         // the residual-branch gradient, unpacked BEFORE the barrier that frees the stage
        f32x2 dot2 = pk2(0.f, 0.f);
#pragma unroll
        for (int it = 0; it < NIT; it++)
            if (it * RB_T + t < nchunk) {
                float fx[8], fdy[8];
                unpack8(sx[it * RB_T + t], fx);
                unpack8(sdy[it * RB_T + t], fdy);
                if constexpr (ADD) {
                    float f[8];
                    unpack8(sres[it * RB_T + t], f);
#pragma unroll
                    for (int k = 0; k < 4; k++) fr[it][k] = pk2(f[2 * k], f[2 * k + 1]);
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const f32x2 dy2 = pk2(fdy[2 * k], fdy[2 * k + 1]);
                    xh[it][k] = mul2(pk2(fx[2 * k], fx[2 * k + 1]), rs2);
                    tg[it][k] = mul2(dy2, g2[it][k]);
                    dot2 = fma2(tg[it][k], xh[it][k], dot2);
                    dg[it][k] = fma2(dy2, xh[it][k], dg[it][k]);
                }
            }
        float d0, d1;
        upk2(dot2, d0, d1);
        const float dot = warp_sum(d0 + d1);
        if ((t & 31) == 0) red[par * (RB_T / 32) + (t >> 5)] = dot;
        __syncthreads();                                             // also: every thread has its chunks of this stage in registers
        if (t == 0 && row + ahead < M) load_row(row + ahead, stage);
        if (++stage == stages) { stage = 0; phase ^= 1; }
        float tot = 0.f;
#pragma unroll
        for (int w = 0; w < RB_T / 32; w++) tot += red[par * (RB_T / 32) + w];
        // dx = rs * tg - (rs * c) * xhat
        const float nc = -(rs * (tot * inv_h));
        const f32x2 nc2 = pk2(nc, nc);
#pragma unroll
        for (int it = 0; it < NIT; it++)
            if (it * RB_T + t < nchunk) {
                uint32_t o[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    float lo, hi;
                    upk2(fma2(xh[it][k], nc2, mul2(tg[it][k], rs2)), lo, hi);
                    if constexpr (ADD) {
                        float rlo, rhi;
                        upk2(fr[it][k], rlo, rhi);
                        lo += rlo;
                        hi += rhi;
                    }
                    o[k] = pack_bf16x2(lo, hi);
                }
                dx[row * nchunk + it * RB_T + t] = make_uint4(o[0], o[1], o[2], o[3]);
            }
    }
    float* out = dgamma_part + static_cast<int64_t>(blockIdx.x) * H;
#pragma unroll
    for (int it = 0; it < NIT; it++)
        if (it * RB_T + t < nchunk) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 4; k++) upk2(dg[it][k], v[2 * k], v[2 * k + 1]);
            float4* o4 = reinterpret_cast<float4*>(out + (it * RB_T + t) * 8);
            o4[0] = make_float4(v[0], v[1], v[2], v[3]);
            o4[1] = make_float4(v[4], v[5], v[6], v[7]);
        }
}

// dgamma[c] = sum over the per-CTA partial rows.  CTA = 32 columns x 32 partial-row lanes: every warp reads 128 B rows, each thread
// keeps four independent accumulators (the one-thread-per-column chain this replaces spent ~30 us on 444 dependent loads).
constexpr int CR_COLS = 32, CR_LANES = 32;     // 1024 threads: <= 10 partial rows per thread at 296 CTAs' worth of partials (8 lanes: 37 rows, 13.8 us)
__global__ void __launch_bounds__(CR_COLS * CR_LANES) colreduce_bf16_kernel(const float* __restrict__ part, int nparts, int H,
                                                                            uint16_t* __restrict__ out) {
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    __shared__ float red[CR_LANES][CR_COLS + 1];
    const int cx = threadIdx.x % CR_COLS, py = threadIdx.x / CR_COLS;
    const int c = blockIdx.x * CR_COLS + cx;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c < H) {
        const float* src = part + c;
        int p = py;
        for (; p + 3 * CR_LANES < nparts; p += 4 * CR_LANES) {
            s0 += src[static_cast<int64_t>(p) * H];
            s1 += src[static_cast<int64_t>(p + CR_LANES) * H];
            s2 += src[static_cast<int64_t>(p + 2 * CR_LANES) * H];
            s3 += src[static_cast<int64_t>(p + 3 * CR_LANES) * H];
        }
        for (; p < nparts; p += CR_LANES) s0 += src[static_cast<int64_t>(p) * H];
    }
    red[py][cx] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (py == 0 && c < H) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < CR_LANES; j++) s += red[j][cx];
        out[c] = static_cast<uint16_t>(pack_bf16x2(s, 0.f) & 0xffffu);
    }
}

int rmsnorm_bwd_grid(int64_t M) {
    const int64_t cap = static_cast<int64_t>(sm_count()) * 4;
    return static_cast<int>(M < cap ? M : cap);
}

// ws: fp32 [rmsnorm_bwd_grid(M) * H] (an upper bound of the grid actually launched)
int launch_rmsnorm_bwd(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma_bf16, float* ws,
                       int64_t M, int64_t H, cudaStream_t stream) {
    LLMFP8_REQUIRE(H % 8 == 0 && H > 0 && H <= 8192, "rmsnorm_bwd: hidden size %lld must be a multiple of 8 and <= 8192", (long long)H);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(dx) |
                    reinterpret_cast<uintptr_t>(ws) | reinterpret_cast<uintptr_t>(dres)) % 16 == 0, "rmsnorm_bwd: misaligned pointers");
    if (M == 0) return 0;
    const int nit = static_cast<int>((H / 8 + RB_T - 1) / RB_T);
    int grid = 0;
    {
        // ring: as many whole-row stages as fit in ~104 KB (two CTAs per SM) or ~208 KB (one CTA: H > 4096), at most 8
        const int64_t stage_bytes = (((dres ? 3 : 2) * H * 2) + 127) / 128 * 128;
        const int64_t budget = nit <= 2 ? 104 * 1024 : 208 * 1024;
        int stages = static_cast<int>(budget / stage_bytes);
        stages = stages > 8 ? 8 : stages;
        LLMFP8_REQUIRE(stages >= 2, "rmsnorm_bwd: hidden size %lld leaves no room for a two-stage ring", (long long)H);
        const int smem = static_cast<int>(stages * stage_bytes + stages * 8 + 2 * (RB_T / 32) * 4 + 128);
#define RMSB_TMA(N, ADD)                                                                                                                       \
    do {                                                                                                                                       \
        auto kern = rmsnorm_bwd_kernel<N, ADD>;                                                                                            \
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);                                         \
        if (e != cudaSuccess) { set_error("rmsnorm_bwd: smem attr: %s", cudaGetErrorString(e)); return (int)e; }                               \
        int per_sm = 0;                                                                                                                        \
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, RB_T, smem) != cudaSuccess || per_sm < 1) per_sm = 1;           \
        const int64_t cap = static_cast<int64_t>(sm_count()) * per_sm;                                                                         \
        grid = static_cast<int>(M < cap ? M : cap);                                                                                            \
        if (grid > rmsnorm_bwd_grid(M)) grid = rmsnorm_bwd_grid(M);                                                                            \
        LLMFP8_LAUNCH(kern, grid, RB_T, smem, stream, (const uint8_t*)dy, (const uint8_t*)x, (const uint4*)gamma, rsigma, (const uint8_t*)dres, \
                      (uint4*)dx, ws, M, (int)H, stages);                                                                                      \
    } while (0)
#define RMSB_TMA_CASE(N)                 \
    case N:                              \
        if (dres) RMSB_TMA(N, true);     \
        else RMSB_TMA(N, false);         \
        break;
        switch (nit) {
            RMSB_TMA_CASE(1) RMSB_TMA_CASE(2) RMSB_TMA_CASE(3) RMSB_TMA_CASE(4)
            default: set_error("rmsnorm_bwd: hidden size too large"); return -1;
        }
#undef RMSB_TMA
#undef RMSB_TMA_CASE
        LLMFP8_LAUNCH_CHECK("rmsnorm_bwd");
        LLMFP8_LAUNCH((colreduce_bf16_kernel), static_cast<unsigned>((H + CR_COLS - 1) / CR_COLS), CR_COLS * CR_LANES, 0, stream, ws, grid, (int)H, static_cast<uint16_t*>(dgamma_bf16));
        LLMFP8_LAUNCH_CHECK("rmsnorm_bwd.dgamma");
        return 0;
    }
}

// ------------------------------------------------------------------------------------------- K5 without caller-provided scratch
// llmfp8_rmsnorm_bwd_noalloc: K5 for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
// The dgamma partial rows ([launched CTAs, H] fp32, every element written by the kernel before the column reduce reads it: nothing has
// to start at zero) live in a file-scope device array instead of a caller workspace.  Every GPU that loads the library holds the array
// until exit; every call on a device shares it, so calls that overlap on that GPU (two streams) race on it — serialise them.
// In host code `g_k5_scratch` is a host-side stand-in, not a device address: the address the kernel gets comes from
// cudaGetSymbolAddress, taken once per device with that device current.  This path calls no cudaMalloc / cudaFree / allocating routine.
constexpr int64_t K5_SCRATCH_FLOATS = 592ll * 4096;                  // 9.25 MiB: rmsnorm_bwd_grid(M) = 4 x 148 partial rows of H <= 4096
static __device__ __align__(256) float g_k5_scratch[K5_SCRATCH_FLOATS];
constexpr int K5_MAX_DEVICES = 64;

int64_t rmsnorm_bwd_noalloc_max_floats() { return K5_SCRATCH_FLOATS; }

int launch_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma_bf16,
                               int64_t M, int64_t H, cudaStream_t stream) {
    LLMFP8_REQUIRE(M >= 0 && H > 0, "rmsnorm_bwd_noalloc: bad sizes M=%lld H=%lld", (long long)M, (long long)H);
    const int64_t need = static_cast<int64_t>(rmsnorm_bwd_grid(M)) * H;
    LLMFP8_REQUIRE(need <= K5_SCRATCH_FLOATS,
                   "rmsnorm_bwd_noalloc: %lld partial rows x H = %lld need %lld floats of scratch, the built-in array holds %lld "
                   "(H <= 4096 at any M; larger H only while rows x H fits) - use llmfp8_rmsnorm_bwd with a caller workspace",
                   (long long)rmsnorm_bwd_grid(M), (long long)H, (long long)need, (long long)K5_SCRATCH_FLOATS);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("rmsnorm_bwd_noalloc: cudaGetDevice: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_REQUIRE(dev >= 0 && dev < K5_MAX_DEVICES, "rmsnorm_bwd_noalloc: device ordinal %d out of range", dev);
    static std::mutex mu;
    static float* dev_ptr[K5_MAX_DEVICES] = {};
    float* ws = nullptr;
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!dev_ptr[dev]) {
            float* p = nullptr;
            e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_k5_scratch);      // this device's copy of the array
            if (e != cudaSuccess || !p) { set_error("rmsnorm_bwd_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return e != cudaSuccess ? (int)e : -1; }
            dev_ptr[dev] = p;
        }
        ws = dev_ptr[dev];
    }
    return launch_rmsnorm_bwd(dy, x, gamma, rsigma, dres, dx, dgamma_bf16, ws, M, H, stream);
}

}  // namespace llmfp8
