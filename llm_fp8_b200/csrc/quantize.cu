// Per-tensor (delayed-scaling) quantise kernels for sm_100a.
//   K2  cast_amax        bf16 [n]   -> fp8 [n], amax                     (3 B/element of HBM traffic)
//   K2b cast_amax_bgrad  bf16 [R,C] -> fp8 [R,C], amax, column sums      (dY cast + bias grad in one pass)
//   K4  amax_scale_update  amax history roll + scale / scale_inv update  (one launch for all fp8 tensors)
// Replaces TE's cast/cast_transpose_dbias + amax_and_scale_update ops reached from te_llama.py:76-80.
// Arithmetic is the oracle's (oracle/oracle_c.c orc_quantize_bf16 / orc_amax_scale_update), bit for bit.
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

// ------------------------------------------------------------------------------------------- K2
// K2 through a shared-memory ring: persistent CTAs (two per SM), each streaming 16 KB chunks of x with 1-D bulk copies (cp.async.bulk,
// mbarrier completion) into a ring that keeps `stages - 1` chunks in flight per CTA; the 256 threads take one 16 B unit per 4 KB sub-block
// from the ring, quantise and store 8 B.  Thread 0 is the producer: the block barrier after the ring reads proves the slot is free.
// (Round 1's register-staged form — 8 loads per thread in flight, then none while the CTA computes and retires — measured 0.832 of the
// HBM peak on the fc1 weight against 0.858 here; at [M, H] sizes both take 20.5 us: a kernel that short is ramp and tail.)
constexpr int CR_CHUNK = 16384;                 // bytes of bf16 input per stage
constexpr int CR_UNITS = CR_CHUNK / 16 / 256;   // 16 B units per thread and stage
template <int FMT>
__global__ void __launch_bounds__(256, 2) cast_amax_kernel(const uint8_t* __restrict__ x, uint2* __restrict__ q, int64_t nbytes,
                                                                const float* __restrict__ scale_ptr, float* __restrict__ amax, int stages) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + stages * CR_CHUNK);
    __shared__ float red[8];
    const int t = threadIdx.x;
    const int64_t nchunks = (nbytes + CR_CHUNK - 1) / CR_CHUNK;
    auto load_chunk = [&](int64_t c, int stage) {                   // thread 0 only
        const int64_t off = c * CR_CHUNK;
        const uint32_t bytes = static_cast<uint32_t>(nbytes - off < CR_CHUNK ? nbytes - off : CR_CHUNK);
        ptx::mbar_arrive_expect_tx(&full[stage], bytes);
        ptx::bulk_load_1d(ring + stage * CR_CHUNK, x + off, bytes, &full[stage]);
    };
    if (t == 0) {
        for (int i = 0; i < stages; i++) ptx::mbar_init(&full[i], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    if (t == 0) {
        int64_t c = blockIdx.x;
        for (int i = 0; i < stages && c < nchunks; i++, c += gridDim.x) load_chunk(c, i);
    }
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const f32x2 scale2 = pk2(scale, scale);
    const int64_t ahead = static_cast<int64_t>(stages) * gridDim.x;
    float m = 0.f;
    int stage = 0, phase = 0;
    for (int64_t c = blockIdx.x; c < nchunks; c += gridDim.x) {
        const int64_t unit0 = c * (CR_CHUNK / 16);
        const int64_t nunits = nbytes / 16;
        ptx::mbar_wait(&full[stage], phase);
        const uint4* src = reinterpret_cast<const uint4*>(ring + stage * CR_CHUNK);
        uint4 v[CR_UNITS];
#pragma unroll
        for (int u = 0; u < CR_UNITS; u++) v[u] = unit0 + u * 256 + t < nunits ? src[u * 256 + t] : make_uint4(0, 0, 0, 0);
        uint2 o[CR_UNITS];
#pragma unroll
        for (int u = 0; u < CR_UNITS; u++) {
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
            o[u] = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
        }
        __syncthreads();                                            // every thread's units of this slot are converted: registers only from here
        if (t == 0 && c + ahead < nchunks) load_chunk(c + ahead, stage);
        if (++stage == stages) { stage = 0; phase ^= 1; }
#pragma unroll
        for (int u = 0; u < CR_UNITS; u++)
            if (unit0 + u * 256 + t < nunits) q[unit0 + u * 256 + t] = o[u];
    }
    if (amax) block_amax_to_global(m, red, amax);
}

int launch_cast_amax(const void* x, void* q, int64_t n, int fmt, const float* scale, float* amax, cudaStream_t stream) {
    LLMFP8_REQUIRE(n % 16 == 0, "cast_amax: element count %lld must be a multiple of 16", (long long)n);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(q)) % 16 == 0, "cast_amax: pointers must be 16 B aligned");
    LLMFP8_REQUIRE(fmt == FMT_E4M3 || fmt == FMT_E5M2, "cast_amax: bad fp8 format %d", fmt);
    if (n == 0) return 0;
    {
        const int stages = 6;                                       // 96 KB per CTA, two CTAs per SM: 160 KB of loads in flight per SM
        const int smem = stages * CR_CHUNK + stages * 8 + 128;
        const int64_t nchunks = (n * 2 + CR_CHUNK - 1) / CR_CHUNK;
        const int64_t cap = static_cast<int64_t>(sm_count()) * 2;
        const int grid = static_cast<int>(nchunks < cap ? nchunks : cap);
        cudaError_t e;
        if (fmt == FMT_E4M3) {
            e = cudaFuncSetAttribute(cast_amax_kernel<FMT_E4M3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) { set_error("cast_amax: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
            LLMFP8_LAUNCH((cast_amax_kernel<FMT_E4M3>), grid, 256, smem, stream, static_cast<const uint8_t*>(x), static_cast<uint2*>(q), n * 2, scale, amax, stages);
        } else {
            e = cudaFuncSetAttribute(cast_amax_kernel<FMT_E5M2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) { set_error("cast_amax: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
            LLMFP8_LAUNCH((cast_amax_kernel<FMT_E5M2>), grid, 256, smem, stream, static_cast<const uint8_t*>(x), static_cast<uint2*>(q), n * 2, scale, amax, stages);
        }
        LLMFP8_LAUNCH_CHECK("cast_amax");
        return 0;
    }
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
// The strip's rows arrive as 32 x 256 boxes by TMA (one cp.async.bulk.tensor per stage, zero-filled past the matrix edge) into a ring that
// keeps `stages - 1` boxes in flight; the grid is one wave of two CTAs per SM.  Measured 0.42 of the HBM peak on dY [6144, 4096] (round 1's
// register-staged form: 0.39) — the 512 B row segments of a strip, not the staging, are what holds this kernel back; a row-streaming form
// with per-CTA partial sums (K5's layout) is the next step, and needs a [grid, C] workspace the C ABI does not hand over today.
constexpr int BGR_ROWS = 32;                                // rows per stage
constexpr int BGR_STAGE = BGR_ROWS * BG_COLS * 2;           // 16 KB
template <int FMT>
__global__ void __launch_bounds__(256, 2) cast_amax_bgrad_kernel(const __grid_constant__ CUtensorMap map_x, uint2* __restrict__ q, int64_t R, int64_t C,
                                                                      int64_t rows_per_cta, const float* __restrict__ scale_ptr, float* __restrict__ amax,
                                                                      float* __restrict__ colsum_ws, int stages) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u);
    uint64_t* full = reinterpret_cast<uint64_t*>(ring + stages * BGR_STAGE);
    __shared__ float red[8];
    __shared__ float part[BG_RL][BG_COLS + 8];
    const int t = threadIdx.x;
    const int cv = t % BG_CV, rl = t / BG_CV;
    const int strip_col = blockIdx.x * BG_COLS;
    const int64_t col0 = static_cast<int64_t>(strip_col) + cv * 8;
    const int64_t r_begin = static_cast<int64_t>(blockIdx.y) * rows_per_cta;
    const int64_t r_end = r_begin + rows_per_cta < R ? r_begin + rows_per_cta : R;
    const int nboxes = r_end > r_begin ? static_cast<int>((r_end - r_begin + BGR_ROWS - 1) / BGR_ROWS) : 0;
    auto load_box = [&](int b, int stage) {                         // thread 0 only
        ptx::mbar_arrive_expect_tx(&full[stage], BGR_STAGE);
        ptx::tma_load_2d(&map_x, &full[stage], ring + stage * BGR_STAGE, strip_col, static_cast<int>(r_begin) + b * BGR_ROWS);
    };
    if (t == 0) {
        ptx::prefetch_tensormap(&map_x);
        for (int i = 0; i < stages; i++) ptx::mbar_init(&full[i], 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh); empty otherwise
    pdl_launch_dependents();
    if (t == 0)
        for (int i = 0; i < stages && i < nboxes; i++) load_box(i, i);
    const float scale = scale_ptr ? __ldg(scale_ptr) : 1.0f;
    const bool col_ok = col0 < C;
    const int64_t rowvec = C / 8;
    float m = 0.f;
    float cs[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    int stage = 0, phase = 0;
    for (int b = 0; b < nboxes; b++) {
        ptx::mbar_wait(&full[stage], phase);
        const uint8_t* src = ring + stage * BGR_STAGE;
        uint4 v[BGR_ROWS / BG_RL];
#pragma unroll
        for (int u = 0; u < BGR_ROWS / BG_RL; u++) v[u] = *reinterpret_cast<const uint4*>(src + (rl + u * BG_RL) * (BG_COLS * 2) + cv * 16);
        uint2 o[BGR_ROWS / BG_RL];
#pragma unroll
        for (int u = 0; u < BGR_ROWS / BG_RL; u++) {
            // rows past r_end inside the box belong to the next CTA (or are zero fill past R): they must not enter this CTA's sums
            const bool mine = r_begin + b * BGR_ROWS + rl + u * BG_RL < r_end;
            float f[8];
            unpack8(mine ? v[u] : make_uint4(0, 0, 0, 0), f);
            m = amax8(f, m);
#pragma unroll
            for (int i = 0; i < 8; i++) cs[i] += f[i];
            o[u] = quant8<FMT>(f, scale);
        }
        __syncthreads();                                            // the slot is in registers everywhere
        if (t == 0 && b + stages < nboxes) load_box(b + stages, stage);
        if (++stage == stages) { stage = 0; phase ^= 1; }
        if (col_ok) {
#pragma unroll
            for (int u = 0; u < BGR_ROWS / BG_RL; u++) {
                const int64_t r = r_begin + b * BGR_ROWS + rl + u * BG_RL;
                if (r < r_end) q[r * rowvec + col0 / 8] = o[u];
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
    LLMFP8_REQUIRE(R < (int64_t(1) << 31) && C < (int64_t(1) << 31), "cast_amax_bgrad: dims too large");
    {
        int64_t yc = (static_cast<int64_t>(sm_count()) * 2) / strips;          // one wave of two CTAs per SM
        yc = yc < 1 ? 1 : yc;
        int64_t rpc = (R + yc - 1) / yc;
        rpc = (rpc + BGR_ROWS - 1) / BGR_ROWS * BGR_ROWS;
        yc = (R + rpc - 1) / rpc;
        CUtensorMap m_x;
        int rc = make_tensor_map_2d(&m_x, x, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, C, R, C * 2, BG_COLS, BGR_ROWS, CU_TENSOR_MAP_SWIZZLE_NONE);
        if (rc) return rc;
        const int stages = 6;
        const int smem = stages * BGR_STAGE + stages * 8 + 128;
        dim3 rgrid(strips, static_cast<unsigned>(yc));
        if (fmt == FMT_E4M3) {
            e = cudaFuncSetAttribute(cast_amax_bgrad_kernel<FMT_E4M3>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) { set_error("cast_amax_bgrad: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
            LLMFP8_LAUNCH((cast_amax_bgrad_kernel<FMT_E4M3>), rgrid, 256, smem, stream, m_x, static_cast<uint2*>(q), R, C, rpc, scale, amax, colsum_ws, stages);
        } else {
            e = cudaFuncSetAttribute(cast_amax_bgrad_kernel<FMT_E5M2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
            if (e != cudaSuccess) { set_error("cast_amax_bgrad: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
            LLMFP8_LAUNCH((cast_amax_bgrad_kernel<FMT_E5M2>), rgrid, 256, smem, stream, m_x, static_cast<uint2*>(q), R, C, rpc, scale, amax, colsum_ws, stages);
        }
        LLMFP8_LAUNCH_CHECK("cast_amax_bgrad");
        LLMFP8_LAUNCH((f32_to_bf16_kernel), static_cast<unsigned>((C + 255) / 256), 256, 0, stream, colsum_ws, static_cast<uint16_t*>(bgrad_bf16), C);
        LLMFP8_LAUNCH_CHECK("cast_amax_bgrad.finalize");
        return 0;
    }
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
