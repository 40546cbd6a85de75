// Activation + MXFP8 block quantiser in one pass for sm_100a (K3m / K3m-bwd) — SURVEY.md §8(f) rank 2.
//   fwd:  fc1_out bf16 [M, 2I] (gate | up)              -> a   = bf16(silu(gate) * up)  -> rowwise + colwise MXFP8 copies of a   [M, I]
//   bwd:  da bf16 [M, I], fc1_out bf16 [M, 2I]          -> dfc1 = bf16(dgate | dup)      -> rowwise + colwise MXFP8 copies of dfc1 [M, 2I],
//                                                          dbias[2I] = column sums of the fp32 (dgate | dup)
// Results are defined as mx_quantize(swiglu_bf16(fc1)) and mx_quantize(dswiglu_bf16(da, fc1)) — the two-kernel sequences of
// functional.MXLayerNormMLPFn, which tests/test_gpu_quantize.py pins bit-exactly to the oracle — so the bf16 rounding of the
// intermediate is kept; what disappears is its round trip through HBM (fwd: 4·M·I of 10.06·M·I bytes, bwd: 8·M·I of 18.1·M·I).
// Replaces TE's swiglu / dswiglu + mxfp8 cast inside LayerNormMLP(activation="swiglu") under te_llama_mxfp8.py:93.
//
// One CTA walks 128 x 128 tiles of the [M, I] index space.  Per tile: TMA loads of the gate, up (and da) tiles -> the activation is
// computed in place (bf16, into the gate / up buffers) -> the 1x32 and 32x1 block passes of mx_quantize.cu run over those buffers ->
// e4m3 tiles staged in shared memory -> TMA store; scale tiles (512 B, pre-swizzled for tcgen05.cp) written with 16 B stores.
// STATUS: written after the round-1 GPU budget was spent — not run on hardware yet.  Opt-in (LLMFP8_MX_FUSED_ACT=1), gated tests.
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

constexpr int AT = 128;                     // tile edge
constexpr int AT_IN_BYTES = AT * AT * 2;    // 32 KB bf16 tile
constexpr int AT_OUT_BYTES = AT * AT;       // 16 KB e4m3 tile

// Rowwise pass over one bf16 tile in shared memory (row pitch 256 B): thread `t` of 256 owns whole 1x32 blocks.
// Same arithmetic and shared-memory access order as mx_quantize_kernel<0> (mx_quantize.cu), which stays untouched.
__device__ __forceinline__ void act_row_pass(const uint8_t* in, uint8_t* out_row, uint8_t* sf_row, int t) {
#pragma unroll 1
    for (int i = 0; i < (AT * AT / 32) / 256; i++) {
        const int idx = i * 256 + t;
        const int r = idx >> 2, blk = idx & 3;
        const int rot = (t >> 1) & 3;
        const uint8_t* src = in + r * (AT * 2) + blk * 64;
        f32x2 f[4][4];
        float am = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int seg = (j + rot) & 3;
            const uint4 v = *reinterpret_cast<const uint4*>(src + seg * 16);
            float x[8];
            unpack8(v, x);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                am = fmaxf(am, fmaxf(fabsf(x[2 * k]), fabsf(x[2 * k + 1])));
                f[j][k] = pk2(x[2 * k], x[2 * k + 1]);
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
            *reinterpret_cast<uint2*>(out_row + r * AT + blk * 32 + seg * 8) = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
        }
        sf_row[(r & 31) * 16 + (r >> 5) * 4 + blk] = static_cast<uint8_t>(e);
    }
}

// Colwise pass: thread `t` of 256 owns 2 adjacent columns x 32 rows = two 32x1 blocks.
__device__ __forceinline__ void act_col_pass(const uint8_t* in, uint8_t* out_col, uint8_t* sf_col, int t) {
    const int cp = t & 63, rb = t >> 6;
    uint32_t pk[32];
#pragma unroll
    for (int r = 0; r < 32; r++) pk[r] = *reinterpret_cast<const uint32_t*>(in + (rb * 32 + r) * (AT * 2) + cp * 4);
    // |x| of both columns at once as u16x2 integers (non-negative bf16 values order like their bit patterns)
    uint32_t mx = 0;
#pragma unroll
    for (int r = 0; r < 32; r++) {
        uint32_t m;
        asm("max.u16x2 %0, %1, %2;" : "=r"(m) : "r"(mx), "r"(pk[r] & 0x7fff7fffu));
        mx = m;
    }
    // NaN patterns (> 0x7f80) must not win: fmaxf in the oracle drops NaN.  Rare: redo both columns the slow way.
    float am0 = bf16lo(mx), am1 = bf16hi(mx);
    if (((mx & 0xffffu) > 0x7f80u) || ((mx >> 16) > 0x7f80u)) {
        am0 = am1 = 0.f;
#pragma unroll
        for (int r = 0; r < 32; r++) {
            am0 = fmaxf(am0, fabsf(bf16lo(pk[r])));
            am1 = fmaxf(am1, fabsf(bf16hi(pk[r])));
        }
    }
    const uint32_t e0 = e8m0_from_amax(am0), e1 = e8m0_from_amax(am1);
    const f32x2 inv2 = pk2(e8m0_inv_scale(e0), e8m0_inv_scale(e1));
#pragma unroll
    for (int r = 0; r < 32; r++) {
        float lo, hi;
        upk2(mul2(pk2(bf16lo(pk[r]), bf16hi(pk[r])), inv2), lo, hi);
        *reinterpret_cast<uint16_t*>(out_col + (rb * 32 + r) * AT + cp * 2) = static_cast<uint16_t>(cvt_fp8x2<FMT_E4M3>(lo, hi));
    }
    const int c0 = cp * 2, c1 = cp * 2 + 1;
    sf_col[(c0 & 31) * 16 + (c0 >> 5) * 4 + rb] = static_cast<uint8_t>(e0);
    sf_col[(c1 & 31) * 16 + (c1 >> 5) * 4 + rb] = static_cast<uint8_t>(e1);
}

// ------------------------------------------------------------------------------------------------ forward: SwiGLU + quantise
struct ActFwdSmem {
    alignas(128) uint8_t g[AT_IN_BYTES];        // gate tile; overwritten in place by a = bf16(silu(gate) * up)
    alignas(128) uint8_t u[AT_IN_BYTES];        // up tile
    alignas(128) uint8_t out_row[AT_OUT_BYTES];
    alignas(128) uint8_t out_col[AT_OUT_BYTES];
    alignas(16) uint8_t sf_row[512];
    alignas(16) uint8_t sf_col[512];
    alignas(8) uint64_t full_g, full_u;
};

// 97 KB of shared memory per CTA: two CTAs per SM, each single-buffered — the two CTAs overlap each other's load and compute phases,
// and within a CTA the up tile of the next tile is requested as soon as the activation pass is done with it, the gate tile once the
// block passes have read the activation.
__global__ void __launch_bounds__(256, 2) mx_swiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_u,
                                                                    const __grid_constant__ CUtensorMap map_row, const __grid_constant__ CUtensorMap map_col,
                                                                    uint8_t* __restrict__ sf_row, uint8_t* __restrict__ sf_col, int M, int I, int do_row,
                                                                    int do_col) {
    extern __shared__ uint8_t smem_raw[];
    ActFwdSmem& s = *reinterpret_cast<ActFwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x;
    const int tiles_c = (I + AT - 1) / AT, tiles_r = (M + AT - 1) / AT;
    const int ntiles = tiles_c * tiles_r;

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        ptx::mbar_init(&s.full_g, 1);
        ptx::mbar_init(&s.full_u, 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();
    if (tid == 0 && static_cast<int>(blockIdx.x) < ntiles) {
        const int t = blockIdx.x;
        ptx::mbar_arrive_expect_tx(&s.full_g, AT_IN_BYTES);
        ptx::tma_load_2d(&map_g, &s.full_g, s.g, (t % tiles_c) * AT, (t / tiles_c) * AT);
        ptx::mbar_arrive_expect_tx(&s.full_u, AT_IN_BYTES);
        ptx::tma_load_2d(&map_u, &s.full_u, s.u, (t % tiles_c) * AT, (t / tiles_c) * AT);
    }

    int it = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, it++) {
        const uint32_t phase = it & 1;
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * AT, col0 = tc * AT;
        const int tn = t + gridDim.x;
        if (tid == 0) ptx::tma_store_wait_read<0>();   // the previous tile's stores no longer read out_row / out_col
        ptx::mbar_wait(&s.full_g, phase);
        ptx::mbar_wait(&s.full_u, phase);

        // activation pass, in place: 2048 16-byte vectors, 8 per thread, consecutive threads on consecutive vectors.
        // Arithmetic = swiglu_quant_kernel<FMT_BF16> (swiglu.cu): a = (g * sigmoid(g)) * up, each product rounded, then RNE to bf16.
#pragma unroll 2
        for (int i = 0; i < (AT * AT / 8) / 256; i++) {
            const int idx = i * 256 + tid;
            const uint4 vg = reinterpret_cast<const uint4*>(s.g)[idx];
            const uint4 vu = reinterpret_cast<const uint4*>(s.u)[idx];
            float g[8], up[8];
            unpack8(vg, g);
            unpack8(vu, up);
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const f32x2 sig = sigmoid2_pinned(g[2 * j], g[2 * j + 1]);
                float lo, hi;
                upk2(mul2(mul2(pk2(g[2 * j], g[2 * j + 1]), sig), pk2(up[2 * j], up[2 * j + 1])), lo, hi);
                o[j] = pack_bf16x2(lo, hi);
            }
            reinterpret_cast<uint4*>(s.g)[idx] = make_uint4(o[0], o[1], o[2], o[3]);
        }
        __syncthreads();                               // S1: activation tile complete, up buffer dead, out buffers free
        if (tid == 0 && tn < ntiles) {
            ptx::mbar_arrive_expect_tx(&s.full_u, AT_IN_BYTES);
            ptx::tma_load_2d(&map_u, &s.full_u, s.u, (tn % tiles_c) * AT, (tn / tiles_c) * AT);
        }
        if (do_row) act_row_pass(s.g, s.out_row, s.sf_row, tid);
        if (do_col) act_col_pass(s.g, s.out_col, s.sf_col, tid);
        ptx::fence_proxy_async();                      // generic writes (out tiles, the in-place activation) before async-proxy reads / writes
        __syncthreads();                               // S2
        if (tid == 0) {
            if (do_row) ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
            if (do_col) ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
            ptx::tma_store_commit();
            if (tn < ntiles) {
                ptx::mbar_arrive_expect_tx(&s.full_g, AT_IN_BYTES);
                ptx::tma_load_2d(&map_g, &s.full_g, s.g, (tn % tiles_c) * AT, (tn / tiles_c) * AT);
            }
        }
        if (do_row && tid >= 32 && tid < 64) {
            const int64_t off = (static_cast<int64_t>(tr) * tiles_c + tc) * 512;
            reinterpret_cast<uint4*>(sf_row + off)[tid - 32] = reinterpret_cast<const uint4*>(s.sf_row)[tid - 32];
        }
        if (do_col && tid >= 64 && tid < 96) {
            const int64_t off = (static_cast<int64_t>(tc) * tiles_r + tr) * 512;
            reinterpret_cast<uint4*>(sf_col + off)[tid - 64] = reinterpret_cast<const uint4*>(s.sf_col)[tid - 64];
        }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// ------------------------------------------------------------------------------------------------ backward: dSwiGLU + quantise (+ dbias)
constexpr int AB_THREADS = 512;             // warps 0-7 quantise the dgate tile, warps 8-15 the dup tile, all 16 run the activation pass
struct ActBwdSmem {
    alignas(128) uint8_t da[AT_IN_BYTES];
    alignas(128) uint8_t g[AT_IN_BYTES];         // gate tile -> dgate (bf16), in place
    alignas(128) uint8_t u[AT_IN_BYTES];         // up tile   -> dup   (bf16), in place
    alignas(128) uint8_t out_row[2][AT_OUT_BYTES];
    alignas(128) uint8_t out_col[2][AT_OUT_BYTES];
    alignas(16) uint8_t sf_row[2][512];
    alignas(16) uint8_t sf_col[2][512];
    float part[AB_THREADS / 32][2 * AT];         // per-warp column sums of the fp32 dgate (0..127) and dup (128..255) values
    alignas(8) uint64_t full_da, full_gu;
};

// Output matrix dfc1 is [M, 2I] = (dgate | dup): the dgate tile of (tr, tc) lands at columns col0, the dup tile at I + col0 (I % 128 == 0,
// so both start on a scale-tile boundary).  ~179 KB of shared memory: one CTA per SM, 16 warps.
__global__ void __launch_bounds__(AB_THREADS, 1) mx_dswiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_da, const __grid_constant__ CUtensorMap map_g,
                                                                            const __grid_constant__ CUtensorMap map_u, const __grid_constant__ CUtensorMap map_row,
                                                                            const __grid_constant__ CUtensorMap map_col, uint8_t* __restrict__ sf_row,
                                                                            uint8_t* __restrict__ sf_col, float* __restrict__ dbias_ws, int M, int I,
                                                                            int do_row, int do_col) {
    extern __shared__ uint8_t smem_raw[];
    ActBwdSmem& s = *reinterpret_cast<ActBwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = I / AT, tiles_r = (M + AT - 1) / AT;      // tiles of the [M, I] index space; the output has 2 * tiles_c tile columns
    const int ntiles = tiles_c * tiles_r;

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_da);
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        ptx::mbar_init(&s.full_da, 1);
        ptx::mbar_init(&s.full_gu, 1);
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();
    if (tid == 0 && static_cast<int>(blockIdx.x) < ntiles) {
        const int t = blockIdx.x;
        const int c = (t % tiles_c) * AT, r = (t / tiles_c) * AT;
        ptx::mbar_arrive_expect_tx(&s.full_da, AT_IN_BYTES);
        ptx::tma_load_2d(&map_da, &s.full_da, s.da, c, r);
        ptx::mbar_arrive_expect_tx(&s.full_gu, 2 * AT_IN_BYTES);
        ptx::tma_load_2d(&map_g, &s.full_gu, s.g, c, r);
        ptx::tma_load_2d(&map_u, &s.full_gu, s.u, c, r);
    }

    const f32x2 one2 = pk2(1.0f, 1.0f);
    int it = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, it++) {
        const uint32_t phase = it & 1;
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * AT, col0 = tc * AT;
        const int tn = t + gridDim.x;
        if (tid == 0) ptx::tma_store_wait_read<0>();   // the previous tile's stores no longer read the out buffers
        ptx::mbar_wait(&s.full_da, phase);
        ptx::mbar_wait(&s.full_gu, phase);

        // activation pass, in place: 2048 16-byte vectors, 4 per thread; thread -> fixed 8-column segment (tid & 15), rows i*32 + tid/16.
        // Arithmetic = dswiglu_quant_kernel<FMT_BF16> (swiglu.cu); the column sums take the fp32 values before the bf16 rounding.
        f32x2 sg[4], su[4];
#pragma unroll
        for (int j = 0; j < 4; j++) sg[j] = su[j] = pk2(0.f, 0.f);
#pragma unroll 1
        for (int i = 0; i < (AT * AT / 8) / AB_THREADS; i++) {
            const int idx = i * AB_THREADS + tid;
            const uint4 vd = reinterpret_cast<const uint4*>(s.da)[idx];
            const uint4 vg = reinterpret_cast<const uint4*>(s.g)[idx];
            const uint4 vu = reinterpret_cast<const uint4*>(s.u)[idx];
            float dy[8], g[8], up[8];
            unpack8(vd, dy);
            unpack8(vg, g);
            unpack8(vu, up);
            uint32_t og[4], ou[4];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const f32x2 g2 = pk2(g[2 * j], g[2 * j + 1]), u2 = pk2(up[2 * j], up[2 * j + 1]), dy2 = pk2(dy[2 * j], dy[2 * j + 1]);
                const f32x2 sig = sigmoid2_pinned(g[2 * j], g[2 * j + 1]);
                const f32x2 tt = fma2(g2, sub2(one2, sig), one2);
                const f32x2 dsilu = mul2(sig, tt);
                const f32x2 dg = mul2(mul2(dy2, u2), dsilu);
                const f32x2 du = mul2(dy2, mul2(g2, sig));
                sg[j] = add2(dg, sg[j]);
                su[j] = add2(du, su[j]);
                float lo, hi;
                upk2(dg, lo, hi);
                og[j] = pack_bf16x2(lo, hi);
                upk2(du, lo, hi);
                ou[j] = pack_bf16x2(lo, hi);
            }
            reinterpret_cast<uint4*>(s.g)[idx] = make_uint4(og[0], og[1], og[2], og[3]);
            reinterpret_cast<uint4*>(s.u)[idx] = make_uint4(ou[0], ou[1], ou[2], ou[3]);
        }
        if (dbias_ws) {
            // lanes l and l ^ 16 hold the same column segment (rows differ): fold them, then one row of `part` per warp
#pragma unroll
            for (int j = 0; j < 4; j++) {
                float glo, ghi, ulo, uhi;
                upk2(sg[j], glo, ghi);
                upk2(su[j], ulo, uhi);
                glo += __shfl_xor_sync(0xffffffffu, glo, 16);
                ghi += __shfl_xor_sync(0xffffffffu, ghi, 16);
                ulo += __shfl_xor_sync(0xffffffffu, ulo, 16);
                uhi += __shfl_xor_sync(0xffffffffu, uhi, 16);
                if (lane < 16) {
                    s.part[warp][lane * 8 + 2 * j] = glo;
                    s.part[warp][lane * 8 + 2 * j + 1] = ghi;
                    s.part[warp][AT + lane * 8 + 2 * j] = ulo;
                    s.part[warp][AT + lane * 8 + 2 * j + 1] = uhi;
                }
            }
        }
        __syncthreads();                               // S1: dgate / dup tiles complete, da buffer dead, out buffers free, part[] complete
        if (tid == 0 && tn < ntiles) {
            ptx::mbar_arrive_expect_tx(&s.full_da, AT_IN_BYTES);
            ptx::tma_load_2d(&map_da, &s.full_da, s.da, (tn % tiles_c) * AT, (tn / tiles_c) * AT);
        }
        if (dbias_ws && tid < 2 * AT) {
            float acc = 0.f;
#pragma unroll
            for (int w = 0; w < AB_THREADS / 32; w++) acc += s.part[w][tid];
            const int col = tid < AT ? col0 + tid : I + col0 + (tid - AT);
            atomicAdd(dbias_ws + col, acc);
        }
        {
            const int grp = tid >> 8, lt = tid & 255;          // grp 0: dgate, grp 1: dup
            const uint8_t* src = grp ? s.u : s.g;
            if (do_row) act_row_pass(src, s.out_row[grp], s.sf_row[grp], lt);
            if (do_col) act_col_pass(src, s.out_col[grp], s.sf_col[grp], lt);
        }
        ptx::fence_proxy_async();
        __syncthreads();                               // S2
        if (tid == 0) {
            if (do_row) {
                ptx::tma_store_2d(&map_row, s.out_row[0], col0, row0);
                ptx::tma_store_2d(&map_row, s.out_row[1], I + col0, row0);
            }
            if (do_col) {
                ptx::tma_store_2d(&map_col, s.out_col[0], col0, row0);
                ptx::tma_store_2d(&map_col, s.out_col[1], I + col0, row0);
            }
            ptx::tma_store_commit();
            if (tn < ntiles) {
                const int c = (tn % tiles_c) * AT, r = (tn / tiles_c) * AT;
                ptx::mbar_arrive_expect_tx(&s.full_gu, 2 * AT_IN_BYTES);
                ptx::tma_load_2d(&map_g, &s.full_gu, s.g, c, r);
                ptx::tma_load_2d(&map_u, &s.full_gu, s.u, c, r);
            }
        }
        // scale tiles: warps 1, 2 -> rowwise (dgate, dup), warps 3, 4 -> colwise (dgate, dup); one 16 B vector per lane
        if (do_row && warp >= 1 && warp <= 2) {
            const int h = warp - 1;
            const int64_t off = (static_cast<int64_t>(tr) * (2 * tiles_c) + h * tiles_c + tc) * 512;
            reinterpret_cast<uint4*>(sf_row + off)[lane] = reinterpret_cast<const uint4*>(s.sf_row[h])[lane];
        }
        if (do_col && warp >= 3 && warp <= 4) {
            const int h = warp - 3;
            const int64_t off = (static_cast<int64_t>(h * tiles_c + tc) * tiles_r + tr) * 512;
            reinterpret_cast<uint4*>(sf_col + off)[lane] = reinterpret_cast<const uint4*>(s.sf_col[h])[lane];
        }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// ------------------------------------------------------------------------------------------------ host side
int launch_f32_to_bf16(const float* in, void* out, int64_t n, cudaStream_t stream);

static int act_common_checks(const char* who, const void* fc1_out, int64_t M, int64_t I, const void* q_row, const void* sf_row, const void* q_col,
                             const void* sf_col, int64_t out_cols) {
    const bool do_row = q_row != nullptr, do_col = q_col != nullptr;
    LLMFP8_REQUIRE(do_row || do_col, "%s: nothing to produce", who);
    LLMFP8_REQUIRE(I % 32 == 0, "%s: ffn size %lld must be a multiple of 32", who, (long long)I);
    LLMFP8_REQUIRE(!do_row || sf_row, "%s: rowwise copy needs a scale buffer", who);
    LLMFP8_REQUIRE(!do_col || (M % 32 == 0 && sf_col), "%s: colwise copy needs rows %% 32 == 0 (got %lld) and a scale buffer", who, (long long)M);
    LLMFP8_REQUIRE(reinterpret_cast<uintptr_t>(fc1_out) % 16 == 0, "%s: fc1_out must be 16 B aligned", who);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(q_row) | reinterpret_cast<uintptr_t>(q_col) | reinterpret_cast<uintptr_t>(sf_row) |
                    reinterpret_cast<uintptr_t>(sf_col)) % 16 == 0, "%s: outputs must be 16 B aligned", who);
    LLMFP8_REQUIRE(M < (1ll << 31) && out_cols < (1ll << 31), "%s: dims too large", who);
    return 0;
}

// a = bf16(silu(gate) * up) of fc1_out [M, 2I] (gate | up), quantised rowwise (q_row, sf_row) and / or colwise (q_col, sf_col) as [M, I].
int launch_mx_swiglu_quantize(const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col, cudaStream_t stream) {
    if (int rc = act_common_checks("mx_swiglu_quantize", fc1_out, M, I, q_row, sf_row, q_col, sf_col, I)) return rc;
    if (M == 0 || I == 0) return 0;
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    const uint16_t* base = static_cast<const uint16_t*>(fc1_out);
    CUtensorMap m_g, m_u, m_row, m_col;
    // gate and up halves as two [M, I] views of the same rows (pitch 2I): reads past column I are zero-filled, never the other half
    int rc = make_tensor_map_2d(&m_g, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_u, base + I, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, I, M, I, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, I, M, I, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int ntiles = static_cast<int>(((M + AT - 1) / AT) * ((I + AT - 1) / AT));
    const int smem = sizeof(ActFwdSmem) + 128;
    const int cap = sm_count() * 2;
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e = cudaFuncSetAttribute(mx_swiglu_quantize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mx_swiglu_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((mx_swiglu_quantize_kernel), grid, 256, smem, stream, m_g, m_u, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, (int)M, (int)I, do_row, do_col);
    LLMFP8_LAUNCH_CHECK("mx_swiglu_quantize");
    return 0;
}

// dfc1 = bf16(dgate | dup) [M, 2I] from da [M, I] and fc1_out [M, 2I], quantised rowwise and / or colwise; dbias_ws fp32 [2I] workspace and
// dbias_bf16 [2I] output go together (both null: no bias gradient).
int launch_mx_dswiglu_quantize(const void* da, const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col,
                               float* dbias_ws, void* dbias_bf16, cudaStream_t stream) {
    if (int rc = act_common_checks("mx_dswiglu_quantize", fc1_out, M, I, q_row, sf_row, q_col, sf_col, 2 * I)) return rc;
    LLMFP8_REQUIRE(I % AT == 0, "mx_dswiglu_quantize: ffn size %lld must be a multiple of 128 (the dup half starts on a scale-tile boundary)", (long long)I);
    LLMFP8_REQUIRE(reinterpret_cast<uintptr_t>(da) % 16 == 0, "mx_dswiglu_quantize: da must be 16 B aligned");
    LLMFP8_REQUIRE((dbias_ws == nullptr) == (dbias_bf16 == nullptr), "mx_dswiglu_quantize: dbias workspace and output go together");
    if (M == 0 || I == 0) return 0;
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    if (dbias_ws) {
        cudaError_t e = cudaMemsetAsync(dbias_ws, 0, sizeof(float) * 2 * I, stream);
        if (e != cudaSuccess) { set_error("mx_dswiglu_quantize: memset: %s", cudaGetErrorString(e)); return (int)e; }
    }
    const uint16_t* base = static_cast<const uint16_t*>(fc1_out);
    CUtensorMap m_da, m_g, m_u, m_row, m_col;
    int rc = make_tensor_map_2d(&m_da, da, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, I * 2, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_g, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_u, base + I, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, AT, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int ntiles = static_cast<int>(((M + AT - 1) / AT) * (I / AT));
    const int smem = sizeof(ActBwdSmem) + 128;
    const int cap = sm_count();
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e = cudaFuncSetAttribute(mx_dswiglu_quantize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mx_dswiglu_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((mx_dswiglu_quantize_kernel), grid, AB_THREADS, smem, stream, m_da, m_g, m_u, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, dbias_ws, (int)M,
                                                                   (int)I, do_row, do_col);
    LLMFP8_LAUNCH_CHECK("mx_dswiglu_quantize");
    if (dbias_ws) return launch_f32_to_bf16(dbias_ws, dbias_bf16, 2 * I, stream);
    return 0;
}

}  // namespace llmfp8
