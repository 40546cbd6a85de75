// Activation + MXFP8 block quantiser in one pass for sm_100a (K3m / K3m-bwd) — SURVEY.md §8(f) rank 2.
//   fwd:  fc1_out bf16 [M, 2I] (gate | up)              -> a   = bf16(silu(gate) * up)  -> rowwise + colwise MXFP8 copies of a   [M, I]
//   bwd:  da bf16 [M, I], fc1_out bf16 [M, 2I]          -> dfc1 = bf16(dgate | dup)      -> rowwise + colwise MXFP8 copies of dfc1 [M, 2I],
//                                                          dbias[2I] = column sums of the fp32 (dgate | dup)
// Results are defined as mx_quantize(swiglu_bf16(fc1)) and mx_quantize(dswiglu_bf16(da, fc1)) — the two-kernel sequences of
// functional.MXLayerNormMLPFn, which tests/test_gpu_quantize.py pins bit-exactly to the oracle — so the bf16 rounding of the
// intermediate is kept; what disappears is its round trip through HBM (fwd: 4·M·I of 10.06·M·I bytes, bwd: 8·M·I of 18.1·M·I).
// Replaces TE's swiglu / dswiglu + mxfp8 cast inside LayerNormMLP(activation="swiglu") under te_llama_mxfp8.py:93.
//
// Persistent, warp-specialised CTAs (one per SM): a producer warp streams input tiles through a TMA ring, 16 consumer warps compute the
// activation in place (bf16, into the gate / up buffers), run the 1x32 and 32x1 block passes of mx_quantize.cu over those buffers, stage
// the e4m3 tiles in shared memory and TMA-store them; scale tiles (512 B, pre-swizzled for tcgen05.cp) leave with vector stores.
// Default path of functional.MXLayerNormMLPFn (LLMFP8_MX_FUSED_ACT=0 restores the two-kernel sequence).
#include <cstdlib>
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

constexpr int AT = 128;                     // tile edge
constexpr int AT_IN_BYTES = AT * AT * 2;    // 32 KB bf16 tile
constexpr int AT_OUT_BYTES = AT * AT;       // 16 KB e4m3 tile

// ------------------------------------------------------------------------------------------------ forward: SwiGLU + quantise
// Warp-specialised like the backward kernel below: a producer warp keeps a 2-deep ring of (gate | up) 128 x 128 bf16 tile pairs in flight
// with TMA; 16 consumer warps compute a = bf16(silu(gate) * up) in place (into the gate buffer), then threads 0..255 run the colwise and
// threads 256..511 the rowwise block pass over it.
constexpr int AF_CONS = 512;
constexpr int AF_THREADS = AF_CONS + 32;
constexpr int AF_STAGES = 2;
struct ActFwdStage {
    alignas(128) uint8_t g[AT_IN_BYTES];        // gate tile; overwritten in place by a = bf16(silu(gate) * up)
    uint8_t u[AT_IN_BYTES];                     // up tile
};
struct ActFwdSmem {
    ActFwdStage st[AF_STAGES];
    alignas(128) uint8_t out_row[AT_OUT_BYTES];
    alignas(128) uint8_t out_col[AT_OUT_BYTES];
    alignas(16) uint8_t sf_row[512];
    alignas(16) uint8_t sf_col[512];
    alignas(8) uint64_t full[AF_STAGES], empty[AF_STAGES];
};

template <class Put>
__device__ __forceinline__ void act_row_pass_f(const uint8_t* in, uint8_t* out_row, int t, Put&& put);
template <class Put>
__device__ __forceinline__ void act_col_pass_f(const uint8_t* in, uint8_t* out_col, int t, Put&& put);

__global__ void __launch_bounds__(AF_THREADS, 1) mx_swiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_u,
                                                                           const __grid_constant__ CUtensorMap map_row, const __grid_constant__ CUtensorMap map_col,
                                                                           uint8_t* __restrict__ sf_row, uint8_t* __restrict__ sf_col, int M, int I, int do_row,
                                                                           int do_col) {
    extern __shared__ uint8_t smem_raw[];
    ActFwdSmem& s = *reinterpret_cast<ActFwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = (I + AT - 1) / AT, tiles_r = (M + AT - 1) / AT;
    const int64_t ntiles = static_cast<int64_t>(tiles_c) * tiles_r;
    const int t0 = static_cast<int>(ntiles * blockIdx.x / gridDim.x), t1 = static_cast<int>(ntiles * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        for (int i = 0; i < AF_STAGES; i++) {
            ptx::mbar_init(&s.full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();

    if (warp == AF_CONS / 32) {
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int t = t0; t < t1; t++) {
                const int c = (t % tiles_c) * AT, r = (t / tiles_c) * AT;
                ptx::mbar_wait(&s.empty[stage], phase ^ 1);
                ptx::mbar_arrive_expect_tx(&s.full[stage], 2 * AT_IN_BYTES);
                ptx::tma_load_2d(&map_g, &s.full[stage], s.st[stage].g, c, r);
                ptx::tma_load_2d(&map_u, &s.full[stage], s.st[stage].u, c, r);
                if (++stage == AF_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    int stage = 0, phase = 0;
    for (int t = t0; t < t1; t++) {
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * AT, col0 = tc * AT;
        if (tid == 0) ptx::tma_store_wait_read<0>();   // the previous tile's stores no longer read out_row / out_col
        ptx::mbar_wait(&s.full[stage], phase);
        ActFwdStage& S = s.st[stage];

        // activation pass, in place: 2048 16-byte vectors, 4 per thread, consecutive threads on consecutive vectors.
        // Arithmetic = swiglu_quant_kernel<FMT_BF16> (swiglu.cu): a = (g * sigmoid(g)) * up, each product rounded, then RNE to bf16.
#pragma unroll 2
        for (int i = 0; i < (AT * AT / 8) / AF_CONS; i++) {
            const int idx = i * AF_CONS + tid;
            const uint4 vg = reinterpret_cast<const uint4*>(S.g)[idx];
            const uint4 vu = reinterpret_cast<const uint4*>(S.u)[idx];
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
            reinterpret_cast<uint4*>(S.g)[idx] = make_uint4(o[0], o[1], o[2], o[3]);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(AF_CONS) : "memory");      // S1: activation tile complete, out buffers and scale staging free
        if (tid < 256) {
            if (do_col) act_col_pass_f(S.g, s.out_col, tid, [&](int c, int rb, uint32_t e) { s.sf_col[(c & 31) * 16 + (c >> 5) * 4 + rb] = static_cast<uint8_t>(e); });
        } else {
            if (do_row) act_row_pass_f(S.g, s.out_row, tid - 256, [&](int r, int blk, uint32_t e) { s.sf_row[(r & 31) * 16 + (r >> 5) * 4 + blk] = static_cast<uint8_t>(e); });
        }
        ptx::fence_proxy_async();                      // generic writes (out tiles) before the async-proxy reads of the stores
        asm volatile("bar.sync 1, %0;" ::"n"(AF_CONS) : "memory");      // S2
        if (tid == 0) {
            if (do_row) ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
            if (do_col) ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
            ptx::tma_store_commit();
            ptx::mbar_arrive(&s.empty[stage]);
        }
        if (do_row && tid >= 32 && tid < 64) {
            const int64_t off = (static_cast<int64_t>(tr) * tiles_c + tc) * 512;
            reinterpret_cast<uint4*>(sf_row + off)[tid - 32] = reinterpret_cast<const uint4*>(s.sf_row)[tid - 32];
        }
        if (do_col && tid >= 64 && tid < 96) {
            const int64_t off = (static_cast<int64_t>(tc) * tiles_r + tr) * 512;
            reinterpret_cast<uint4*>(sf_col + off)[tid - 64] = reinterpret_cast<const uint4*>(s.sf_col)[tid - 64];
        }
        if (++stage == AF_STAGES) { stage = 0; phase ^= 1; }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// ------------------------------------------------------------------------------------------------ backward: dSwiGLU + quantise (+ dbias)
// Warp-specialised: one producer warp keeps a 3-deep ring of input tiles (da, gate, up: 64 x 128 bf16 each) in flight with TMA while 16
// consumer warps work — so the loads of tiles n+1, n+2 run under the arithmetic of tile n (the first version of this kernel was single
// buffered: per tile 4.3 us of transfer + 4 us of issue-bound arithmetic back to back = 0.48 of HBM bandwidth).
//   consumers, per tile:  activation pass in place (gate -> dgate, up -> dup, bf16; gate | up are adjacent, so together they form one
//   virtual 128 x 128 tile V: rows 0..63 dgate, rows 64..127 dup) -> barrier -> threads 0..255 run the 32x1 (colwise) block pass over V,
//   threads 256..511 the 1x32 (rowwise) one -> barrier -> four TMA stores (dgate half at column col0, dup half at I + col0), the ring
//   slot goes back to the producer, scale bytes go out from their staging area.
// Each CTA walks a contiguous range of tiles in COLUMN-major order, so the dbias column sums stay in registers from tile to tile and are
// folded / atomically added only when the tile column changes (2-3 times per CTA) instead of once per tile.
constexpr int ABR = 64;                          // tile rows of the [M, I] index space
constexpr int AB_CONS = 512;                     // consumer threads
constexpr int AB_THREADS = AB_CONS + 32;         // + the producer warp
constexpr int AB_STAGES = 3;
constexpr int AB_IN = ABR * AT * 2;              // 16 KB: one 64 x 128 bf16 tile
struct ActBwdStage {
    alignas(128) uint8_t g[AB_IN];               // gate -> dgate  \ contiguous: V
    uint8_t u[AB_IN];                            // up   -> dup    /
    uint8_t da[AB_IN];
};
struct ActBwdSmem {
    ActBwdStage st[AB_STAGES];
    alignas(128) uint8_t out_row[AT_OUT_BYTES];  // V as e4m3, rowwise blocks
    alignas(128) uint8_t out_col[AT_OUT_BYTES];  // V as e4m3, colwise blocks
    alignas(16) uint8_t sf_row[2][32][8];        // [half][row & 31][(row >> 5) * 4 + k-block]: 8 consecutive bytes of the swizzled scale tile
    alignas(16) uint8_t sf_col[2][AT][2];        // [half][column][row block]: 2 consecutive bytes of the swizzled scale tile
    float part[AB_CONS / 32][2 * AT];            // per-warp column sums of the fp32 dgate (0..127) and dup (128..255) values
    alignas(8) uint64_t full[AB_STAGES], empty[AB_STAGES];
};
static_assert(offsetof(ActBwdStage, u) == AB_IN && offsetof(ActBwdStage, da) == 2 * AB_IN, "gate | up must be adjacent");

__device__ __forceinline__ void bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(AB_CONS) : "memory"); }

// rowwise / colwise block passes over a 128 x 128 bf16 tile (row pitch 256 B) with the scale byte handed to `put`; arithmetic and
// shared-memory access order are act_row_pass / act_col_pass's (and mx_quantize_kernel<0>'s)
template <class Put>
__device__ __forceinline__ void act_row_pass_f(const uint8_t* in, uint8_t* out_row, int t, Put&& put) {
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
        put(r, blk, e);
    }
}
template <class Put>
__device__ __forceinline__ void act_col_pass_f(const uint8_t* in, uint8_t* out_col, int t, Put&& put) {
    const int cp = t & 63, rb = t >> 6;
    uint32_t pk[32];
#pragma unroll
    for (int r = 0; r < 32; r++) pk[r] = *reinterpret_cast<const uint32_t*>(in + (rb * 32 + r) * (AT * 2) + cp * 4);
    uint32_t mx = 0;
#pragma unroll
    for (int r = 0; r < 32; r++) {
        uint32_t m;
        asm("max.u16x2 %0, %1, %2;" : "=r"(m) : "r"(mx), "r"(pk[r] & 0x7fff7fffu));
        mx = m;
    }
    float am0 = bf16lo(mx), am1 = bf16hi(mx);
    if (((mx & 0xffffu) > 0x7f80u) || ((mx >> 16) > 0x7f80u)) {      // NaN patterns must not win (fmaxf in the oracle drops NaN): rare slow path
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
    put(cp * 2, rb, e0);
    put(cp * 2 + 1, rb, e1);
}

// Output matrix dfc1 is [M, 2I] = (dgate | dup): the dgate tile of (tr, tc) lands at columns col0, the dup tile at I + col0 (I % 128 == 0,
// so both start on a scale-tile boundary).  Row tiles are 64 rows: two of them fill one 128-row scale tile, and the row-tile count is
// rounded up to even so that every scale tile is written whole (rows past M come zero-filled from TMA and leave clipped).
__global__ void __launch_bounds__(AB_THREADS, 1) mx_dswiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_da, const __grid_constant__ CUtensorMap map_g,
                                                                            const __grid_constant__ CUtensorMap map_u, const __grid_constant__ CUtensorMap map_row,
                                                                            const __grid_constant__ CUtensorMap map_col, uint8_t* __restrict__ sf_row,
                                                                            uint8_t* __restrict__ sf_col, float* __restrict__ dbias_ws, int M, int I,
                                                                            int do_row, int do_col) {
    extern __shared__ uint8_t smem_raw[];
    ActBwdSmem& s = *reinterpret_cast<ActBwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = I / AT, tiles_r128 = (M + AT - 1) / AT, tiles_r = 2 * tiles_r128;
    const int64_t ntiles = static_cast<int64_t>(tiles_c) * tiles_r;
    const int t0 = static_cast<int>(ntiles * blockIdx.x / gridDim.x), t1 = static_cast<int>(ntiles * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_da);
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        for (int i = 0; i < AB_STAGES; i++) {
            ptx::mbar_init(&s.full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();

    if (warp == AB_CONS / 32) {
        // ================================ producer: one lane keeps the ring full
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int t = t0; t < t1; t++) {
                const int c = (t / tiles_r) * AT, r = (t % tiles_r) * ABR;
                ptx::mbar_wait(&s.empty[stage], phase ^ 1);
                ptx::mbar_arrive_expect_tx(&s.full[stage], 3 * AB_IN);
                ptx::tma_load_2d(&map_g, &s.full[stage], s.st[stage].g, c, r);
                ptx::tma_load_2d(&map_u, &s.full[stage], s.st[stage].u, c, r);
                ptx::tma_load_2d(&map_da, &s.full[stage], s.st[stage].da, c, r);
                if (++stage == AB_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    // ==================================== consumers
    const f32x2 one2 = pk2(1.0f, 1.0f), zero2 = pk2(0.f, 0.f);
    f32x2 sg[4], su[4];
#pragma unroll
    for (int j = 0; j < 4; j++) sg[j] = su[j] = zero2;
    auto flush_dbias = [&](int tcf) {
        // lanes l and l ^ 16 hold the same 8-column segment (rows differ): fold them, one row of `part` per warp, then 256 threads add up
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
            sg[j] = su[j] = zero2;
        }
        bar_consumers();
        if (tid < 2 * AT) {
            float acc = 0.f;
#pragma unroll
            for (int w = 0; w < AB_CONS / 32; w++) acc += s.part[w][tid];
            const int col = tid < AT ? tcf * AT + tid : I + tcf * AT + (tid - AT);
            atomicAdd(dbias_ws + col, acc);
        }
        bar_consumers();
    };

    int stage = 0, phase = 0, cur_tc = -1;
    for (int t = t0; t < t1; t++) {
        const int tc = t / tiles_r, tr = t % tiles_r;
        const int row0 = tr * ABR, col0 = tc * AT;
        if (dbias_ws && tc != cur_tc) {
            if (cur_tc >= 0) flush_dbias(cur_tc);
            cur_tc = tc;
        }
        if (tid == 0) ptx::tma_store_wait_read<0>();   // the previous tile's stores no longer read the out buffers
        ptx::mbar_wait(&s.full[stage], phase);
        ActBwdStage& S = s.st[stage];

        // activation pass, in place: 1024 16-byte vectors, 2 per thread; thread -> fixed 8-column segment (tid & 15), rows tid / 16 and + 32.
        // Arithmetic = dswiglu_quant_kernel<FMT_BF16> (swiglu.cu); the column sums take the fp32 values before the bf16 rounding.
#pragma unroll
        for (int i = 0; i < (ABR * AT / 8) / AB_CONS; i++) {
            const int idx = i * AB_CONS + tid;
            const uint4 vd = reinterpret_cast<const uint4*>(S.da)[idx];
            const uint4 vg = reinterpret_cast<const uint4*>(S.g)[idx];
            const uint4 vu = reinterpret_cast<const uint4*>(S.u)[idx];
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
            reinterpret_cast<uint4*>(S.g)[idx] = make_uint4(og[0], og[1], og[2], og[3]);
            reinterpret_cast<uint4*>(S.u)[idx] = make_uint4(ou[0], ou[1], ou[2], ou[3]);
        }
        bar_consumers();                               // S1: V complete, out buffers and scale staging free
        if (tid < 256) {
            if (do_col) act_col_pass_f(S.g, s.out_col, tid, [&](int c, int rb, uint32_t e) { s.sf_col[rb >> 1][c][rb & 1] = static_cast<uint8_t>(e); });
        } else {
            if (do_row) act_row_pass_f(S.g, s.out_row, tid - 256, [&](int r, int blk, uint32_t e) {
                s.sf_row[r >> 6][r & 31][((r >> 5) & 1) * 4 + blk] = static_cast<uint8_t>(e);
            });
        }
        ptx::fence_proxy_async();                      // generic writes (out tiles) before the async-proxy reads of the stores
        bar_consumers();                               // S2: out tiles and scale staging complete, the ring slot is dead
        if (tid == 0) {
            if (do_row) {
                ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
                ptx::tma_store_2d(&map_row, s.out_row + ABR * AT, I + col0, row0);
            }
            if (do_col) {
                ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
                ptx::tma_store_2d(&map_col, s.out_col + ABR * AT, I + col0, row0);
            }
            ptx::tma_store_commit();
            ptx::mbar_arrive(&s.empty[stage]);
        }
        // scale bytes: this 64-row tile fills half of a 128-row scale tile (row groups 0-1 or 2-3 of the swizzled layout)
        const int upper = (row0 & 64) ? 1 : 0;
        if (do_row && tid >= 32 && tid < 96) {
            const int j = tid - 32, h = j >> 5, q = j & 31;
            const int64_t off = (static_cast<int64_t>(row0 >> 7) * (2 * tiles_c) + h * tiles_c + tc) * 512 + q * 16 + upper * 8;
            *reinterpret_cast<uint2*>(sf_row + off) = *reinterpret_cast<const uint2*>(s.sf_row[h][q]);
        }
        if (do_col && tid >= 256) {
            const int j = tid - 256, h = j >> 7, c = j & 127;
            const int64_t off = (static_cast<int64_t>(h * tiles_c + tc) * tiles_r128 + (row0 >> 7)) * 512 + (c & 31) * 16 + (c >> 5) * 4 + upper * 2;
            *reinterpret_cast<uint16_t*>(sf_col + off) = *reinterpret_cast<const uint16_t*>(s.sf_col[h][c]);
        }
        if (++stage == AB_STAGES) { stage = 0; phase ^= 1; }
    }
    if (dbias_ws && cur_tc >= 0) flush_dbias(cur_tc);
    if (tid == 0) ptx::tma_store_wait<0>();
}

// ------------------------------------------------------------------------------------------------ host side
int launch_f32_to_bf16(const float* in, void* out, int64_t n, cudaStream_t stream);
// mx_act_quantize2.cu: the two-CTAs-per-SM form of both kernels (same bytes).  A/B switch for one measurement: LLMFP8_ACT_V1=1 keeps this file's kernels.
int launch_mx_swiglu_quantize2(const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col, cudaStream_t stream);
int launch_mx_dswiglu_quantize2(const void* da, const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col,
                                float* dbias_ws, cudaStream_t stream);
static bool act_v1() {
    static const bool v1 = [] { const char* e = getenv("LLMFP8_ACT_V1"); return e && e[0] == '1'; }();
    return v1;
}

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
    if (!act_v1()) return launch_mx_swiglu_quantize2(fc1_out, M, I, q_row, sf_row, q_col, sf_col, stream);
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
    const int cap = sm_count();
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e = cudaFuncSetAttribute(mx_swiglu_quantize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mx_swiglu_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((mx_swiglu_quantize_kernel), grid, AF_THREADS, smem, stream, m_g, m_u, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, (int)M, (int)I, do_row, do_col);
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
    if (!act_v1()) {
        if (int rc = launch_mx_dswiglu_quantize2(da, fc1_out, M, I, q_row, sf_row, q_col, sf_col, dbias_ws, stream)) return rc;
        if (dbias_ws) return launch_f32_to_bf16(dbias_ws, dbias_bf16, 2 * I, stream);
        return 0;
    }
    const uint16_t* base = static_cast<const uint16_t*>(fc1_out);
    CUtensorMap m_da, m_g, m_u, m_row, m_col;
    int rc = make_tensor_map_2d(&m_da, da, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, I * 2, AT, ABR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_g, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, ABR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_u, base + I, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, ABR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, ABR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, ABR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int64_t ntiles64 = 2 * ((M + AT - 1) / AT) * (I / AT);
    const int ntiles = ntiles64 > (1 << 30) ? (1 << 30) : static_cast<int>(ntiles64);
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
