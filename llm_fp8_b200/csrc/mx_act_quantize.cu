// Activation + MXFP8 block quantiser in one pass for sm_100a (K3m / K3m-bwd) — SURVEY.md §8(f) rank 2.
//   fwd:  fc1_out bf16 [M, 2I] (gate | up)              -> a   = bf16(silu(gate) * up)  -> rowwise + colwise MXFP8 copies of a   [M, I]
//   bwd:  da bf16 [M, I], fc1_out bf16 [M, 2I]          -> dfc1 = bf16(dgate | dup)      -> rowwise + colwise MXFP8 copies of dfc1 [M, 2I],
//                                                          dbias[2I] = column sums of the fp32 (dgate | dup)
// Results are defined as mx_quantize(swiglu_bf16(fc1)) and mx_quantize(dswiglu_bf16(da, fc1)) — the two-kernel sequences that
// tests/test_gpu_quantize.py pins bit-exactly to the oracle — so the bf16 rounding of the intermediate is kept; what disappears is its
// round trip through HBM (fwd: 4·M·I of 10.06·M·I bytes, bwd: 8·M·I of 18.1·M·I).
// Replaces TE's swiglu / dswiglu + mxfp8 cast inside LayerNormMLP(activation="swiglu") under te_llama_mxfp8.py:93.
//
// Persistent, warp-specialised CTAs, TWO per SM: a producer warp streams input tiles through a TMA ring, 8 consumer warps compute the
// activation in place (bf16, into the gate / up buffers), run the 1x32 and 32x1 block passes over those buffers (128 threads each), stage
// the e4m3 tiles in shared memory and TMA-store them; scale bytes (pre-swizzled for tcgen05.cp) leave with vector stores.
// The first form of these kernels ran one 544-thread CTA per SM on 128-row tiles: ncu showed it issue-bound in PHASES (all 16 consumer
// warps in the activation pass — FMA / XU pipes, the pinned sigmoid — then all in the block passes — ALU / LSU — with two CTA-wide barriers
// per tile; `math_pipe_throttle` and `not_selected` the top stalls at 62 % issue utilisation).  With half the consumers and half the tile
// per CTA two CTAs share an SM and drift apart: measured 114.8 -> 108.5 us (fwd) and 189.6 -> 185.3 us (bwd) at the Llama-3.1-8B shape,
// same bytes (profiles/r02/NOTES.md, call 22).
//   fwd tile: 64 x 128 of [M, I]; the e4m3 tiles are staged IN PLACE over the `up` buffer (dead after the activation pass), so a ring slot
//             is 32 KB and three of them fit twice per SM; the slot goes back to the producer once its stores have been read.
//   bwd tile: 32 x 128 of [M, I]; gate | up adjacent = the virtual tile V (rows 0..31 dgate, 32..63 dup), separate 2 x 8 KB out buffers.
// A 64-row (32-row) tile fills half (a quarter) of a 128-row swizzled scale tile; row-tile counts are rounded up to whole scale tiles.
// Default path of functional.MXLayerNormMLPFn (LLMFP8_MX_FUSED_ACT=0 restores the two-kernel sequence).
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

namespace act2 {
constexpr int AT = 128;                     // tile columns
constexpr int VR = 64;                      // rows of the tile the block passes run over
constexpr int CONS = 256;                   // consumer threads
constexpr int THREADS = CONS + 32;          // + the producer warp
constexpr int PASS_T = CONS / 2;            // threads per block pass
constexpr int V_OUT = VR * AT;              // 8 KB e4m3 tile

__device__ __forceinline__ void bar_consumers() { asm volatile("bar.sync 1, %0;" ::"n"(CONS) : "memory"); }

// rowwise (1x32) block pass over a 64 x 128 bf16 tile (row pitch 256 B) by 128 threads: one block per thread and iteration, its four 16 B
// segments read in a lane-rotated order (bank-conflict free).  Arithmetic = mx_quantize_kernel<0>.
template <class Put>
__device__ __forceinline__ void row_pass(const uint8_t* in, uint8_t* out_row, int t, Put&& put) {
#pragma unroll 1
    for (int i = 0; i < (VR * AT / 32) / PASS_T; i++) {
        const int idx = i * PASS_T + t;
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
// colwise (32x1) block pass by 128 threads: thread = 2 adjacent columns x one 32-row block
template <class Put>
__device__ __forceinline__ void col_pass(const uint8_t* in, uint8_t* out_col, int t, Put&& put) {
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

// ------------------------------------------------------------------------------------------------ forward
constexpr int F_STAGES = 3;
constexpr int F_IN = VR * AT * 2;               // 16 KB: 64 x 128 bf16
struct FwdStage {
    alignas(128) uint8_t g[F_IN];               // gate; overwritten in place by a = bf16(silu(gate) * up)
    uint8_t u[F_IN];                            // up; after the activation pass: out_row (8 KB) | out_col (8 KB)
};
struct FwdSmem {
    FwdStage st[F_STAGES];
    alignas(16) uint8_t sf_row[32][8];          // [row & 31][(row >> 5) * 4 + k-block]: 8 consecutive bytes of the swizzled scale tile
    alignas(16) uint8_t sf_col[AT][2];          // [column][row block]: 2 consecutive bytes
    alignas(8) uint64_t full[F_STAGES], empty[F_STAGES];
};

__global__ void __launch_bounds__(THREADS, 2) mx_swiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_g, const __grid_constant__ CUtensorMap map_u,
                                                                         const __grid_constant__ CUtensorMap map_row, const __grid_constant__ CUtensorMap map_col,
                                                                         uint8_t* __restrict__ sf_row, uint8_t* __restrict__ sf_col, int M, int I, int do_row,
                                                                         int do_col) {
    extern __shared__ uint8_t smem_raw[];
    FwdSmem& s = *reinterpret_cast<FwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = (I + AT - 1) / AT, tiles_r128 = (M + AT - 1) / AT, tiles_r = 2 * tiles_r128;
    const int64_t ntiles = static_cast<int64_t>(tiles_c) * tiles_r;
    const int t0 = static_cast<int>(ntiles * blockIdx.x / gridDim.x), t1 = static_cast<int>(ntiles * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        for (int i = 0; i < F_STAGES; i++) {
            ptx::mbar_init(&s.full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();

    if (warp == CONS / 32) {
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int t = t0; t < t1; t++) {
                const int c = (t % tiles_c) * AT, r = (t / tiles_c) * VR;
                ptx::mbar_wait(&s.empty[stage], phase ^ 1);
                ptx::mbar_arrive_expect_tx(&s.full[stage], 2 * F_IN);
                ptx::tma_load_2d(&map_g, &s.full[stage], s.st[stage].g, c, r);
                ptx::tma_load_2d(&map_u, &s.full[stage], s.st[stage].u, c, r);
                if (++stage == F_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    int stage = 0, phase = 0, draining = -1;        // draining: the slot whose e4m3 tiles are being read by the previous tile's TMA stores
    for (int t = t0; t < t1; t++) {
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * VR, col0 = tc * AT;
        if (tid == 0 && draining >= 0) {
            ptx::tma_store_wait_read<0>();
            ptx::mbar_arrive(&s.empty[draining]);
        }
        ptx::mbar_wait(&s.full[stage], phase);
        FwdStage& S = s.st[stage];

        // activation pass, in place: 1024 16-byte vectors, 4 per thread.  Arithmetic = swiglu_quant_kernel<FMT_BF16> (swiglu.cu).
#pragma unroll 2
        for (int i = 0; i < (VR * AT / 8) / CONS; i++) {
            const int idx = i * CONS + tid;
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
        bar_consumers();                               // S1: activation tile complete, `up` is dead, the scale staging of the previous tile has left
        uint8_t* out_row = S.u;
        uint8_t* out_col = S.u + V_OUT;
        if (tid < PASS_T) {
            if (do_col) col_pass(S.g, out_col, tid, [&](int c, int rb, uint32_t e) { s.sf_col[c][rb] = static_cast<uint8_t>(e); });
        } else {
            if (do_row) row_pass(S.g, out_row, tid - PASS_T, [&](int r, int blk, uint32_t e) { s.sf_row[r & 31][(r >> 5) * 4 + blk] = static_cast<uint8_t>(e); });
        }
        ptx::fence_proxy_async();                      // generic writes (out tiles) before the async-proxy reads of the stores
        bar_consumers();                               // S2
        if (tid == 0) {
            if (do_row) ptx::tma_store_2d(&map_row, out_row, col0, row0);
            if (do_col) ptx::tma_store_2d(&map_col, out_col, col0, row0);
            ptx::tma_store_commit();
        }
        draining = stage;
        // scale bytes: this 64-row tile fills half of a 128-row scale tile (row groups 0-1 or 2-3 of the swizzled layout)
        const int upper = (row0 & 64) ? 1 : 0;
        if (do_row && tid >= 32 && tid < 64) {
            const int q = tid - 32;
            const int64_t off = (static_cast<int64_t>(row0 >> 7) * tiles_c + tc) * 512 + q * 16 + upper * 8;
            *reinterpret_cast<uint2*>(sf_row + off) = *reinterpret_cast<const uint2*>(s.sf_row[q]);
        }
        if (do_col && tid >= PASS_T) {
            const int c = tid - PASS_T;
            const int64_t off = (static_cast<int64_t>(tc) * tiles_r128 + (row0 >> 7)) * 512 + (c & 31) * 16 + (c >> 5) * 4 + upper * 2;
            *reinterpret_cast<uint16_t*>(sf_col + off) = *reinterpret_cast<const uint16_t*>(s.sf_col[c]);
        }
        if (++stage == F_STAGES) { stage = 0; phase ^= 1; }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// ------------------------------------------------------------------------------------------------ backward
constexpr int BR = 32;                           // tile rows of the [M, I] index space
constexpr int B_STAGES = 3;
constexpr int B_IN = BR * AT * 2;                // 8 KB: 32 x 128 bf16
struct BwdStage {
    alignas(128) uint8_t g[B_IN];                // gate -> dgate  \ contiguous: V
    uint8_t u[B_IN];                             // up   -> dup    /
    uint8_t da[B_IN];
};
struct BwdSmem {
    BwdStage st[B_STAGES];
    alignas(128) uint8_t out_row[V_OUT];         // V as e4m3, rowwise blocks
    alignas(128) uint8_t out_col[V_OUT];         // V as e4m3, colwise blocks
    alignas(16) uint8_t sf_row[2][32][4];        // [half][row & 31][k-block]: 4 consecutive bytes of the swizzled scale tile
    alignas(16) uint8_t sf_col[2][AT];           // [half][column]: one byte each
    float part[CONS / 32][2 * AT];               // per-warp column sums of the fp32 dgate (0..127) and dup (128..255) values
    alignas(8) uint64_t full[B_STAGES], empty[B_STAGES];
};
static_assert(offsetof(BwdStage, u) == B_IN && offsetof(BwdStage, da) == 2 * B_IN, "gate | up must be adjacent");

// Output matrix dfc1 is [M, 2I] = (dgate | dup): the dgate tile of (tr, tc) lands at columns col0, the dup tile at I + col0 (I % 128 == 0).
// Each CTA walks a contiguous range of tiles in COLUMN-major order, so the dbias column sums stay in registers from tile to tile and are
// folded / atomically added only when the tile column changes.
__global__ void __launch_bounds__(THREADS, 2) mx_dswiglu_quantize_kernel(const __grid_constant__ CUtensorMap map_da, const __grid_constant__ CUtensorMap map_g,
                                                                          const __grid_constant__ CUtensorMap map_u, const __grid_constant__ CUtensorMap map_row,
                                                                          const __grid_constant__ CUtensorMap map_col, uint8_t* __restrict__ sf_row,
                                                                          uint8_t* __restrict__ sf_col, float* __restrict__ dbias_ws, int M, int I,
                                                                          int do_row, int do_col) {
    extern __shared__ uint8_t smem_raw[];
    BwdSmem& s = *reinterpret_cast<BwdSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = I / AT, tiles_r128 = (M + AT - 1) / AT, tiles_r = 4 * tiles_r128;
    const int64_t ntiles = static_cast<int64_t>(tiles_c) * tiles_r;
    const int t0 = static_cast<int>(ntiles * blockIdx.x / gridDim.x), t1 = static_cast<int>(ntiles * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_da);
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_u);
        for (int i = 0; i < B_STAGES; i++) {
            ptx::mbar_init(&s.full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): the prologue above ran under the previous kernel's tail
    pdl_launch_dependents();

    if (warp == CONS / 32) {
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int t = t0; t < t1; t++) {
                const int c = (t / tiles_r) * AT, r = (t % tiles_r) * BR;
                ptx::mbar_wait(&s.empty[stage], phase ^ 1);
                ptx::mbar_arrive_expect_tx(&s.full[stage], 3 * B_IN);
                ptx::tma_load_2d(&map_g, &s.full[stage], s.st[stage].g, c, r);
                ptx::tma_load_2d(&map_u, &s.full[stage], s.st[stage].u, c, r);
                ptx::tma_load_2d(&map_da, &s.full[stage], s.st[stage].da, c, r);
                if (++stage == B_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

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
        {
            float acc = 0.f;
#pragma unroll
            for (int w = 0; w < CONS / 32; w++) acc += s.part[w][tid];
            const int col = tid < AT ? tcf * AT + tid : I + tcf * AT + (tid - AT);
            atomicAdd(dbias_ws + col, acc);
        }
        bar_consumers();
    };

    int stage = 0, phase = 0, cur_tc = -1;
    for (int t = t0; t < t1; t++) {
        const int tc = t / tiles_r, tr = t % tiles_r;
        const int row0 = tr * BR, col0 = tc * AT;
        if (dbias_ws && tc != cur_tc) {
            if (cur_tc >= 0) flush_dbias(cur_tc);
            cur_tc = tc;
        }
        if (tid == 0) ptx::tma_store_wait_read<0>();   // the previous tile's stores no longer read the out buffers
        ptx::mbar_wait(&s.full[stage], phase);
        BwdStage& S = s.st[stage];

        // activation pass, in place: 512 16-byte vectors, 2 per thread; thread -> fixed 8-column segment (tid & 15), rows tid / 16 and + 16.
        // Arithmetic = dswiglu_quant_kernel<FMT_BF16> (swiglu.cu); the column sums take the fp32 values before the bf16 rounding.
#pragma unroll
        for (int i = 0; i < (BR * AT / 8) / CONS; i++) {
            const int idx = i * CONS + tid;
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
        if (tid < PASS_T) {
            if (do_col) col_pass(S.g, s.out_col, tid, [&](int c, int rb, uint32_t e) { s.sf_col[rb][c] = static_cast<uint8_t>(e); });
        } else {
            if (do_row) row_pass(S.g, s.out_row, tid - PASS_T, [&](int r, int blk, uint32_t e) { s.sf_row[r >> 5][r & 31][blk] = static_cast<uint8_t>(e); });
        }
        ptx::fence_proxy_async();                      // generic writes (out tiles) before the async-proxy reads of the stores
        bar_consumers();                               // S2: out tiles and scale staging complete, the ring slot is dead
        if (tid == 0) {
            if (do_row) {
                ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
                ptx::tma_store_2d(&map_row, s.out_row + BR * AT, I + col0, row0);
            }
            if (do_col) {
                ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
                ptx::tma_store_2d(&map_col, s.out_col + BR * AT, I + col0, row0);
            }
            ptx::tma_store_commit();
            ptx::mbar_arrive(&s.empty[stage]);
        }
        // scale bytes: this 32-row tile fills ONE row group (quarter) of a 128-row scale tile
        const int quarter = (row0 >> 5) & 3;
        if (do_row && tid >= 32 && tid < 96) {
            const int j = tid - 32, h = j >> 5, q = j & 31;
            const int64_t off = (static_cast<int64_t>(row0 >> 7) * (2 * tiles_c) + h * tiles_c + tc) * 512 + q * 16 + quarter * 4;
            *reinterpret_cast<uint32_t*>(sf_row + off) = *reinterpret_cast<const uint32_t*>(s.sf_row[h][q]);
        }
        if (do_col) {
            const int h = tid >> 7, c = tid & 127;
            const int64_t off = (static_cast<int64_t>(h * tiles_c + tc) * tiles_r128 + (row0 >> 7)) * 512 + (c & 31) * 16 + (c >> 5) * 4 + quarter;
            sf_col[off] = s.sf_col[h][c];
        }
        if (++stage == B_STAGES) { stage = 0; phase ^= 1; }
    }
    if (dbias_ws && cur_tc >= 0) flush_dbias(cur_tc);
    if (tid == 0) ptx::tma_store_wait<0>();
}
}  // namespace act2

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
    using namespace act2;
    if (int rc = act_common_checks("mx_swiglu_quantize", fc1_out, M, I, q_row, sf_row, q_col, sf_col, I)) return rc;
    if (M == 0 || I == 0) return 0;
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    const uint16_t* base = static_cast<const uint16_t*>(fc1_out);
    CUtensorMap m_g, m_u, m_row, m_col;
    // gate and up halves as two [M, I] views of the same rows (pitch 2I): reads past column I are zero-filled, never the other half
    int rc = make_tensor_map_2d(&m_g, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, VR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_u, base + I, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, VR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, I, M, I, AT, VR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, I, M, I, AT, VR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int64_t ntiles64 = 2 * ((M + AT - 1) / AT) * ((I + AT - 1) / AT);
    const int ntiles = ntiles64 > (1 << 30) ? (1 << 30) : static_cast<int>(ntiles64);
    const int smem = sizeof(FwdSmem) + 128;
    const int cap = 2 * sm_count();
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e = cudaFuncSetAttribute(mx_swiglu_quantize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mx_swiglu_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((mx_swiglu_quantize_kernel), grid, THREADS, smem, stream, m_g, m_u, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, (int)M, (int)I, do_row, do_col);
    LLMFP8_LAUNCH_CHECK("mx_swiglu_quantize");
    return 0;
}

// dfc1 = bf16(dgate | dup) [M, 2I] from da [M, I] and fc1_out [M, 2I], quantised rowwise and / or colwise; dbias_ws fp32 [2I] workspace and
// dbias_bf16 [2I] output go together (both null: no bias gradient).
int launch_mx_dswiglu_quantize(const void* da, const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col,
                               float* dbias_ws, void* dbias_bf16, cudaStream_t stream) {
    using namespace act2;
    if (int rc = act_common_checks("mx_dswiglu_quantize", fc1_out, M, I, q_row, sf_row, q_col, sf_col, 2 * I)) return rc;
    LLMFP8_REQUIRE(I % AT == 0, "mx_dswiglu_quantize: ffn size %lld must be a multiple of 128 (the dup half starts on a scale-tile boundary)", (long long)I);
    LLMFP8_REQUIRE(reinterpret_cast<uintptr_t>(da) % 16 == 0, "mx_dswiglu_quantize: da must be 16 B aligned");
    LLMFP8_REQUIRE((dbias_ws == nullptr) == (dbias_bf16 == nullptr), "mx_dswiglu_quantize: dbias workspace and output go together");
    if (M == 0 || I == 0) return 0;
    if (dbias_ws) {
        cudaError_t e0 = cudaMemsetAsync(dbias_ws, 0, sizeof(float) * 2 * I, stream);
        if (e0 != cudaSuccess) { set_error("mx_dswiglu_quantize: memset: %s", cudaGetErrorString(e0)); return (int)e0; }
    }
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    const uint16_t* base = static_cast<const uint16_t*>(fc1_out);
    CUtensorMap m_da, m_g, m_u, m_row, m_col;
    int rc = make_tensor_map_2d(&m_da, da, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, I * 2, AT, BR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_g, base, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, BR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_u, base + I, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, I, M, 2 * I * 2, AT, BR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_row, do_row ? q_row : q_col, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, BR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&m_col, do_col ? q_col : q_row, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, 2 * I, M, 2 * I, AT, BR, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    const int64_t ntiles64 = 4 * ((M + AT - 1) / AT) * (I / AT);
    const int ntiles = ntiles64 > (1 << 30) ? (1 << 30) : static_cast<int>(ntiles64);
    const int smem = sizeof(BwdSmem) + 128;
    const int cap = 2 * sm_count();
    const int grid = ntiles < cap ? ntiles : cap;
    cudaError_t e = cudaFuncSetAttribute(mx_dswiglu_quantize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("mx_dswiglu_quantize: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((mx_dswiglu_quantize_kernel), grid, THREADS, smem, stream, m_da, m_g, m_u, m_row, m_col, (uint8_t*)sf_row, (uint8_t*)sf_col, dbias_ws, (int)M,
                                                                    (int)I, do_row, do_col);
    LLMFP8_LAUNCH_CHECK("mx_dswiglu_quantize");
    if (dbias_ws) return launch_f32_to_bf16(dbias_ws, dbias_bf16, 2 * I, stream);
    return 0;
}

}  // namespace llmfp8
