// Fused AdamW + MXFP8 requantise for sm_100a ("quantize W once per optimizer step, not per forward").
// The reference steps torch.optim.AdamW over its bf16 parameters every step (train_fp8.py:200-204, 288-291) and TE then re-casts every weight in
// the next forward.  Here the optimizer's own pass over a GEMM weight also emits the weight's MXFP8 copies (rowwise for fprop, colwise for
// dgrad; E8M0 scales pre-swizzled) from the tile it already holds in shared memory, so the casts — 148 us per decoder layer at Llama-3.1-8B,
// 3.5 % of the step — leave the forward:  per parameter element 8 B read (w, g, m, v in bf16) + 6 B written (w, m, v) + 2.06 B of copies, against
// a separate 2 B re-read + 2.06 B for the cast.
//   m  = beta1 m + (1 - beta1) g            v = beta2 v + (1 - beta2) g^2                  (fp32 arithmetic, bf16 state as torch keeps it for
//   w  = w (1 - lr wd) - (lr / bc1) m / (sqrt(v) / sqrt(bc2) + eps)                          bf16 parameters; decoupled weight decay)
//   copies = mx_quantize(bf16(w))  — bit-identical to llmfp8_mx_quantize of the updated weight (same block passes as mx_act_quantize.cu).
// Warp-specialised like the fused activation kernels: a producer warp streams (w, g, m, v) 64 x 128 tiles through a 2-deep TMA ring, 8 consumer
// warps update in place, run the block passes, and TMA-store w, m, v and the e4m3 tiles.
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

namespace {

constexpr int OT = 128;                      // tile columns
constexpr int OR_ = 64;                      // tile rows
constexpr int O_IN = OR_ * OT * 2;           // 16 KB: one bf16 tile
constexpr int O_CONS = 256;
constexpr int O_THREADS = O_CONS + 32;
constexpr int O_STAGES = 2;

struct OptStage {
    alignas(128) uint8_t w[O_IN];
    uint8_t g[O_IN];
    uint8_t m[O_IN];
    uint8_t v[O_IN];
};
struct OptSmem {
    OptStage st[O_STAGES];
    alignas(128) uint8_t out_row[OR_ * OT];
    alignas(128) uint8_t out_col[OR_ * OT];
    alignas(16) uint8_t sf_row[32][8];       // [row & 31][(row >> 5) * 4 + k-block]
    alignas(16) uint8_t sf_col[OT][2];       // [column][row block]
    alignas(8) uint64_t full[O_STAGES], empty[O_STAGES];
};

struct AdamArgs {
    float lr_wd;        // 1 - lr * weight_decay
    float step_size;    // lr / bias_correction1
    float beta1, beta2, one_m_beta1, one_m_beta2;
    float inv_sqrt_bc2; // 1 / sqrt(bias_correction2)
    float eps;
};

__device__ __forceinline__ void bar_cons() { asm volatile("bar.sync 1, %0;" ::"n"(O_CONS) : "memory"); }

}  // namespace

__global__ void __launch_bounds__(O_THREADS, 1) adamw_mx_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_g,
                                                                const __grid_constant__ CUtensorMap map_m, const __grid_constant__ CUtensorMap map_v,
                                                                const __grid_constant__ CUtensorMap map_row, const __grid_constant__ CUtensorMap map_col,
                                                                uint8_t* __restrict__ sf_row, uint8_t* __restrict__ sf_col, int R, int C, int do_row,
                                                                int do_col, const AdamArgs a) {
    extern __shared__ uint8_t smem_raw[];
    OptSmem& s = *reinterpret_cast<OptSmem*>(smem_raw + ((128u - (ptx::smem_u32(smem_raw) & 127u)) & 127u));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tiles_c = (C + OT - 1) / OT, tiles_r128 = (R + 127) / 128, tiles_r = 2 * tiles_r128;   // 64-row tiles, whole 128-row scale tiles
    const int64_t ntiles = static_cast<int64_t>(tiles_c) * tiles_r;
    const int t0 = static_cast<int>(ntiles * blockIdx.x / gridDim.x), t1 = static_cast<int>(ntiles * (blockIdx.x + 1) / gridDim.x);

    if (tid == 0) {
        ptx::prefetch_tensormap(&map_w);
        ptx::prefetch_tensormap(&map_g);
        ptx::prefetch_tensormap(&map_m);
        ptx::prefetch_tensormap(&map_v);
        for (int i = 0; i < O_STAGES; i++) {
            ptx::mbar_init(&s.full[i], 1);
            ptx::mbar_init(&s.empty[i], 1);
        }
        ptx::fence_barrier_init();
    }
    __syncthreads();
    pdl_wait();
    pdl_launch_dependents();

    if (warp == O_CONS / 32) {
        if (lane == 0) {
            int stage = 0, phase = 0;
            for (int t = t0; t < t1; t++) {
                const int c = (t % tiles_c) * OT, r = (t / tiles_c) * OR_;
                ptx::mbar_wait(&s.empty[stage], phase ^ 1);
                ptx::mbar_arrive_expect_tx(&s.full[stage], 4 * O_IN);
                ptx::tma_load_2d(&map_w, &s.full[stage], s.st[stage].w, c, r);
                ptx::tma_load_2d(&map_g, &s.full[stage], s.st[stage].g, c, r);
                ptx::tma_load_2d(&map_m, &s.full[stage], s.st[stage].m, c, r);
                ptx::tma_load_2d(&map_v, &s.full[stage], s.st[stage].v, c, r);
                if (++stage == O_STAGES) { stage = 0; phase ^= 1; }
            }
        }
        return;
    }

    int stage = 0, phase = 0, prev_stage = -1;
    for (int t = t0; t < t1; t++) {
        const int tr = t / tiles_c, tc = t % tiles_c;
        const int row0 = tr * OR_, col0 = tc * OT;
        if (tid == 0) {
            // the previous tile's stores read its ring slot (w, m, v are updated in place) and the out tiles: once they have been read the slot
            // goes back to the producer and the out tiles may be overwritten
            ptx::tma_store_wait_read<0>();
            if (prev_stage >= 0) ptx::mbar_arrive(&s.empty[prev_stage]);
        }
        ptx::mbar_wait(&s.full[stage], phase);
        OptStage& S = s.st[stage];

        // update pass, in place: 1024 16-byte vectors per tensor, 4 per thread
#pragma unroll 2
        for (int i = 0; i < (OR_ * OT / 8) / O_CONS; i++) {
            const int idx = i * O_CONS + tid;
            float w[8], g[8], m[8], v[8];
            unpack8(reinterpret_cast<const uint4*>(S.w)[idx], w);
            unpack8(reinterpret_cast<const uint4*>(S.g)[idx], g);
            unpack8(reinterpret_cast<const uint4*>(S.m)[idx], m);
            unpack8(reinterpret_cast<const uint4*>(S.v)[idx], v);
#pragma unroll
            for (int j = 0; j < 8; j++) {
                m[j] = __fmaf_rn(a.beta1, m[j], __fmul_rn(a.one_m_beta1, g[j]));
                v[j] = __fmaf_rn(a.beta2, v[j], __fmul_rn(a.one_m_beta2, __fmul_rn(g[j], g[j])));
                const float denom = __fmaf_rn(__fsqrt_rn(v[j]), a.inv_sqrt_bc2, a.eps);
                w[j] = __fmaf_rn(-a.step_size, __fdiv_rn(m[j], denom), __fmul_rn(w[j], a.lr_wd));
            }
            reinterpret_cast<uint4*>(S.w)[idx] = make_uint4(pack_bf16x2(w[0], w[1]), pack_bf16x2(w[2], w[3]), pack_bf16x2(w[4], w[5]), pack_bf16x2(w[6], w[7]));
            reinterpret_cast<uint4*>(S.m)[idx] = make_uint4(pack_bf16x2(m[0], m[1]), pack_bf16x2(m[2], m[3]), pack_bf16x2(m[4], m[5]), pack_bf16x2(m[6], m[7]));
            reinterpret_cast<uint4*>(S.v)[idx] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        }
        bar_cons();                                    // S1: the updated tile is complete; out tiles and scale staging are free (thread 0 waited above)
        const uint8_t* in = S.w;
        if (tid < 128) {
            if (do_col) {
                // thread: 2 adjacent columns x 32 rows = two 32x1 blocks (arithmetic and access order of mx_quantize_kernel<0>)
                const int cp = tid & 63, rb = tid >> 6;
                uint32_t pk[32];
#pragma unroll
                for (int r = 0; r < 32; r++) pk[r] = *reinterpret_cast<const uint32_t*>(in + (rb * 32 + r) * (OT * 2) + cp * 4);
                uint32_t mx = 0;
#pragma unroll
                for (int r = 0; r < 32; r++) {
                    uint32_t q;
                    asm("max.u16x2 %0, %1, %2;" : "=r"(q) : "r"(mx), "r"(pk[r] & 0x7fff7fffu));
                    mx = q;
                }
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
                    *reinterpret_cast<uint16_t*>(s.out_col + (rb * 32 + r) * OT + cp * 2) = static_cast<uint16_t>(cvt_fp8x2<FMT_E4M3>(lo, hi));
                }
                s.sf_col[cp * 2][rb] = static_cast<uint8_t>(e0);
                s.sf_col[cp * 2 + 1][rb] = static_cast<uint8_t>(e1);
            }
        } else if (do_row) {
            // thread: two whole 1x32 blocks
            const int lt = tid - 128;
#pragma unroll 1
            for (int i = 0; i < 2; i++) {
                const int idx = i * 128 + lt;
                const int r = idx >> 2, blk = idx & 3;
                const int rot = (lt >> 1) & 3;
                const uint8_t* src = in + r * (OT * 2) + blk * 64;
                f32x2 f[4][4];
                float am = 0.f;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int seg = (j + rot) & 3;
                    float x[8];
                    unpack8(*reinterpret_cast<const uint4*>(src + seg * 16), x);
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
                    *reinterpret_cast<uint2*>(s.out_row + r * OT + blk * 32 + seg * 8) = make_uint2(h[0] | (h[1] << 16), h[2] | (h[3] << 16));
                }
                s.sf_row[r & 31][(r >> 5) * 4 + blk] = static_cast<uint8_t>(e);
            }
        }
        ptx::fence_proxy_async();                      // generic writes (w, m, v in place; out tiles) before the async-proxy reads of the stores
        bar_cons();                                    // S2
        if (tid == 0) {
            ptx::tma_store_2d(&map_w, S.w, col0, row0);
            ptx::tma_store_2d(&map_m, S.m, col0, row0);
            ptx::tma_store_2d(&map_v, S.v, col0, row0);
            if (do_row) ptx::tma_store_2d(&map_row, s.out_row, col0, row0);
            if (do_col) ptx::tma_store_2d(&map_col, s.out_col, col0, row0);
            ptx::tma_store_commit();
        }
        const int upper = (row0 & 64) ? 1 : 0;         // this 64-row tile fills row groups 0-1 or 2-3 of its 128-row scale tile
        if (do_row && tid >= 32 && tid < 64) {
            const int q = tid - 32;
            const int64_t off = (static_cast<int64_t>(row0 >> 7) * tiles_c + tc) * 512 + q * 16 + upper * 8;
            *reinterpret_cast<uint2*>(sf_row + off) = *reinterpret_cast<const uint2*>(s.sf_row[q]);
        }
        if (do_col && tid >= 128) {
            const int c = tid - 128;
            const int64_t off = (static_cast<int64_t>(tc) * tiles_r128 + (row0 >> 7)) * 512 + (c & 31) * 16 + (c >> 5) * 4 + upper * 2;
            *reinterpret_cast<uint16_t*>(sf_col + off) = *reinterpret_cast<const uint16_t*>(s.sf_col[c]);
        }
        prev_stage = stage;
        if (++stage == O_STAGES) { stage = 0; phase ^= 1; }
    }
    if (tid == 0) ptx::tma_store_wait<0>();
}

// w, g, m, v: bf16 [R, C] contiguous (m, v: the optimizer state, updated in place like w).  q_row / q_col (+ scales): the MXFP8 copies of the
// UPDATED weight (either may be null).  bias_corr{1,2} = 1 - beta^step, computed by the host.
int launch_adamw_mx(void* w, const void* g, void* m, void* v, int64_t R, int64_t C, float lr, float beta1, float beta2, float eps, float weight_decay,
                    float bias_corr1, float bias_corr2, void* q_row, void* sf_row, void* q_col, void* sf_col, cudaStream_t stream) {
    const int do_row = q_row != nullptr, do_col = q_col != nullptr;
    LLMFP8_REQUIRE(w && g && m && v, "adamw_mx: null tensor");
    LLMFP8_REQUIRE(C % 32 == 0 && R % 32 == 0, "adamw_mx: rows %lld and cols %lld must be multiples of 32", (long long)R, (long long)C);
    LLMFP8_REQUIRE(!do_row || sf_row, "adamw_mx: rowwise copy needs a scale buffer");
    LLMFP8_REQUIRE(!do_col || sf_col, "adamw_mx: colwise copy needs a scale buffer");
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v) |
                    reinterpret_cast<uintptr_t>(q_row) | reinterpret_cast<uintptr_t>(q_col) | reinterpret_cast<uintptr_t>(sf_row) |
                    reinterpret_cast<uintptr_t>(sf_col)) % 16 == 0, "adamw_mx: tensors must be 16 B aligned");
    LLMFP8_REQUIRE(bias_corr1 > 0.f && bias_corr2 > 0.f, "adamw_mx: bias corrections must be positive (step >= 1)");
    LLMFP8_REQUIRE(R < (1ll << 31) && C < (1ll << 31), "adamw_mx: dims too large");
    if (R == 0 || C == 0) return 0;
    CUtensorMap mw, mg, mm, mv, mrow, mcol;
    int rc;
    const void* srcs[4] = {w, g, m, v};
    CUtensorMap* maps[4] = {&mw, &mg, &mm, &mv};
    for (int i = 0; i < 4; i++) {
        rc = make_tensor_map_2d(maps[i], srcs[i], CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, C, R, C * 2, OT, OR_, CU_TENSOR_MAP_SWIZZLE_NONE);
        if (rc) return rc;
    }
    rc = make_tensor_map_2d(&mrow, do_row ? q_row : (do_col ? q_col : w), CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, C, R, C, OT, OR_, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    rc = make_tensor_map_2d(&mcol, do_col ? q_col : (do_row ? q_row : w), CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, C, R, C, OT, OR_, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
    AdamArgs a;
    a.lr_wd = 1.0f - lr * weight_decay;
    a.step_size = lr / bias_corr1;
    a.beta1 = beta1; a.beta2 = beta2; a.one_m_beta1 = 1.0f - beta1; a.one_m_beta2 = 1.0f - beta2;
    a.inv_sqrt_bc2 = 1.0f / sqrtf(bias_corr2);
    a.eps = eps;
    const int64_t ntiles = 2 * ((R + 127) / 128) * ((C + OT - 1) / OT);
    const int smem = sizeof(OptSmem) + 128;
    const int cap = sm_count();
    const int grid = ntiles < cap ? static_cast<int>(ntiles) : cap;
    cudaError_t e = cudaFuncSetAttribute(adamw_mx_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) { set_error("adamw_mx: smem attr: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH((adamw_mx_kernel), grid, O_THREADS, smem, stream, mw, mg, mm, mv, mrow, mcol, (uint8_t*)sf_row, (uint8_t*)sf_col, (int)R, (int)C, do_row, do_col, a);
    LLMFP8_LAUNCH_CHECK("adamw_mx");
    return 0;
}

}  // namespace llmfp8
