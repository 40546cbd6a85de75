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
// Persistent over (128*CG) x BN output tiles, grouped rasterisation so a wave's A/B panels stay L2-resident.
// CG = 2 (the default for large problems) runs CTA PAIRS (2-CTA clusters, tcgen05 cta_group::2): each CTA of the pair stages its own
// 128 rows of A and HALF of the B tile, the leader CTA issues one M=256 MMA that reads both halves, and each CTA keeps its
// 128 x BN accumulator in its own TMEM.  Shared-memory traffic per MMA cycle drops from 96 B/clk (1-CTA 128x256) to 64 B/clk —
// the 1-CTA kernel measured shared-memory bound (profiles/r01_*: BN sweep).
// BN (the UMMA N) is a RUNTIME parameter in {128, 192, 256}: the host picks it per problem to minimise wave quantisation
// (tiles / 148 CTAs).  MXFP8 keeps 64 TMEM columns for scale factors: at BN <= 192 two accumulators fit beside them, at BN = 256 the two
// accumulators overlap by 64 columns (OVL below) — either way the epilogue of tile t overlaps the mainloop of tile t+1.
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

constexpr int BM = 128;          // tile rows per CTA (UMMA M = 128 * CG)
constexpr int BK = 128;          // k elements per pipeline stage (= one 128 B swizzle row when K-major)
constexpr int UMMA_K = 32;       // k per tcgen05.mma for 8-bit operands
constexpr int GEMM_THREADS = 352;   // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quadrant), warp 10 tile scheduler
constexpr int EPI_WARPS = 8;
constexpr int NUM_WARPS = GEMM_THREADS / 32;
constexpr int CLC_ST = 2;           // cluster-launch-control queries in flight
constexpr int RING = 4;             // tile ids queued per CTA (the producer runs up to ~3 tiles ahead of the epilogue)
constexpr int RING_READERS = 10;    // warps 0..9 of the CTA
constexpr int CTRL_BYTES = 512;     // barriers + CLC responses + TMEM pointer
constexpr int GROUP_M = 16;      // rasterisation group, in 128-row blocks
constexpr int MAX_STAGES = 8;
constexpr int SF_COL0 = 448;     // TMEM columns 448..511: four rotating scale-factor buffers of 16 columns (SFA 4 + SFB up to 12), so the
constexpr int SF_BUFS = 4;       // smem->TMEM copy for k-block i+1 never waits for the MMAs of k-block i to finish reading theirs
constexpr int SMEM_BUDGET = 227 * 1024;
constexpr int EPI_BUF_BYTES = 4096;          // one TMA-store box: 32 rows x 128 B
constexpr int EPI_BYTES = EPI_WARPS * EPI_BUF_BYTES;

struct GemmParams {
    void* d;                 // output, row-major [M, N], leading dimension ldd (elements)
    int64_t ldd;
    int M, N, K;
    const float* a_scale_inv;   // per-tensor only
    const float* b_scale_inv;
    const void* bias;           // bf16 [N] or null
    const void* residual;       // bf16 [M, N] (leading dimension ldr) added in fp32 before the bf16 rounding, or null (bf16 output, staged path)
    int64_t ldr;
    int accumulate;             // D += result
    int out_fp32;               // 0: bf16 output, 1: fp32 output
    uint32_t idesc;
    int bn;                     // tile columns = UMMA N (128 | 192 | 256)
    int b_rows_cta;             // B rows staged per CTA: bn / CG (K-major) or that rounded up to 128 (MN-major chunks)
    int stages;                 // smem ring depth
    int acc_stages;             // TMEM accumulator buffers (2; 1 for MX with bn > 192 when acc_overlap is off)
    int debug;                  // experiments only: bit 0 = skip the per-stage scale copies (smem -> TMEM), bit 1 = skip the scale TMA loads,
                                // bit 3 = static persistent grid, bit 4 = single accumulator for MX at 256 columns (no OVL),
                                // bit 5 = single accumulator for the fused-residual MX kernel (no RES + OVL), bit 6 = at most 5 ring stages, bit 7 = epilogue warps sleep between polls of their accumulator barrier,
                                // bit 8 = overlapped accumulators 192 columns apart (64 shared) instead of 224 (32 shared)
    int dynamic;                // 1: tiles handed out by cluster launch control (grid = one cluster per tile); 0: static round-robin over a persistent grid
    int sf_tiles;               // MX: 512 B scale tiles per stage = 1 (SFA) + 2 or 3 (SFB)
    int acc_overlap;            // MX at 256 columns: two accumulators that share TMEM columns (OVL kernels); debug bit 4 turns it off
    int ovl_stride;             // OVL kernels: column distance of the two accumulators — 224 (32 shared columns, two scale buffers at 480..511; the default) or,
                                // debug bit 8, 192 (64 shared columns, four scale buffers at 448..511: the first form)
    // cross-entropy epilogues of the bf16 lm_head GEMM (EPI 1 / 2)
    const int64_t* ce_labels;   // [M] class index per row
    const float* ce_lse;        // EPI 2: [M] log-sum-exp of each row's logits
    const float* ce_gscale;     // EPI 2: device scalar, upstream gradient / number of counted rows
    float2* ce_partials;        // EPI 1: [2 * num_n, M] (maximum, sum of exp(v - maximum)) per warp half and column tile
    float* ce_label_logit;      // EPI 1: [M]
    int64_t ce_ignore;          // rows with this label contribute no loss and no gradient
    unsigned long long* trace;  // experiments only (llmfp8_gemm_trace): clock64 stamps of cluster 0's MMA thread and one epilogue warp, 16 per tile
};

__device__ __forceinline__ float bf16_round(float x) { return __uint_as_float(pack_bf16x2(x, 0.f) << 16); }
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
constexpr float CE_LOG2E = 1.4426950408889634f;

// Shared-memory control block of one CTA.
struct GemmCtrl {
    uint64_t full[MAX_STAGES], empty[MAX_STAGES];
    uint64_t tmem_full[2], tmem_empty[2];
    uint64_t clc_full[CLC_ST], clc_empty[CLC_ST];     // CLC responses: scheduler warps only
    uint64_t ring_full[RING], ring_empty[RING];       // tile ids re-broadcast by this CTA's scheduler warp to its other warps
    alignas(16) uint8_t clc_resp[CLC_ST][16];
    int ring_tile[RING];
    uint32_t tmem_ptr;
    alignas(8) uint64_t ovl_free;                     // OVL kernels only: the epilogue has drained the TMEM columns both accumulators share
};
static_assert(sizeof(GemmCtrl) <= CTRL_BYTES, "control block");

// Tile sequence of one warp role.  Static mode: tile += units.  Dynamic mode: the cluster's first tile is its own blockIdx; every
// further tile is a not-yet-launched cluster that the scheduler warp cancelled through cluster launch control — so CTAs that start
// late (SMs held by a concurrent NCCL kernel) or run slow simply take fewer tiles, instead of stretching the kernel by a whole wave.
// The scheduler warp of each CTA reads the (multicast) CLC response and hands the tile id to the other warps through a small ring
// in shared memory, so the TMA / MMA warps never touch the async proxy fence the response read needs.
struct TileSched {
    GemmCtrl* c;
    int r, ph, dynamic, num_tiles, num_units;
    __device__ __forceinline__ void init(GemmCtrl* ctrl, int dyn, int ntiles, int nunits) {
        c = ctrl; r = 0; ph = 0; dynamic = dyn; num_tiles = ntiles; num_units = nunits;
    }
    // whole warp calls, AFTER it has issued the work of tile `cur`; -> true and `next` when another tile follows
    __device__ __forceinline__ bool advance(int cur, int& next) {
        if (!dynamic) {
            next = cur + num_units;
            return next < num_tiles;
        }
        ptx::mbar_wait(&c->ring_full[r], ph);
        const int t = *reinterpret_cast<volatile int*>(&c->ring_tile[r]);
        __syncwarp();
        if ((threadIdx.x & 31) == 0) ptx::mbar_arrive(&c->ring_empty[r]);
        if (++r == RING) { r = 0; ph ^= 1; }
        next = t;
        return t >= 0;
    }
};

template <int CG>
__device__ __forceinline__ void tma2d(const CUtensorMap* m, uint64_t* bar, uint32_t bar_leader, void* dst, int c0, int c1) {
    if constexpr (CG == 1) ptx::tma_load_2d(m, bar, dst, c0, c1);
    else ptx::tma_load_2d_cg2(m, bar_leader, dst, c0, c1);
}
template <int CG>
__device__ __forceinline__ void tma3d(const CUtensorMap* m, uint64_t* bar, uint32_t bar_leader, void* dst, int c0, int c1, int c2) {
    if constexpr (CG == 1) ptx::tma_load_3d(m, bar, dst, c0, c1, c2);
    else ptx::tma_load_3d_cg2(m, bar_leader, dst, c0, c1, c2);
}

// RES: fused residual add in the staged bf16 epilogue (fprop layouts only) — a separate instantiation so the 32 prefetch registers
// it needs never weigh on the plain kernels.
// OVL (MX at 256 columns without accumulate; debug bit 4 selects the single-accumulator kernel instead): two accumulators of 256 columns
// laid p.ovl_stride apart instead of one.  224 apart (the default): [0, 256) and [224, 480) share columns [224, 256), two rotating 16-column
// scale buffers live in [480, 512).  192 apart (debug bit 8, the first form): [0, 256) / [192, 448), four scale buffers in [448, 512).  The
// epilogue of a tile drains the shared columns first and signals `ovl_free`, after which the MMAs of the next tile may start while the
// rest of the previous accumulator is still being read: the exposed epilogue shrinks from a whole tile's read-out to the shared chunk —
// 32 lanes x 32 columns x 4 B per warp at the 16 B per cycle a TMEM lane quarter delivers = 256 cycles (64 columns: 512).  Measured, same
// box: qkv fprop 2476 -> 2505 TF, proj 2092 -> 2104, fc1 2987 -> 2994, in-step GEMM launches 277.1 -> 275.9 us (profiles/r02/NOTES.md, call 28).
// BF: bf16 operands (kind::f16, fp32 accumulate) through the same pipeline — a stage still holds 128 B of k per row, i.e. 64 k-elements,
// one tcgen05.mma covers 16 of them (32 B), MN-major operands come in 64-element (128 B) swizzle atoms.  The lm_head of the reference
// (HF nn.Linear in bf16, outside TE: train_fp8.py:278) runs on it.  EPI selects what the epilogue does with the fp32 accumulator:
//   0  scale / bias / residual -> D (everything above)
//   1  cross-entropy statistics of the bf16-rounded logits, no D: per (row, 64-column chunk set of this warp) the running maximum and
//      sum of exponentials -> p.ce_partials[(n_blk * 2 + half) * M + row], and the label's logit -> p.ce_label_logit[row]
//   2  d(logits) = (softmax - onehot) * ce_gscale[0] for rows whose label is not ce_ignore, 0 otherwise -> D (bf16), from p.ce_lse
template <int CG, bool A_MN, bool B_MN, bool MX, bool RES = false, bool OVL = false, bool BF = false, int EPI = 0>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
fp8_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const __grid_constant__ CUtensorMap map_sfa,
                const __grid_constant__ CUtensorMap map_sfb, const __grid_constant__ CUtensorMap map_d, const GemmParams p) {
    constexpr int A_BYTES = BM * BK;
    constexpr int ELT = BF ? 2 : 1;                  // operand element size in bytes
    constexpr int BKE = BK / ELT;                    // k elements per stage (a stage row is always 128 B)
    constexpr int CW = 128 / ELT;                    // MN-major operands: elements per 128 B swizzle atom
    static_assert(!(BF && MX) && !(EPI != 0 && (!BF || A_MN || B_MN || RES || OVL)), "unsupported combination");
    constexpr int TM = BM * CG;                      // output rows per tile (per CTA pair when CG == 2)
    constexpr int TMEM_COLS = 512;
    const int BN = p.bn;
    const int STAGES = p.stages, ACC_STAGES = p.acc_stages;
    const int stage_bytes = A_BYTES + p.b_rows_cta * BK;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (STS, not generic ST)
    uint8_t* epi_base = smem + STAGES * stage_bytes;          // 8 epilogue warps x 4 KB, 1024 B aligned (128B-swizzled TMA-store boxes)
    uint8_t* sf_base = epi_base + EPI_BYTES;
    const int sf_stage_bytes = MX ? 512 * p.sf_tiles : 0;
    GemmCtrl* ctrl = reinterpret_cast<GemmCtrl*>(sf_base + STAGES * sf_stage_bytes);
    uint64_t* full_bar = ctrl->full;
    uint64_t* empty_bar = ctrl->empty;
    uint64_t* tmem_full = ctrl->tmem_full;
    uint64_t* tmem_empty = ctrl->tmem_empty;
    uint32_t* tmem_ptr_smem = &ctrl->tmem_ptr;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CG == 2 ? ptx::cluster_ctarank() : 0u;    // 0 = leader (issues the MMAs)
    const int unit = blockIdx.x / CG, num_units = gridDim.x / CG;     // dynamic mode: unit = this cluster's own (first) tile
    const int num_m = (p.M + TM - 1) / TM, num_n = (p.N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_kb = (p.K + BKE - 1) / BKE;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&map_a);
        ptx::prefetch_tensormap(&map_b);
        if (MX) {
            ptx::prefetch_tensormap(&map_sfa);
            ptx::prefetch_tensormap(&map_sfb);
        }
        for (int s = 0; s < STAGES; s++) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < ACC_STAGES; s++) {
            ptx::mbar_init(&tmem_full[s], 1);
            ptx::mbar_init(&tmem_empty[s], EPI_WARPS * CG);
        }
        if constexpr (OVL) ptx::mbar_init(&ctrl->ovl_free, ((p.out_fp32 && p.ovl_stride == 192) ? EPI_WARPS : EPI_WARPS / 2) * CG);   // the warps that own a shared chunk
        for (int s = 0; s < CLC_ST; s++) {
            ptx::mbar_init(&ctrl->clc_full[s], 1);
            ptx::mbar_init(&ctrl->clc_empty[s], CG);                 // the scheduler warp of every CTA of the pair reads each response
        }
        for (int s = 0; s < RING; s++) {
            ptx::mbar_init(&ctrl->ring_full[s], 1);
            ptx::mbar_init(&ctrl->ring_empty[s], RING_READERS);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 1) ptx::tmem_alloc_cg<CG>(tmem_ptr_smem, TMEM_COLS);
    ptx::tc_fence_before();
    if constexpr (CG == 2) ptx::cluster_sync();      // peer barriers must be initialised before anything signals them
    else __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    pdl_wait();                // LLMFP8_PDL build only (common.cuh): barrier init, descriptor prefetch, TMEM allocation and the cluster
    pdl_launch_dependents();   // sync above ran under the previous kernel's tail; nothing before this line touches global memory

    constexpr int GROUP = GROUP_M / CG;
    auto tile_coords = [&](int tile, int& m_blk, int& n_blk) {
        const int per_group = GROUP * num_n;
        const int group = tile / per_group;
        const int first_m = group * GROUP;
        const int gsize = (num_m - first_m) < GROUP ? (num_m - first_m) : GROUP;
        const int in_group = tile - group * per_group;
        m_blk = first_m + in_group % gsize;
        n_blk = in_group / gsize;
    };

    TileSched sched;
    sched.init(ctrl, p.dynamic, num_tiles, num_units);

    if (warp == 0) {
        // ================================ TMA producer (one per CTA: own A rows, own part of B, own scale tiles)
        {
            int stage = 0, phase = 0;
            const int b_chunks = p.b_rows_cta / CW;     // MN-major B: 128-byte-wide column chunks (the last one may over-fetch)
            const int bn_cta = BN / CG;
            const bool sf_load = MX && !(p.debug & 2);
            const uint32_t tx_bytes = CG * (stage_bytes + (sf_load ? sf_stage_bytes : 0));
            int tile = unit, next_tile;
            for (bool more = true; more; tile = next_tile) {
                int m_blk, n_blk;
                tile_coords(tile, m_blk, n_blk);
                const int m0 = m_blk * TM + rank * BM, n0 = n_blk * BN, nb0 = n0 + rank * bn_cta;
                for (int kb = 0; kb < num_kb; kb++) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);     // all lanes poll (keeps the warp converged and its values uniform)
                    if (ptx::elect_one_sync()) {
                    uint8_t* sa = smem + stage * stage_bytes;
                    uint8_t* sb = sa + A_BYTES;
                    uint64_t* fb = &full_bar[stage];
                    const uint32_t fb_leader = CG == 2 ? ptx::mapa(ptx::smem_u32(fb), 0) : 0u;
                    if (rank == 0) ptx::mbar_arrive_expect_tx(fb, tx_bytes);
                    const int k0 = kb * BKE;
                    if (!A_MN) {
                        tma2d<CG>(&map_a, fb, fb_leader, sa, k0, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / CW; j++) tma2d<CG>(&map_a, fb, fb_leader, sa + j * (BKE * 128), m0 + j * CW, k0);
                    }
                    if (!B_MN) {
                        tma2d<CG>(&map_b, fb, fb_leader, sb, k0, nb0);
                    } else {
                        for (int j = 0; j < b_chunks; j++) tma2d<CG>(&map_b, fb, fb_leader, sb + j * (BKE * 128), nb0 + j * CW, k0);
                    }
                    if (sf_load) {
                        uint8_t* ssf = sf_base + stage * sf_stage_bytes;
                        tma3d<CG>(&map_sfa, fb, fb_leader, ssf, 0, kb, m_blk * CG + static_cast<int>(rank));
                        tma3d<CG>(&map_sfb, fb, fb_leader, ssf + 512, 0, kb, n0 >> 7);     // two (three at 224 columns) 128-row scale tiles, zero-filled past the end
                    }
                    }
                    __syncwarp();
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                more = sched.advance(tile, next_tile);
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (one thread of the leader CTA)
        if (rank == 0) {
            int stage = 0, phase = 0, acc = 0, acc_phase = 0, sfi = 0;
            int tiles_done = 0;                 // OVL: tiles this cluster has issued so far
            int trace_n = 0;
            constexpr uint32_t A_LBO = A_MN ? (BKE * 128) : 16, B_LBO = B_MN ? (BKE * 128) : 16;      // MN-major: distance between 128 B atoms
            constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K / ELT * 128) : UMMA_K, B_KSTEP = B_MN ? (UMMA_K / ELT * 128) : UMMA_K;   // bytes per MMA k-step (32 B of k)
            int tile = unit, next_tile;
            for (bool more = true; more; tile = next_tile) {
                int m_blk, n_blk;
                tile_coords(tile, m_blk, n_blk);
                // MX: the scale of column n lives in TMEM column (sf buffer) + 4 + (n - 128 * (n0 / 128)) / 32: a tile that starts 64 rows
                // into a 128-row scale tile (BN = 192) skips the leading 32-row groups by advancing the column address
                const uint32_t sfb_off = 4 + (((n_blk * BN) & 127) >> 5);
                unsigned long long* tr = (p.trace && blockIdx.x == 0 && trace_n < 64) ? p.trace + 16 * trace_n : nullptr;
                trace_n++;
                if (tr && lane == 0) tr[0] = clock64();
                ptx::mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                if (tr && lane == 0) tr[1] = clock64();
                if constexpr (OVL) {
                    if (tiles_done > 0) ptx::mbar_wait(&ctrl->ovl_free, (tiles_done - 1) & 1);   // previous tile's shared columns are in registers
                    tiles_done++;
                }
                if (tr && lane == 0) tr[2] = clock64();
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * (OVL ? p.ovl_stride : BN);
                const uint32_t sf_col0 = (OVL && p.ovl_stride == 224) ? 480u : static_cast<uint32_t>(SF_COL0);
                const int sf_mask = (OVL && p.ovl_stride == 224) ? 1 : SF_BUFS - 1;     // 32 columns hold two 16-column scale buffers
                for (int kb = 0; kb < num_kb; kb++) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    if (tr && lane == 0 && kb == 0) tr[3] = clock64();
                    ptx::tc_fence_after();
                    const uint32_t sa = ptx::smem_u32(smem + stage * stage_bytes);
                    const uint32_t sb = sa + A_BYTES;
                    const uint32_t tmem_sf = tmem_base + sf_col0 + 16 * sfi;
                    if (ptx::elect_one_sync()) {
                    if (MX && !((p.debug & 1) && (kb | tile) != 0)) {
                        const uint32_t ssf = ptx::smem_u32(sf_base + stage * sf_stage_bytes);
                        ptx::tc_cp_32x128b_warpx4_cg<CG>(tmem_sf, ptx::make_smem_desc(ssf, 0, 128, 0));
                        ptx::tc_cp_32x128b_warpx4_cg<CG>(tmem_sf + 4, ptx::make_smem_desc(ssf + 512, 0, 128, 0));
                        ptx::tc_cp_32x128b_warpx4_cg<CG>(tmem_sf + 8, ptx::make_smem_desc(ssf + 1024, 0, 128, 0));
                        if (p.sf_tiles == 4) ptx::tc_cp_32x128b_warpx4_cg<CG>(tmem_sf + 12, ptx::make_smem_desc(ssf + 1536, 0, 128, 0));
                    }
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; k++) {
                        const uint64_t da = ptx::make_smem_desc(sa + k * A_KSTEP, A_LBO, 1024, 2);
                        const uint64_t db = ptx::make_smem_desc(sb + k * B_KSTEP, B_LBO, 1024, 2);
                        const uint32_t accum = (kb | k) != 0;
                        if (MX) {
                            const uint32_t idesc = p.idesc | (static_cast<uint32_t>(k) << 4) | (static_cast<uint32_t>(k) << 29);
                            ptx::mma_mxf8f6f4_cg<CG>(tmem_d, da, db, idesc, accum, tmem_sf, tmem_sf + sfb_off);
                        } else if (BF) {
                            ptx::mma_f16_cg<CG>(tmem_d, da, db, p.idesc, accum);
                        } else {
                            ptx::mma_f8f6f4_cg<CG>(tmem_d, da, db, p.idesc, accum);
                        }
                    }
                    ptx::tc_commit_cg<CG>(&empty_bar[stage]);   // smem slot (in both CTAs) reusable once these MMAs have read it
                    }
                    __syncwarp();
                    if (MX && !(p.debug & 4)) sfi = (sfi + 1) & sf_mask;
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (tr && lane == 0) tr[4] = clock64();
                if (ptx::elect_one_sync()) ptx::tc_commit_cg<CG>(&tmem_full[acc]);         // accumulator complete -> epilogue warps of both CTAs
                __syncwarp();
                if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
                more = sched.advance(tile, next_tile);
            }
        } else {
            // the peer's MMA warp issues nothing, but it is one of the readers every CLC response waits for
            int tile = unit, next_tile;
            while (sched.advance(tile, next_tile)) tile = next_tile;
        }
    } else if (warp == NUM_WARPS - 1) {
        // ================================ tile scheduler.  Per round: (leader) wait until both CTAs' scheduler warps have read the
        // response slot, arm this CTA's barrier for the 16 response bytes, (leader) issue the query — the response is multicast to both
        // CTAs; read it (async-proxy fenced), free the slot, push the tile id into this CTA's ring.  Stops after the first failed query.
        if (p.dynamic) {
            const uint32_t clc_empty_leader = CG == 2 ? ptx::mapa(ptx::smem_u32(&ctrl->clc_empty[0]), 0) : 0u;
            int cs = 0, cph = 0, rr = 0, rph = 0;
            for (bool more = true; more;) {
                if (rank == 0) ptx::mbar_wait(&ctrl->clc_empty[cs], cph ^ 1);
                if (ptx::elect_one_sync()) {
                    ptx::mbar_arrive_expect_tx(&ctrl->clc_full[cs], 16);
                    if (rank == 0) ptx::clc_try_cancel<CG>(ctrl->clc_resp[cs], &ctrl->clc_full[cs]);
                }
                __syncwarp();
                ptx::mbar_wait(&ctrl->clc_full[cs], cph);
                uint32_t x;
                more = ptx::clc_query(ctrl->clc_resp[cs], x);
                ptx::fence_proxy_async();        // the response read is ordered before the async-proxy write of the next query into this slot
                __syncwarp();
                if (lane == 0) {
                    if constexpr (CG == 2) ptx::mbar_arrive_cluster(clc_empty_leader + 8 * cs);
                    else ptx::mbar_arrive(&ctrl->clc_empty[cs]);
                }
                if (++cs == CLC_ST) { cs = 0; cph ^= 1; }
                ptx::mbar_wait(&ctrl->ring_empty[rr], rph ^ 1);
                if (lane == 0) {
                    ctrl->ring_tile[rr] = more ? static_cast<int>(x) / CG : -1;
                    ptx::mbar_arrive(&ctrl->ring_full[rr]);          // release: the tile id is visible to whoever acquires this phase
                }
                __syncwarp();
                if (++rr == RING) { rr = 0; rph ^= 1; }
            }
        }
    } else {
        // ================================ epilogue: TMEM -> registers -> (alpha, bias, accumulate) -> global
        const int quad = warp & 3;            // TMEM lane quadrant this warp may read
        const int half = (warp - 2) >> 2;     // two warps per quadrant split the column chunks (even / odd)
        int acc = 0, acc_phase = 0;
        float alpha = 1.0f;
        if (!MX && !BF) alpha = __fmul_rn(__ldg(p.a_scale_inv), __ldg(p.b_scale_inv));
        const uint32_t tmem_empty_leader = CG == 2 ? ptx::mapa(ptx::smem_u32(&tmem_empty[0]), 0) : 0u;
        auto release_acc = [&](int a) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if constexpr (CG == 2) ptx::mbar_arrive_cluster(tmem_empty_leader + a * 8);
                else ptx::mbar_arrive(&tmem_empty[a]);
            }
        };
        if (!p.accumulate) {
            // ---- staged path: 32 x 128 B boxes (64 bf16 or 32 fp32 columns) -> swizzled smem -> TMA store (coalesced, clips partial tiles)
            uint8_t* stg = epi_base + (warp - 2) * EPI_BUF_BYTES;
            int etrace_n = 0;
            const int ccols = p.out_fp32 ? 32 : 64;
            const int nchunks = (BN + ccols - 1) / ccols;
            // column offset of chunk c inside the tile.  A tile width that is not a multiple of the chunk (224 = 3.5 x 64) lets the LAST chunk
            // start early so that it ends with the tile: it re-stores 32 columns its neighbour chunk stores too — the same accumulators, the
            // same bytes — instead of needing a second, narrower store box
            auto coff = [&](int c) { return c * ccols + ccols <= BN ? c * ccols : BN - ccols; };
            if (lane == 0) ptx::prefetch_tensormap(&map_d);
            int tile = unit, next_tile;
            for (bool more = true; more; tile = next_tile) {
                int m_blk, n_blk;
                tile_coords(tile, m_blk, n_blk);
                const int row0 = m_blk * TM + static_cast<int>(rank) * BM + quad * 32;
                const int n0 = n_blk * BN;
                // cross-entropy epilogues: this lane's row
                int64_t ce_lab = -1;
                float ce_m = -INFINITY, ce_s = 0.f, ce_ll = 0.f, ce_lse = 0.f, ce_gs = 0.f;
                bool ce_hit = false;
                if constexpr (EPI != 0) {
                    if (row0 + lane < p.M) {
                        ce_lab = p.ce_labels[row0 + lane];
                        if constexpr (EPI == 2) {
                            ce_lse = p.ce_lse[row0 + lane];
                            ce_gs = ce_lab != p.ce_ignore ? __ldg(p.ce_gscale) : 0.f;
                        }
                    }
                }
                // fused residual: this lane's 128 B row segment of R for one chunk lives in rv.  Software-pipelined through that single
                // buffer: the first chunk is requested BEFORE waiting for the accumulator (its DRAM latency hides under the main loop's
                // tail — the MX kernel at 256 columns has one accumulator, so its epilogue is exposed), the next chunk right after the
                // current one has been consumed.
                uint4 rv[RES ? 8 : 1];
                auto load_residual = [&](int c) {
                    const int row = row0 + lane;
                    const int col = n0 + coff(c);
                    const uint16_t* rp = static_cast<const uint16_t*>(p.residual) + static_cast<int64_t>(row) * p.ldr + col;
#pragma unroll
                    for (int g = 0; g < (RES ? 8 : 1); g++)
                        rv[g] = (row < p.M && col + g * 8 < p.N) ? __ldg(reinterpret_cast<const uint4*>(rp + g * 8)) : make_uint4(0, 0, 0, 0);
                };
                unsigned long long* etr = (p.trace && blockIdx.x == 0 && (warp == 2 || warp == 6) && etrace_n < 64) ? p.trace + 16 * etrace_n + (warp == 2 ? 5 : 10) : nullptr;
                etrace_n++;
                if constexpr (RES) load_residual((OVL && acc == 0) ? half + 2 * (((nchunks - half + 1) >> 1) - 1) : half);   // the chunk the walk below starts with
                if (p.debug & 128) ptx::mbar_wait_backoff(&tmem_full[acc], acc_phase, 128);      // experiment: sleeping instead of spinning epilogue warps
                else ptx::mbar_wait(&tmem_full[acc], acc_phase);
                ptx::tc_fence_after();
                if (etr && lane == 0) etr[0] = clock64();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * (OVL ? p.ovl_stride : BN);
                // OVL: this warp's chunks are walked so that the one lying in the shared columns [192, 256) comes first — the LAST chunk of
                // the accumulator at column 0 (walk down), the FIRST chunk of the one at column 192 (walk up).  bf16 output: 64-column
                // chunks, the shared one belongs to one `half`; fp32 output: 32-column chunks, each half owns one of the two shared chunks.
                const int nmine = (nchunks - half + 1) >> 1;
                const bool walk_down = OVL && acc == 0;
                const bool narrow = OVL && p.ovl_stride == 224;     // 32 shared columns: half of a bf16 chunk, or one fp32 chunk
                const bool owns_shared = OVL && ((p.out_fp32 && !narrow) || half == (acc == 0 ? 1 : 0));
#pragma unroll 1
                for (int c = walk_down ? half + 2 * (nmine - 1) : half, n = 0; OVL ? (n < nmine) : (c < nchunks); c += walk_down ? -2 : 2, n++) {
                    const int col = n0 + coff(c);
                    uint32_t r0[32], r1[32];
                    auto free_shared = [&]() {                          // the shared columns are in registers: the next tile's MMAs may overwrite them
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if constexpr (CG == 2) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&ctrl->ovl_free), 0));
                            else ptx::mbar_arrive(&ctrl->ovl_free);
                        }
                    };
                    if (OVL && narrow && n == 0 && owns_shared && !p.out_fp32) {
                        // the shared 32 columns are ONE half of this 64-column chunk (the upper half of the accumulator at column 0, the lower
                        // half of the one at column 224): pull that half alone, hand the columns over, then fetch the other half
                        if (walk_down) ptx::tmem_ld_32x32b_x32(taddr + coff(c) + 32, r1);
                        else ptx::tmem_ld_32x32b_x32(taddr + coff(c), r0);
                        ptx::tmem_ld_wait();
                        free_shared();
                        if (walk_down) ptx::tmem_ld_32x32b_x32(taddr + coff(c), r0);
                        else ptx::tmem_ld_32x32b_x32(taddr + coff(c) + 32, r1);
                        if (lane == 0) ptx::tma_store_wait_read<0>();
                        ptx::tmem_ld_wait();
                        if (etr && lane == 0) etr[1] = clock64();
                    } else {
                        ptx::tmem_ld_32x32b_x32(taddr + coff(c), r0);
                        if (!p.out_fp32) ptx::tmem_ld_32x32b_x32(taddr + coff(c) + 32, r1);
                        if (lane == 0) ptx::tma_store_wait_read<0>();      // the box previously stored from this warp's buffer has been read
                        ptx::tmem_ld_wait();
                        if (etr && lane == 0 && n < 2) etr[1 + n] = clock64();
                        if constexpr (OVL) {
                            if (n == 0 && owns_shared) free_shared();
                        }
                    }
                    if (OVL ? (n + 1 == nmine) : (c + 2 >= nchunks)) release_acc(acc);           // this warp's share of the accumulator is in registers
                    else __syncwarp();
                    if constexpr (EPI == 1) {
                        // statistics of the bf16-rounded logits (HF rounds the lm_head output to bf16 before the fp32 cross-entropy)
                        if (col < p.N) {
                            const int ncol = p.N - col < 64 ? p.N - col : 64;
                            const int lj = (ce_lab >= col && ce_lab < col + ncol) ? static_cast<int>(ce_lab - col) : -1;
                            float cm = -INFINITY;
#pragma unroll
                            for (int j = 0; j < 64; j++) {
                                const float v = bf16_round(__uint_as_float(j < 32 ? r0[j] : r1[j - 32]));
                                if (j < ncol) cm = fmaxf(cm, v);
                                if (j == lj) { ce_ll = v; ce_hit = true; }
                            }
                            float cs = 0.f;
#pragma unroll
                            for (int j = 0; j < 64; j++) {
                                const float v = bf16_round(__uint_as_float(j < 32 ? r0[j] : r1[j - 32]));
                                if (j < ncol) cs += ex2_approx((v - cm) * CE_LOG2E);
                            }
                            const float nm = fmaxf(ce_m, cm);
                            ce_s = ce_s * ex2_approx((ce_m - nm) * CE_LOG2E) + cs * ex2_approx((cm - nm) * CE_LOG2E);
                            ce_m = nm;
                        }
                        continue;
                    }
                    uint8_t* sbuf = stg + lane * 128;
                    const uint32_t sw = lane & 7;
                    if (col < p.N) {
                        if (p.out_fp32) {
#pragma unroll
                            for (int g = 0; g < 8; g++) {
                                float4 o = make_float4(__uint_as_float(r0[g * 4]) * alpha, __uint_as_float(r0[g * 4 + 1]) * alpha,
                                                       __uint_as_float(r0[g * 4 + 2]) * alpha, __uint_as_float(r0[g * 4 + 3]) * alpha);
                                if (p.bias && col + g * 4 < p.N) {
                                    const uint2 bv = __ldg(reinterpret_cast<const uint2*>(static_cast<const uint16_t*>(p.bias) + col + g * 4));
                                    o.x += bf16lo(bv.x); o.y += bf16hi(bv.x); o.z += bf16lo(bv.y); o.w += bf16hi(bv.y);
                                }
                                *reinterpret_cast<float4*>(sbuf + ((g ^ sw) << 4)) = o;
                            }
                        } else {
#pragma unroll
                            for (int g = 0; g < 8; g++) {
                                float v[8];
#pragma unroll
                                for (int j = 0; j < 8; j++) v[j] = __uint_as_float(g < 4 ? r0[g * 8 + j] : r1[(g - 4) * 8 + j]) * alpha;
                                if (p.bias && col + g * 8 < p.N) {
                                    float bf[8];
                                    unpack8(__ldg(reinterpret_cast<const uint4*>(static_cast<const uint16_t*>(p.bias) + col + g * 8)), bf);
#pragma unroll
                                    for (int j = 0; j < 8; j++) v[j] += bf[j];
                                }
                                if constexpr (RES) {
                                    float rf[8];
                                    unpack8(rv[g], rf);
#pragma unroll
                                    for (int j = 0; j < 8; j++) v[j] += rf[j];
                                }
                                if constexpr (EPI == 2) {
#pragma unroll
                                    for (int j = 0; j < 8; j++) {
                                        const float pj = ex2_approx((bf16_round(v[j]) - ce_lse) * CE_LOG2E);
                                        v[j] = (pj - (static_cast<int64_t>(col + g * 8 + j) == ce_lab ? 1.f : 0.f)) * ce_gs;
                                    }
                                }
                                uint4 o;
                                o.x = pack_bf16x2(v[0], v[1]); o.y = pack_bf16x2(v[2], v[3]);
                                o.z = pack_bf16x2(v[4], v[5]); o.w = pack_bf16x2(v[6], v[7]);
                                *reinterpret_cast<uint4*>(sbuf + ((g ^ sw) << 4)) = o;
                            }
                        }
                        ptx::fence_proxy_async();
                        __syncwarp();
                        if (lane == 0) {
                            ptx::tma_store_2d(&map_d, stg, col, row0);
                            ptx::tma_store_commit();
                        }
                    }
                    if constexpr (RES) {
                        // rv is free again: next chunk's segment, under the store above.  Outside the `col < N` guard: the downward walk of
                        // the OVL kernel meets the chunks past the matrix edge FIRST, and the in-range ones after them still need their segment.
                        if (OVL ? (n + 1 < nmine) : (c + 2 < nchunks)) load_residual(c + (walk_down ? -2 : 2));
                    }
                }
                if (etr && lane == 0) etr[3] = clock64();
                if constexpr (EPI == 1) {
                    if (row0 + lane < p.M) {
                        p.ce_partials[static_cast<int64_t>(n_blk * 2 + half) * p.M + row0 + lane] = make_float2(ce_m, ce_s);
                        if (ce_hit) p.ce_label_logit[row0 + lane] = ce_ll;
                    }
                }
                if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
                more = sched.advance(tile, next_tile);
            }
            if (lane == 0) ptx::tma_store_wait<0>();
        } else {
        int tile = unit, next_tile;
        for (bool more = true; more; tile = next_tile) {
            int m_blk, n_blk;
            tile_coords(tile, m_blk, n_blk);
            const int row = m_blk * TM + static_cast<int>(rank) * BM + quad * 32 + lane;
            const int n0 = n_blk * BN;
            ptx::mbar_wait(&tmem_full[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
#pragma unroll 1
            for (int c = half; c < (BN >> 5); c += 2) {
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
            if (lane == 0) {
                if constexpr (CG == 2) ptx::mbar_arrive_cluster(tmem_empty_leader + acc * 8);
                else ptx::mbar_arrive(&tmem_empty[acc]);
            }
            if (++acc == ACC_STAGES) { acc = 0; acc_phase ^= 1; }
            more = sched.advance(tile, next_tile);
        }
        }
    }

    ptx::tc_fence_before();
    if constexpr (CG == 2) ptx::cluster_sync();      // neither CTA may retire while its peer can still signal its barriers / read its smem
    else __syncthreads();
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_cg<CG>(tmem_base, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------ host side
// Tuning / test knobs (llmfp8_gemm_force_bn / _force_cg / _debug).  Process-wide by nature (they pick which kernel every thread's next
// launch gets), so they are inert unless the process opted in with LLMFP8_ALLOW_TUNING=1 before the first call — the test suite and the
// tools/ scripts do, a training process does not, and then no caller can perturb the launches of the autograd thread.  Atomics: a setter
// racing a launch yields the old or the new value, never a torn one.
static std::atomic<int> g_bn_override_a{0}, g_cg_override_a{0}, g_debug_a{0};
static bool tuning_allowed() {
    static const bool ok = [] { const char* e = getenv("LLMFP8_ALLOW_TUNING"); return e && e[0] == '1'; }();
    return ok;
}
static void set_knob(std::atomic<int>& k, int v, const char* what) {
    if (v != 0 && !tuning_allowed()) { set_error("%s ignored: set LLMFP8_ALLOW_TUNING=1 before loading the library (tuning / tests only)", what); return; }
    k.store(v, std::memory_order_relaxed);
}
static std::atomic<unsigned long long*> g_trace{nullptr};
void gemm_trace(unsigned long long* buf) { if (buf == nullptr || tuning_allowed()) g_trace.store(buf, std::memory_order_relaxed); }
void gemm_debug(int d) { set_knob(g_debug_a, d, "llmfp8_gemm_debug"); }
void gemm_force_bn(int bn) { set_knob(g_bn_override_a, bn, "llmfp8_gemm_force_bn"); }
void gemm_force_cg(int cg) { set_knob(g_cg_override_a, cg, "llmfp8_gemm_force_cg"); }

template <int CG, bool A_MN, bool B_MN, bool MX, bool RES = false, bool OVL = false, bool BF = false, int EPI = 0>
static int launch_variant(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa, const CUtensorMap& msb, const CUtensorMap& md,
                          const GemmParams& p, int grid, int smem_bytes, cudaStream_t stream) {
    auto kern = fp8_gemm_kernel<CG, A_MN, B_MN, MX, RES, OVL, BF, EPI>;
    static bool configured[64] = {};   // per device (the attribute lives in the device's context); benign race: the call is idempotent
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = -1;
    if (dev < 0 || !configured[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET);
        if (e != cudaSuccess) { set_error("fp8_gemm: smem attribute (%d B): %s", SMEM_BUDGET, cudaGetErrorString(e)); return (int)e; }
        if (dev >= 0) configured[dev] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
#ifdef LLMFP8_PDL
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.numAttrs = 2;
#endif
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, ma, mb, msa, msb, md, p);
    if (e != cudaSuccess) { set_error("fp8_gemm: launch failed: %s", cudaGetErrorString(e)); return (int)e; }
    LLMFP8_LAUNCH_CHECK("fp8_gemm");
    return 0;
}

template <int CG>
static int dispatch(bool a_mn, bool b_mn, bool mx, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& msa, const CUtensorMap& msb,
                    const CUtensorMap& md, const GemmParams& p, int grid, int smem_bytes, cudaStream_t stream) {
    if (p.residual) {
        if (a_mn || b_mn) { set_error("fp8_gemm: the fused residual is built for the fprop layout only (A, B K-major)"); return -1; }
        if (mx && p.acc_overlap) return launch_variant<CG, false, false, true, true, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        if (mx) return launch_variant<CG, false, false, true, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        return launch_variant<CG, false, false, false, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    }
    if (p.acc_overlap) {       // MX, 256 columns, two overlapping accumulators
        if (!a_mn && !b_mn) return launch_variant<CG, false, false, true, false, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        if (!a_mn && b_mn) return launch_variant<CG, false, true, true, false, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        return launch_variant<CG, true, true, true, false, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    }
    if (!mx) {
        if (!a_mn && !b_mn) return launch_variant<CG, false, false, false>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        if (!a_mn && b_mn) return launch_variant<CG, false, true, false>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        if (a_mn && b_mn) return launch_variant<CG, true, true, false>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
        return launch_variant<CG, true, false, false>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    }
    if (!a_mn && !b_mn) return launch_variant<CG, false, false, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    if (!a_mn && b_mn) return launch_variant<CG, false, true, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    return launch_variant<CG, true, true, true>(ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
}

// Tile width: minimise (waves x per-tile cost).  Narrow tiles cost more per column: their shared-memory traffic per MMA cycle is
// higher (measured, profiles/r01_gemm_bn_sweep: 192 runs 10-25 % and 128 30-35 % below 256 in the 1-CTA kernel).
static int pick_bn(int M, int N, int K, bool mx, bool b_mn, int cg, int units) {
    const double k_cycles = 15500.0 * K / 4096.0;             // issue span of one 256 x 256 tile's main loop (measured at K = 4096)
    if (const int f = g_bn_override_a.load(std::memory_order_relaxed)) return f;
    const int cand[4] = {256, 224, 192, 128};
    const double eff1[4] = {1.0, 0.94, b_mn ? 0.76 : 0.88, 0.70};
    const double eff2[4] = {1.0, 0.97, 0.90, 0.80};
    int best = 0;
    double best_cost = 1e300;
    const int num_m = (M + BM * cg - 1) / (BM * cg);
    for (int i = 0; i < 4; i++) {
        const int bn = cand[i];
        if (cg == 2 && bn == 192 && b_mn) continue;           // 96 MN-major columns per CTA is not a whole swizzle chunk
        // 224 columns (two whole accumulators beside the scale factors, no hand-over) is built and bit-identical to 256 on its FIRST column
        // tile — and faults (misaligned address) on every other one: tiles then start 96 / 64 / 32 rows into a 128-row scale tile, and the
        // block-scaled MMA takes its SFB operand from EVEN TMEM columns only (192-column tiles start at offsets 0 / 64: columns 0 / 2, fine;
        // 224 needs columns 3 and 1; tcgen05.cp cannot shift by one column).  Measured on hardware (profiles/r02/g_gemm224_offsets.txt).
        // Never picked: what remains is the forced single-tile case the test keeps alive for the 3-tile SFB / half-chunk code under it.
        if (bn == 224) continue;
        if (bn > 128 && N <= bn - 64) continue;               // do not pad tiny N
        double e = cg == 2 ? eff2[i] : eff1[i];
        if (mx) {
            // 512 TMEM columns = accumulators + 64 columns of scale factors.  At 256 the two accumulators overlap by 64 columns: the next
            // tile's MMAs wait until the previous tile has COMPLETED and its shared columns are drained — the tensor pipe runs empty once per
            // tile (clock trace, profiles/r02/f_gemm_clock_trace_fc1.txt: ~2700 of ~19100 cycles per tile at K = 4096).  At 224 two whole
            // accumulators fit ([0, 224), [224, 448)) and the pipe never drains; narrower tiles pay more scale copies per FLOP.
            if (bn == 256) e *= k_cycles / (k_cycles + 2700.0);
            else if (bn < 224) e *= 0.85;
        }
        const long tiles = static_cast<long>(num_m) * ((N + bn - 1) / bn);
        // tiles are handed to whichever cluster is free (cluster launch control): the cost is work plus half a tile of tail, not whole waves
        const double cost = (static_cast<double>(tiles) / units + (tiles % units ? 0.5 : 0.0)) * bn / e;
        if (cost < best_cost) { best_cost = cost; best = bn; }
    }
    return best ? best : 128;
}

int make_tensor_map_3d(CUtensorMap* map, const void* base, CUtensorMapDataType dtype, int64_t d0, int64_t d1, int64_t d2, int64_t stride1_bytes,
                       int64_t stride2_bytes, int b0, int b1, int b2);

// a: fp8 bytes. !a_mn: stored [M, K] (lda = row pitch in bytes >= K);  a_mn: stored [K, M] (lda >= M).  Same for b with N.
int launch_fp8_gemm(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd,
                    int out_fp32, int M, int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* sfa, const void* sfb,
                    const void* bias, const void* residual, int64_t ldr, int accumulate, int max_ctas, cudaStream_t stream) {
    const bool mx = sfa != nullptr || sfb != nullptr;
    LLMFP8_REQUIRE(!residual || (!accumulate && !out_fp32), "fp8_gemm: a fused residual needs a bf16 output without accumulate");
    LLMFP8_REQUIRE(!residual || (ldr % 8 == 0 && ldr >= N && reinterpret_cast<uintptr_t>(residual) % 16 == 0 && residual != d),
                   "fp8_gemm: residual must be 16 B aligned, distinct from D, with ldr %% 8 == 0 and ldr >= N (ldr=%lld)", (long long)ldr);
    LLMFP8_REQUIRE(M > 0 && N > 0 && K > 0, "fp8_gemm: empty problem M=%d N=%d K=%d", M, N, K);
    LLMFP8_REQUIRE((a_fmt == FMT_E4M3 || a_fmt == FMT_E5M2) && (b_fmt == FMT_E4M3 || b_fmt == FMT_E5M2), "fp8_gemm: bad operand formats");
    LLMFP8_REQUIRE(mx ? (sfa && sfb) : (a_scale_inv && b_scale_inv), "fp8_gemm: %s", mx ? "MX needs both scale tensors" : "per-tensor needs both scale_inv pointers");
    LLMFP8_REQUIRE(lda % 16 == 0 && ldb % 16 == 0, "fp8_gemm: operand pitches must be multiples of 16 B (lda=%lld ldb=%lld)", (long long)lda, (long long)ldb);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(d)) % 16 == 0, "fp8_gemm: pointers must be 16 B aligned");
    LLMFP8_REQUIRE(N % 8 == 0 && ldd % 8 == 0, "fp8_gemm: N and ldd must be multiples of 8 (N=%d ldd=%lld)", N, (long long)ldd);
    LLMFP8_REQUIRE(!mx || K % 32 == 0, "fp8_gemm: MXFP8 needs K %% 32 == 0 (K=%d)", K);
    LLMFP8_REQUIRE(!mx || (reinterpret_cast<uintptr_t>(sfa) | reinterpret_cast<uintptr_t>(sfb)) % 16 == 0, "fp8_gemm: scale tensors must be 16 B aligned");
    LLMFP8_REQUIRE(!(mx && (a_mn != b_mn) && a_mn), "fp8_gemm: operand layout combination (A MN-major, B K-major) is not built");

    const int g_debug = g_debug_a.load(std::memory_order_relaxed), g_cg_override = g_cg_override_a.load(std::memory_order_relaxed);   // one snapshot per launch
    int grid = sm_count();
    if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
    // CTA pairs for anything with at least one full wave of 256-row tiles; the single-CTA kernel for small problems
    int cg = g_cg_override ? g_cg_override : ((M > BM && grid >= 2 && static_cast<long>((M + 255) / 256) * ((N + 255) / 256) >= grid / 4) ? 2 : 1);
    if (grid < 2) cg = 1;
    if (cg == 2) grid &= ~1;
    const int units = grid / cg;
    const int BN = pick_bn(M, N, K, mx, b_mn != 0, cg, units);
    LLMFP8_REQUIRE(BN == 128 || BN == 192 || BN == 256 || (BN == 224 && mx), "fp8_gemm: tile width %d is not built", BN);
    LLMFP8_REQUIRE(cg == 1 || cg == 2, "fp8_gemm: cta group %d is not built", cg);

    GemmParams p = {};
    p.d = d; p.ldd = ldd; p.M = M; p.N = N; p.K = K;
    p.a_scale_inv = a_scale_inv; p.b_scale_inv = b_scale_inv; p.bias = bias; p.residual = residual; p.ldr = ldr; p.accumulate = accumulate; p.out_fp32 = out_fp32;
    p.idesc = mx ? ptx::make_idesc_mxf8(a_fmt, b_fmt, a_mn, b_mn, BM * cg, BN, 0, 0) : ptx::make_idesc_f8(a_fmt, b_fmt, a_mn, b_mn, BM * cg, BN);
    p.bn = BN;
    p.debug = g_debug;
    p.trace = g_trace.load(std::memory_order_relaxed);
    const int bn_cta = BN / cg;
    p.b_rows_cta = b_mn ? (bn_cta + 127) / 128 * 128 : bn_cta;
    // MX at 256 columns, staged epilogue, no fused residual: two overlapped accumulators (OVL kernels); debug bit 4 falls back to one
    // (debug bit 5 keeps the fused-residual kernel on a single accumulator)
    p.acc_overlap = (!(g_debug & 16) && mx && BN == 256 && !accumulate && (!residual || !(g_debug & 32))) ? 1 : 0;
    p.ovl_stride = p.acc_overlap ? ((g_debug & 256) ? 192 : 224) : 0;
    p.acc_stages = (mx && BN > 224 && !p.acc_overlap) ? 1 : 2;      // 224 columns: two whole accumulators beside the scale factors
    p.sf_tiles = mx ? (BN == 224 ? 4 : 3) : 0;
    const int stage_bytes = BM * BK + p.b_rows_cta * BK + 512 * p.sf_tiles;
    const int fixed = 1024 /*alignment slack*/ + CTRL_BYTES /*barriers, CLC responses*/ + EPI_BYTES /*epilogue staging*/;
    int stages = (SMEM_BUDGET - fixed) / stage_bytes;
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    if ((g_debug & 64) && stages > 5) stages = 5;      // experiment: the per-tensor kernel on the ring depth the MX kernel gets
    p.stages = stages;
    const int smem_bytes = fixed + stages * stage_bytes;

    CUtensorMap ma, mb, msa, msb, md;
    int rc;
    // output boxes of 32 rows x 128 B for the staged epilogue
    if (out_fp32) rc = make_tensor_map_2d(&md, d, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, N, M, ldd * 4, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&md, d, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, N, M, ldd * 2, 64, 32, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    // K-major operand [rows, K]: inner dim = K, box = {128 k, tile rows}.   MN-major operand [K, rows]: inner = rows, box = {128 rows, 128 k}.
    if (!a_mn) rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, K, M, lda, BK, BM, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, K, lda, 128, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn) rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, K, N, ldb, BK, bn_cta, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, N, K, ldb, 128, BK, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (mx) {
        // swizzled E8M0 tensors viewed as [row tiles][k tiles][128 x u32]: one 512 B tile per (128 rows, 128 k)
        const int64_t kt = (K + 127) / 128;
        rc = make_tensor_map_3d(&msa, sfa, CU_TENSOR_MAP_DATA_TYPE_UINT32, 128, kt, (M + 127) / 128, 512, 512 * kt, 128, 1, 1);
        if (rc) return rc;
        rc = make_tensor_map_3d(&msb, sfb, CU_TENSOR_MAP_DATA_TYPE_UINT32, 128, kt, (N + 127) / 128, 512, 512 * kt, 128, 1, p.sf_tiles - 1);
        if (rc) return rc;
    } else {
        msa = ma;
        msb = mb;
    }

    const int num_tiles = ((M + BM * cg - 1) / (BM * cg)) * ((N + BN - 1) / BN);
    // dynamic (default): one cluster per tile, handed out by cluster launch control; static (debug bit 3, or a caller-imposed CTA
    // budget): persistent grid, round-robin tiles
    p.dynamic = ((g_debug & 8) || max_ctas > 0) ? 0 : 1;
    if (p.dynamic) grid = num_tiles * cg;
    else if (num_tiles * cg < grid) grid = num_tiles * cg;
    if (cg == 2) return dispatch<2>(a_mn, b_mn, mx, ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
    return dispatch<1>(a_mn, b_mn, mx, ma, mb, msa, msb, md, p, grid, smem_bytes, stream);
}

// ------------------------------------------------------------------------------------------------ bf16 GEMM (kind::f16) and the lm_head epilogues
template <int CG>
static int dispatch_bf(bool a_mn, bool b_mn, int epi, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& md, const GemmParams& p, int grid,
                       int smem_bytes, cudaStream_t stream) {
    if (epi == 1) return launch_variant<CG, false, false, false, false, false, true, 1>(ma, mb, ma, mb, md, p, grid, smem_bytes, stream);
    if (epi == 2) return launch_variant<CG, false, false, false, false, false, true, 2>(ma, mb, ma, mb, md, p, grid, smem_bytes, stream);
    if (!a_mn && !b_mn) return launch_variant<CG, false, false, false, false, false, true>(ma, mb, ma, mb, md, p, grid, smem_bytes, stream);
    if (!a_mn && b_mn) return launch_variant<CG, false, true, false, false, false, true>(ma, mb, ma, mb, md, p, grid, smem_bytes, stream);
    if (a_mn && b_mn) return launch_variant<CG, true, true, false, false, false, true>(ma, mb, ma, mb, md, p, grid, smem_bytes, stream);
    set_error("bf16_gemm: operand layout combination (A MN-major, B K-major) is not built");
    return -1;
}

// D[M, N] = A B^T (+ bias) (+ D) with bf16 operands.  lda / ldb: row pitches in ELEMENTS of the stored matrices ([M, K] or, MN-major, [K, M]).
// epi 1 / 2: the cross-entropy epilogues (fprop layout, 256-column tiles so that the partial layout is fixed: ce_partials is [2 * ceil(N / 256), M]).
int launch_bf16_gemm(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* d, int64_t ldd, int out_fp32, int M, int N, int K,
                     const void* bias, int accumulate, int max_ctas, int epi, const int64_t* ce_labels, const float* ce_lse, const float* ce_gscale,
                     void* ce_partials, float* ce_label_logit, int64_t ce_ignore, cudaStream_t stream) {
    LLMFP8_REQUIRE(M > 0 && N > 0 && K > 0, "bf16_gemm: empty problem M=%d N=%d K=%d", M, N, K);
    LLMFP8_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "bf16_gemm: operand pitches must be multiples of 8 elements (lda=%lld ldb=%lld)", (long long)lda, (long long)ldb);
    LLMFP8_REQUIRE((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) % 16 == 0, "bf16_gemm: operands must be 16 B aligned");
    LLMFP8_REQUIRE(epi == 1 || (d && reinterpret_cast<uintptr_t>(d) % 16 == 0 && N % 8 == 0 && ldd % 8 == 0), "bf16_gemm: output must be 16 B aligned, N and ldd multiples of 8");
    LLMFP8_REQUIRE(epi == 0 || (!a_mn && !b_mn && !accumulate && !out_fp32 && !bias && ce_labels), "lm_head epilogues: fprop layout, bf16 output, labels required");
    LLMFP8_REQUIRE(epi != 1 || (ce_partials && ce_label_logit), "lm_head statistics need their output buffers");
    LLMFP8_REQUIRE(epi != 2 || (ce_lse && ce_gscale), "lm_head gradient needs lse and the gradient scale");
    const int g_debug = g_debug_a.load(std::memory_order_relaxed), g_cg_override = g_cg_override_a.load(std::memory_order_relaxed);
    int grid = sm_count();
    if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
    int cg = g_cg_override ? g_cg_override : ((M > BM && grid >= 2 && static_cast<long>((M + 255) / 256) * ((N + 255) / 256) >= grid / 4) ? 2 : 1);
    if (grid < 2) cg = 1;
    if (cg == 2) grid &= ~1;
    const int units = grid / cg;
    const int BN = epi ? 256 : pick_bn(M, N, K, false, b_mn != 0, cg, units);
    GemmParams p = {};
    p.d = d; p.ldd = ldd; p.M = M; p.N = N; p.K = K;
    p.bias = bias; p.accumulate = accumulate; p.out_fp32 = out_fp32;
    p.idesc = ptx::make_idesc_f8(1 /*bf16*/, 1, a_mn, b_mn, BM * cg, BN);      // kind::f16 shares the field layout; format 1 = bf16
    p.bn = BN;
    p.debug = g_debug;
    const int bn_cta = BN / cg;
    p.b_rows_cta = b_mn ? (bn_cta + 63) / 64 * 64 : bn_cta;
    p.acc_overlap = 0;
    p.ovl_stride = 0;
    p.acc_stages = 2;
    p.ce_labels = ce_labels; p.ce_lse = ce_lse; p.ce_gscale = ce_gscale; p.ce_partials = static_cast<float2*>(ce_partials); p.ce_label_logit = ce_label_logit;
    p.ce_ignore = ce_ignore;
    const int stage_bytes = BM * BK + p.b_rows_cta * BK;
    const int fixed = 1024 + CTRL_BYTES + EPI_BYTES;
    int stages = (SMEM_BUDGET - fixed) / stage_bytes;
    if (stages > MAX_STAGES) stages = MAX_STAGES;
    p.stages = stages;
    const int smem_bytes = fixed + stages * stage_bytes;
    CUtensorMap ma, mb, md;
    int rc;
    if (epi == 1) {
        rc = 0;          // no D: the store map is never used (the A map stands in below)
    } else if (out_fp32) {
        rc = make_tensor_map_2d(&md, d, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, N, M, ldd * 4, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B);
    } else {
        rc = make_tensor_map_2d(&md, d, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, N, M, ldd * 2, 64, 32, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    if (rc) return rc;
    // K-major [rows, K]: box = {64 k, tile rows};  MN-major [K, rows]: box = {64 rows, 64 k}
    if (!a_mn) rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, K, M, lda * 2, BK / 2, BM, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&ma, a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, M, K, lda * 2, 64, BK / 2, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (!b_mn) rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, K, N, ldb * 2, BK / 2, bn_cta, CU_TENSOR_MAP_SWIZZLE_128B);
    else rc = make_tensor_map_2d(&mb, b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, N, K, ldb * 2, 64, BK / 2, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    if (epi == 1) md = ma;
    const int num_tiles = ((M + BM * cg - 1) / (BM * cg)) * ((N + BN - 1) / BN);
    p.dynamic = ((g_debug & 8) || max_ctas > 0) ? 0 : 1;
    if (p.dynamic) grid = num_tiles * cg;
    else if (num_tiles * cg < grid) grid = num_tiles * cg;
    if (cg == 2) return dispatch_bf<2>(a_mn != 0, b_mn != 0, epi, ma, mb, md, p, grid, smem_bytes, stream);
    return dispatch_bf<1>(a_mn != 0, b_mn != 0, epi, ma, mb, md, p, grid, smem_bytes, stream);
}

}  // namespace llmfp8
