// lm_head + cross-entropy loss: the bf16 [M, H] x [V, H]^T head the reference runs outside TE every step
// (HF LlamaForCausalLM.lm_head + CrossEntropyLoss reached from train_fp8.py:278) without ever materialising the [M, V] logits in forward:
//   1. lm_head_ce_stats  — the bf16 tcgen05 GEMM of gemm.cu with epilogue 1: per row and per (column tile, warp half) the maximum and the
//                          sum of exponentials of the bf16-rounded logits, plus the label's logit;
//   2. ce_reduce         — this file: folds a row's partials into its log-sum-exp, loss_row = lse - logit[label], deterministic mean over
//                          the rows whose label is not `ignore_index`;
//   3. lm_head_dlogits   — the same GEMM with epilogue 2: d(logits) = (softmax - onehot) * (upstream gradient / counted rows), bf16,
//                          from which dX and dW follow as two more bf16 GEMMs (llmfp8_bf16_gemm, dgrad / wgrad layouts).
#include "common.cuh"

namespace llmfp8 {

constexpr int CE_THREADS = 128;

__global__ void __launch_bounds__(CE_THREADS) ce_reduce_kernel(const float2* __restrict__ partials, int n_part, int M, const float* __restrict__ label_logit,
                                                               const int64_t* __restrict__ labels, int64_t ignore, float* __restrict__ lse,
                                                               float2* __restrict__ block_ws) {
    pdl_wait();
    pdl_launch_dependents();
    const int row = blockIdx.x * CE_THREADS + threadIdx.x;
    float loss = 0.f, cnt = 0.f;
    if (row < M) {
        float m = -INFINITY, s = 0.f;
        for (int j = 0; j < n_part; j++) {                 // consecutive threads read consecutive rows of partial j: coalesced
            const float2 p = partials[static_cast<int64_t>(j) * M + row];
            if (p.y > 0.f) {
                const float nm = fmaxf(m, p.x);
                s = s * expf(m - nm) + p.y * expf(p.x - nm);
                m = nm;
            }
        }
        const float l = m + logf(s);
        lse[row] = l;
        if (labels[row] != ignore) {
            loss = l - label_logit[row];
            cnt = 1.f;
        }
    }
    __shared__ float2 red[CE_THREADS / 32];
    loss = warp_sum(loss);
    cnt = warp_sum(cnt);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = make_float2(loss, cnt);
    __syncthreads();
    if (threadIdx.x == 0) {
        float2 t = red[0];
        for (int w = 1; w < CE_THREADS / 32; w++) { t.x += red[w].x; t.y += red[w].y; }
        block_ws[blockIdx.x] = t;
    }
}

// fixed summation order over the blocks: the loss is reproducible run to run
__global__ void ce_finalize_kernel(const float2* __restrict__ block_ws, int nblocks, float* __restrict__ out /* loss, counted rows */) {
    pdl_wait();
    pdl_launch_dependents();
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        double s = 0.0, c = 0.0;
        for (int i = 0; i < nblocks; i++) { s += block_ws[i].x; c += block_ws[i].y; }
        out[0] = c > 0.0 ? static_cast<float>(s / c) : 0.f;
        out[1] = static_cast<float>(c);
    }
}

int ce_reduce_ws_floats(int M) { return 2 * ((M + CE_THREADS - 1) / CE_THREADS); }

int launch_ce_reduce(const void* partials, int n_part, int M, const float* label_logit, const int64_t* labels, int64_t ignore, float* lse, float* block_ws,
                     float* out, cudaStream_t stream) {
    LLMFP8_REQUIRE(partials && label_logit && labels && lse && block_ws && out, "ce_reduce: null pointer");
    LLMFP8_REQUIRE(n_part > 0 && M > 0, "ce_reduce: empty problem");
    const int nblocks = (M + CE_THREADS - 1) / CE_THREADS;
    LLMFP8_LAUNCH(ce_reduce_kernel, nblocks, CE_THREADS, 0, stream, static_cast<const float2*>(partials), n_part, M, label_logit, labels, ignore, lse,
                  reinterpret_cast<float2*>(block_ws));
    LLMFP8_LAUNCH_CHECK("ce_reduce");
    LLMFP8_LAUNCH(ce_finalize_kernel, 1, 32, 0, stream, reinterpret_cast<const float2*>(block_ws), nblocks, out);
    LLMFP8_LAUNCH_CHECK("ce_finalize");
    return 0;
}

}  // namespace llmfp8
