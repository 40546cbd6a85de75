/*
 * llmfp8_b200.h — C ABI of libllmfp8_b200.so: the B200 (sm_100a) FP8 / MXFP8 linear-layer hot path.
 *
 * This is the boundary a Transformer-Engine-style framework binds: the reference (xuanvinh1997/llm-fp8) reaches
 * these operations through `transformer_engine.pytorch` modules constructed at te_llama.py:45-63 and executed under
 * `te.pytorch.fp8_autocast` at te_llama.py:76-80 (hybrid: te_llama_hybrid.py:75-79, MXFP8: te_llama_mxfp8.py:86-94).
 * TE's own binding layer for them is its `transformer_engine_torch` (tex) extension; each entry point below names
 * the tex operation it stands in for.  INTEGRATION.md shows the ctypes stubs.
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a CUDA device pointer owned by the caller unless noted;
 *   - `stream` is a cudaStream_t passed as void*; every call is asynchronous on it, nothing synchronises;
 *   - return 0 = ok, <0 = contract violation (alignment, shape, format), >0 = cudaError_t;
 *     llmfp8_last_error() returns a thread-local description of the last failure;
 *   - no CPU fallback and no library (cuBLASLt) fallback: an unsupported call is an error;
 *   - re-entrant: no mutable global state besides immutable per-process caches, safe from the autograd thread;
 *   - fp8 format enum: 0 = E4M3 (e4m3fn, max 448), 1 = E5M2 (max 57344);
 *   - bf16 tensors are row-major, 16 B aligned, inner dimension a multiple of 8 (16 for the flat cast).
 */
#ifndef LLMFP8_B200_H
#define LLMFP8_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LLMFP8_E4M3 0
#define LLMFP8_E5M2 1
#define LLMFP8_ABI_VERSION 1

int llmfp8_abi_version(void);
const char* llmfp8_last_error(void);
/* number of SMs of the current device (cached) */
int llmfp8_sm_count(void);
/* number of kernels this library has launched in this process (monotonic; bench.py reports the per-step delta) */
uint64_t llmfp8_launch_count(void);

/* ---- per-tensor (DelayedScaling) quantise kernels ------------------------------------------------------------ */

/* K2: q = sat_rne_fp8(x * *scale), *amax = max(*amax, max|x|).   [tex.cast_to_fp8 / quantize]
 * x bf16 [n] (n % 16 == 0), q u8 [n]; scale, amax fp32 device scalars (amax may be NULL). */
int llmfp8_quant_cast_amax(const void* x, void* q, int64_t n, int fmt, const float* scale, float* amax, void* stream);

/* K2 on a column slice: x is [rows, cols] with row pitch ldx elements (ldx >= cols, ldx % 8 == 0, cols % 16 == 0), q dense [rows, cols].
 * Lets Linear consume a view such as qkv[:, :q_out] (te_llama.py:45-56: the attention output / query slice) without a copy. */
int llmfp8_quant_cast_amax_2d(const void* x, int64_t ldx, void* q, int64_t rows, int64_t cols, int fmt, const float* scale, float* amax, void* stream);

/* K2b: as K2 on a [rows, cols] matrix, plus bgrad[c] = sum_r x[r, c] (bf16 out).   [tex.fp8_cast_transpose_bgrad_fused]
 * colsum_ws: fp32 [cols] scratch. */
int llmfp8_quant_cast_amax_bgrad(const void* x, void* q, int64_t rows, int64_t cols, int fmt, const float* scale, float* amax,
                                 float* colsum_ws, void* bgrad, void* stream);

/* K1: y = (x * rsigma) * gamma with rsigma = 1/sqrt(mean(x^2) + eps); q = sat_rne_fp8(y * *scale); amax over |y|.
 * [tex.rmsnorm_fwd with fp8 output, inside LayerNormLinear / LayerNormMLP]   x bf16 [M, H], gamma bf16 [H], H % 8 == 0, H <= 8192. */
int llmfp8_rmsnorm_quant_fwd(const void* x, const void* gamma, void* q, float* rsigma, int64_t M, int64_t H, float eps, int fmt,
                             const float* scale, float* amax, void* stream);

/* K5: RMSNorm backward.   [tex.rmsnorm_bwd]   dx bf16 [M, H], dgamma bf16 [H]; ws fp32 [llmfp8_rmsnorm_bwd_ws_rows(M) * H]. */
int llmfp8_rmsnorm_bwd(const void* dy, const void* x, const void* gamma, const float* rsigma, void* dx, void* dgamma, float* ws, int64_t M,
                       int64_t H, void* stream);
int llmfp8_rmsnorm_bwd_ws_rows(int64_t M);
/* K5 with the residual-branch gradient folded in: dx = rmsnorm_bwd(...) + dres (dres bf16 [M, H]) — the `h + f(norm(h))` pattern of
 * TELlamaDecoderLayer.forward (te_llama.py:78, 81) then needs no separate gradient-accumulation pass. */
int llmfp8_rmsnorm_bwd_add(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma, float* ws,
                           int64_t M, int64_t H, void* stream);

/* K5 for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
 * reference.  Same results as llmfp8_rmsnorm_bwd (dres == NULL) / llmfp8_rmsnorm_bwd_add (dres != NULL), but
 * the fp32 workspace is a static device array inside the library (llmfp8_rmsnorm_bwd_noalloc_max_floats() = 592 * 4096 floats, 9.25 MiB, held by every GPU
 * that loads the library until exit) instead of a caller buffer, and the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limit: llmfp8_rmsnorm_bwd_ws_rows(M) * H <= that maximum (any M at H <= 4096); larger problems return -1 with a message in llmfp8_last_error().
 * The device of `stream` must be the calling thread's current device.  All calls on a device share the array: calls that OVERLAP on the GPU (different
 * streams) race on it and must be serialised by the caller; the array keeps the previous call's partial sums, which the kernel overwrites before reading. */
int llmfp8_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma, int64_t M,
                               int64_t H, void* stream);
int64_t llmfp8_rmsnorm_bwd_noalloc_max_floats(void);

/* K3: a = silu(gate) * up over fc1_out bf16 [M, 2I] (gate = [:, :I], up = [:, I:]); q u8 [M, I].   [tex.swiglu + cast] */
int llmfp8_swiglu_quant_fwd(const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax, void* stream);

/* K3 bwd: d_fc1 = dswiglu(da, fc1_out) quantised to q u8 [M, 2I]; dbias bf16 [2I] = column sums (dbias_ws fp32 [2I]; both NULL = skip).
 * [tex.dswiglu + fp8_cast_transpose_bgrad_dgelu-style fusion] */
int llmfp8_dswiglu_quant_bwd(const void* da, const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax,
                             float* dbias_ws, void* dbias, void* stream);

/* K4: for every fp8 tensor slot t: amax = max_l hist[l, t]; scale = fmax[t] / amax / 2^margin (unchanged when amax is 0 or
 * non-finite); scale_inv = 1 / scale; roll the history by one and zero row 0.   [tex.fused_amax_and_scale_update_after_reduction] */
int llmfp8_amax_scale_update(float* amax_history, float* scale, float* scale_inv, const float* fmax, int history_len, int num_tensors,
                             int margin, void* stream);

/* ---- MXFP8 (MXFP8BlockScaling) quantiser ---------------------------------------------------------------------- */

/* K2m / K1m: x bf16 [rows, cols] -> rowwise copy (1x32 blocks along cols) and/or colwise copy (32x1 blocks along rows), e4m3 data in
 * the layout of x, E8M0 scales in the 128x4-tile swizzled layout (llmfp8_mx_scale_bytes).  A NULL q pointer skips that copy.
 * rsigma/gamma non-NULL: quantise y = (x * rsigma[r]) * gamma[c] instead (RMSNorm fused; rsigma from llmfp8_rmsnorm_rsigma).
 * [tex.quantize with an MXFP8 quantizer; rmsnorm_fwd with MXFP8 output] */
int llmfp8_mx_quantize(const void* x, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col, const float* rsigma,
                       const void* gamma, void* stream);
/* same with an explicit row pitch of x (elements; >= cols, % 8 == 0): quantises a column slice of a wider matrix in place */
int llmfp8_mx_quantize_ld(const void* x, int64_t ldx, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col,
                          const float* rsigma, const void* gamma, void* stream);
/* llmfp8_mx_quantize_ld of a gradient dY [rows, cols] (plain mode, colwise copy required) that also emits the bias gradient: bgrad_ws
 * fp32 [cols] workspace, bgrad_bf16 [cols] = column sums of dY — taken from the tiles while they are in shared memory instead of a second
 * pass over dY.  [tex.bgrad_quantize with an MXFP8 quantizer] */
int llmfp8_mx_quantize_bgrad(const void* x, int64_t ldx, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col,
                             float* bgrad_ws, void* bgrad_bf16, void* stream);
int llmfp8_rmsnorm_rsigma(const void* x, float* rsigma, int64_t M, int64_t H, float eps, void* stream);
/* K3m: SwiGLU + MXFP8 quantiser in one pass.  fc1_out bf16 [M, 2I] (gate | up) -> a = bf16(silu(gate) * up) [M, I], never written to
 * memory -> rowwise and/or colwise MXFP8 copies of a (layouts as llmfp8_mx_quantize).  Bit-identical to llmfp8_swiglu_quant_fwd(fmt = bf16)
 * followed by llmfp8_mx_quantize.  I % 32 == 0; colwise needs M % 32 == 0.
 * [LayerNormMLP(activation="swiglu") under MXFP8BlockScaling: TE's swiglu + mxfp8 cast, te_llama_mxfp8.py:93]
 * Status: not yet run on hardware (written after the round-1 GPU budget was spent); the Python host selects it only with
 * LLMFP8_MX_FUSED_ACT=1. */
int llmfp8_mx_swiglu_quantize(const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col, void* stream);
/* K3m backward: da bf16 [M, I], fc1_out bf16 [M, 2I] -> dfc1 = bf16(dgate | dup) [M, 2I], never written to memory -> rowwise and/or
 * colwise MXFP8 copies of dfc1; dbias[2I] = column sums of the fp32 (dgate | dup) (dbias_ws fp32 [2I] workspace and dbias bf16 [2I] go
 * together, both NULL = no bias gradient).  Bit-identical (data, scales) to llmfp8_dswiglu_quant_bwd(fmt = bf16) + llmfp8_mx_quantize.
 * I % 128 == 0; colwise needs M % 32 == 0.  Same status as llmfp8_mx_swiglu_quantize. */
int llmfp8_mx_dswiglu_quantize(const void* da, const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col,
                               float* dbias_ws, void* dbias, void* stream);
/* bytes of the swizzled scale tensor of a matrix with `rows` rows and `k` elements along the blocked dimension */
int64_t llmfp8_mx_scale_bytes(int64_t rows, int64_t k);

/* ---- GEMMs ------------------------------------------------------------------------------------------------------ */

/* D[M, N] = (a_scale_inv * b_scale_inv) * sum_k A[m, k] B[n, k] (+ bias[n]) (+ D if accumulate).   [tex.generic_gemm -> cuBLASLt]
 *  a_mn = 0: A stored [M, K] (k contiguous, lda >= K bytes);  a_mn = 1: A stored [K, M] (m contiguous, lda >= M).  Likewise b / N.
 *  out_fp32 = 0: D bf16, 1: D fp32; ldd in elements.  max_ctas <= 0: one CTA per SM. */
int llmfp8_fp8_gemm(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd,
                    int out_fp32, int M, int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* bias, int accumulate,
                    int max_ctas, void* stream);

/* D bf16 [M, N] = (a_scale_inv * b_scale_inv) * A B^T (+ bias[n]) + R[m, n]: the residual add of `h + proj(...)` / `h + fc2(...)`
 * (te_llama.py:78, 81) fused into the epilogue — R bf16 with leading dimension ldr, added in fp32 before the single bf16 rounding. */
int llmfp8_fp8_gemm_res(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd, int M,
                        int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* bias, const void* residual, int64_t ldr,
                        int max_ctas, void* stream);

/* bf16 GEMM on the same tcgen05 pipeline (kind::f16, fp32 accumulate): D[M, N] = A B^T (+ bias[n]) (+ D).  lda / ldb: row pitches in ELEMENTS
 * (multiples of 8) of the stored matrices — [M, K] / [N, K], or [K, M] / [K, N] when a_mn / b_mn (MN-major: dgrad and wgrad read the forward's
 * tensors in place).  The one bf16 linear on the reference's step is the lm_head (HF nn.Linear outside TE: train_fp8.py:278). */
int llmfp8_bf16_gemm(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* d, int64_t ldd, int out_fp32, int M, int N, int K,
                     const void* bias, int accumulate, int max_ctas, void* stream);
/* lm_head + cross-entropy without materialising the [M, V] logits in forward:
 *   llmfp8_lm_head_ce_stats : logits = x[M, H] w[V, H]^T in the GEMM, rounded to bf16 as HF's lm_head output is; the epilogue keeps, per row and
 *                             per (256-column tile, warp half), the maximum and the sum of exponentials -> partials float2 [llmfp8_lm_head_partials(V), M],
 *                             and the label's logit -> label_logit [M] (rows whose label is outside [0, V) are left untouched);
 *   llmfp8_ce_reduce        : lse [M], loss_and_count = {mean over rows with label != ignore_index of (lse - label_logit), number of such rows};
 *                             block_ws: llmfp8_ce_reduce_ws_floats(M) floats; deterministic summation order;
 *   llmfp8_lm_head_dlogits  : the GEMM again with d(logits)[M, V] = (exp(logit - lse) - [col == label]) * gscale[0] (0 for ignored rows), bf16;
 *                             dX and dW are then llmfp8_bf16_gemm(dlogits, w MN-major) and llmfp8_bf16_gemm(dlogits MN-major, x MN-major).
 * [HF LlamaForCausalLM: lm_head + CrossEntropyLoss on the bf16 logits upcast to fp32] */
int llmfp8_lm_head_partials(int V);
int llmfp8_lm_head_ce_stats(const void* x, int64_t ldx, const void* w, int64_t ldw, int M, int V, int H, const int64_t* labels, void* partials,
                            float* label_logit, void* stream);
int llmfp8_ce_reduce_ws_floats(int M);
int llmfp8_ce_reduce(const void* partials, int n_partials, int M, const float* label_logit, const int64_t* labels, int64_t ignore_index, float* lse,
                     float* block_ws, float* loss_and_count, void* stream);
int llmfp8_lm_head_dlogits(const void* x, int64_t ldx, const void* w, int64_t ldw, int M, int V, int H, const int64_t* labels, int64_t ignore_index,
                           const float* lse, const float* gscale, void* dlogits, int64_t ldd, void* stream);

/* Fused AdamW + MXFP8 requantise: one pass over a GEMM weight [rows, cols] (bf16; rows, cols % 32 == 0) that applies
 * torch.optim.AdamW's update (decoupled weight decay; the reference's optimizer, train_fp8.py:200-204) with bf16 state exp_avg / exp_avg_sq — all
 * three updated in place, fp32 arithmetic — and emits the MXFP8 copies of the UPDATED weight (q_row / q_col + swizzled E8M0 scales, layouts of
 * llmfp8_mx_quantize; either may be NULL), bit-identical to llmfp8_mx_quantize of it: the next forward finds its weight casts done.
 * bias_correction{1,2} = 1 - beta{1,2}^step, computed by the host. */
int llmfp8_adamw_mx_step(void* w, const void* grad, void* exp_avg, void* exp_avg_sq, int64_t rows, int64_t cols, float lr, float beta1, float beta2, float eps,
                         float weight_decay, float bias_correction1, float bias_correction2, void* q_row, void* sf_row, void* q_col, void* sf_col, void* stream);

/* Tuning / tests only: force the GEMM tile width (128 | 192 | 256; 0 = automatic).  The three knobs below are process-wide, so they
 * are INERT (non-zero values are ignored, llmfp8_last_error says so) unless the process was started with LLMFP8_ALLOW_TUNING=1 — the test
 * suite and tools/ set it, a training process does not; values are atomics (a racing launch sees the old or the new value). */
void llmfp8_gemm_force_bn(int bn);
/* likewise the CTA group: 1 = single-CTA kernel, 2 = CTA pairs (cta_group::2), 0 = automatic */
void llmfp8_gemm_force_cg(int cg);
/* experiments only: bit 0 skips the MX scale smem->TMEM copies, bit 1 the scale loads (results become wrong); bit 3 = static
 * persistent grid instead of cluster launch control, bit 4 = single TMEM accumulator for MXFP8 at 256-column tiles (results unchanged),
 * bit 5 = single accumulator under the fused-residual MXFP8 epilogue (default: overlapped accumulators there too), bit 6 = at most 5
 * ring stages, bit 7 = epilogue warps sleep between polls of their accumulator barrier, bit 8 = overlapped MX accumulators 192 instead of 224 columns apart */
void llmfp8_gemm_debug(int flags);
/* experiments only (LLMFP8_ALLOW_TUNING=1): device buffer of 64 x 16 u64 that the first cluster of every following FP8 GEMM launch fills with clock64
 * stamps of its MMA thread (slots 0-4: tile start, accumulator free, shared columns free, first stage full, last MMA issued) and of two epilogue
 * warps (slots 5-8 and 10-13: accumulator full, first / second chunk loaded, tile done); NULL switches it off (tools/gemm_trace.py) */
void llmfp8_gemm_trace(unsigned long long* device_buf);

/* Block-scaled variant: sfa / sfb are swizzled E8M0 tensors whose blocks run along k.   [tex.generic_gemm with MXFP8 tensors] */
int llmfp8_mxfp8_gemm(const void* a, int64_t lda, int a_mn, const void* sfa, const void* b, int64_t ldb, int b_mn, const void* sfb, void* d,
                      int64_t ldd, int out_fp32, int M, int N, int K, const void* bias, int accumulate, int max_ctas, void* stream);

/* block-scaled GEMM with the fused residual (see llmfp8_fp8_gemm_res) */
int llmfp8_mxfp8_gemm_res(const void* a, int64_t lda, int a_mn, const void* sfa, const void* b, int64_t ldb, int b_mn, const void* sfb, void* d,
                          int64_t ldd, int M, int N, int K, const void* bias, const void* residual, int64_t ldr, int max_ctas, void* stream);

/* ---- data-parallel exchanges ---------------------------------------------------------------- */

/* The only two collectives on the path, for hosts without torch.distributed.  NCCL is dlopen'ed ("libnccl.so.2") on first use.
 * Stand in for DDP's gradient all-reduce (train_multi_gpu.py:355-378) and TE's amax reduction at fp8_autocast exit (te_llama.py:76-80).
 *   llmfp8_nccl_unique_id : rank 0 fills a 128-byte id; the host hands it to every rank (env rendezvous, file, MPI ...).
 *   llmfp8_nccl_init      : collective; binds the communicator to the calling thread's current CUDA device; *comm is opaque.
 *   ..._allreduce_max_f32 : in place, n floats — the delayed-scaling amax pool, once per direction per step.
 *   ..._allreduce_sum_bf16: in place, n bf16 — a weight gradient, straight out of the wgrad GEMM (scale the loss by 1/world for a mean:
 *                           SUM is the reduction NVSwitch runs in the fabric).
 * Return 0, < 0 (contract; -4 = NCCL not loadable) or 1000 + ncclResult_t. */
int llmfp8_nccl_unique_id(void* id128);
int llmfp8_nccl_init(const void* id128, int rank, int world, void** comm);
int llmfp8_nccl_allreduce_max_f32(void* comm, float* buf, int64_t n, void* stream);
int llmfp8_nccl_allreduce_sum_bf16(void* comm, void* buf, int64_t n, void* stream);
int llmfp8_nccl_destroy(void* comm);

/* Weight-gradient SUM all-reduce inside the NVSwitch fabric (multimem.ld_reduce / multimem.st), in place, for gradients that every
 * rank keeps at the SAME offset of a symmetric, multicast-mapped buffer (the host maps it: cuMulticastCreate / cuMemMap, or
 * torch.distributed._symmetric_memory).  Replaces the bucketed NCCL all-reduce DDP runs behind backward (train_multi_gpu.py:355-378)
 * for the large gradients: each rank reduces 1/world of the buffer through the switch and broadcasts it, on <= max_ctas CTAs of 128 threads (0 = 32),
 * small enough to sit beside a resident GEMM CTA on the same SM.
 *   multicast_ptr    : multicast (not unicast) address of the gradient, 16-byte aligned; numel bf16, a multiple of 8.
 *   signal_pads      : DEVICE array of `world` pointers, entry r = rank r's zero-initialised signal pad (peer-mapped), each
 *                      signal_pad_bytes >= 4 * world * llmfp8_nvls_grid(numel, world, max_ctas); the slots are 0 again on return,
 *                      so one pad serves every launch of a stream.  All ranks must issue the same launches in the same order.
 *   timeout_ms       : a rank whose peers do not arrive gives up after this long (0 = 10 s, < 0 = never) and counts it;
 *                      llmfp8_nvls_timeouts (synchronises the device) returns the count — non-zero means results are invalid.
 * Opt-in through dp.GradAllReducer(nvls=...) / LLMFP8_NVLS_REDUCE=1. */
int llmfp8_nvls_allreduce_sum_bf16(void* multicast_ptr, int64_t numel, void* const* signal_pads, int64_t signal_pad_bytes, int rank, int world,
                                   int max_ctas, int timeout_ms, void* stream);
int llmfp8_nvls_grid(int64_t numel, int world, int max_ctas);
/* llmfp8_nccl_allreduce_max_f32 + llmfp8_amax_scale_update in ONE launch, for an amax history [history_len, num_tensors] that every rank
 * keeps in symmetric, multicast-mapped memory: each slot's row-0 amax is maximised over ranks in the switch (multimem.ld_reduce.max.u32 on
 * the bit patterns of the non-negative floats) and the scale / scale_inv / history roll of llmfp8_amax_scale_update follow in the same
 * kernel — what TE does at fp8_autocast exit (te_llama.py:76-80) with an NCCL collective and a separate kernel.  row0_multicast: multicast
 * address of the history's row 0; signal pads, timeout and the not-yet-run-on-hardware status as above (opt-in: LLMFP8_NVLS_AMAX=1). */
int llmfp8_nvls_amax_scale_update(float* amax_history, const void* row0_multicast, float* scale, float* scale_inv, const float* fmax, int history_len,
                                  int num_tensors, int margin, void* const* signal_pads, int64_t signal_pad_bytes, int rank, int world, int timeout_ms,
                                  void* stream);
int llmfp8_nvls_timeouts(unsigned int* count);
/* A timed-out meeting is sticky: every later llmfp8_nvls_* launch on the device returns at once (and counts itself) until the host has
 * re-zeroed the signal pads and called llmfp8_nvls_reset.  llmfp8_nvls_timeouts_async copies the counter into PINNED host memory on
 * `stream` without synchronising — the once-per-step poll dp.GradAllReducer.finish() and fp8.ScalePool.flush() use. */
int llmfp8_nvls_timeouts_async(unsigned int* pinned_host_count, void* stream);
int llmfp8_nvls_reset(void);

#ifdef __cplusplus
}
#endif
#endif /* LLMFP8_B200_H */
