// extern "C" surface of libllmfp8_b200.so (declared in include/llmfp8_b200.h) + process-wide plumbing:
// thread-local error string, SM-count cache, TMA descriptor encoder resolved through the runtime.
#include <cstdarg>
#include <atomic>
#include <cstdio>
#include <mutex>

#include "../../include/llmfp8_b200.h"
#include "common.cuh"
#include "tensormap.cuh"

namespace llmfp8 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return kNumSMsDefault;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = kNumSMsDefault;
        cached[dev] = n;
    }
    return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// cuTensorMapEncodeTiled is a DRIVER entry point: unlike the runtime calls around it, it does not bind the device's primary context to the
// calling thread, and returns CUDA_ERROR_INVALID_CONTEXT on a thread that has none yet — e.g. autograd's backward thread when the first thing
// a process ever runs in backward is one of our TMA-fed kernels.  One cheap runtime call per thread establishes it.
static void ensure_context() {
    static thread_local bool ready = false;
    if (!ready) {
        cudaFree(nullptr);
        ready = true;
    }
}

int make_tensor_map_2d(CUtensorMap* map, const void* base, CUtensorMapDataType dtype, int elem_bytes, int64_t dim0, int64_t dim1,
                       int64_t pitch_bytes, int box0, int box1, CUtensorMapSwizzle swizzle) {
    ensure_context();
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
        return -2;
    }
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(dim0), static_cast<cuuint64_t>(dim1)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(pitch_bytes)};
    cuuint32_t box[2] = {static_cast<cuuint32_t>(box0), static_cast<cuuint32_t>(box1)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, dtype, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): base=%p dims=(%lld,%lld) pitch=%lld box=(%d,%d) elem=%d", (int)r, base, (long long)dim0,
                  (long long)dim1, (long long)pitch_bytes, box0, box1, elem_bytes);
        return -3;
    }
    return 0;
}

int make_tensor_map_3d(CUtensorMap* map, const void* base, CUtensorMapDataType dtype, int64_t d0, int64_t d1, int64_t d2, int64_t stride1_bytes,
                       int64_t stride2_bytes, int b0, int b1, int b2) {
    ensure_context();
    EncodeTiledFn fn = encode_fn();
    if (!fn) {
        set_error("cuTensorMapEncodeTiled is not available from the CUDA driver");
        return -2;
    }
    cuuint64_t dims[3] = {static_cast<cuuint64_t>(d0), static_cast<cuuint64_t>(d1), static_cast<cuuint64_t>(d2)};
    cuuint64_t strides[2] = {static_cast<cuuint64_t>(stride1_bytes), static_cast<cuuint64_t>(stride2_bytes)};
    cuuint32_t box[3] = {static_cast<cuuint32_t>(b0), static_cast<cuuint32_t>(b1), static_cast<cuuint32_t>(b2)};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(map, dtype, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(3d) failed (%d): base=%p dims=(%lld,%lld,%lld) strides=(%lld,%lld) box=(%d,%d,%d)", (int)r, base, (long long)d0,
                  (long long)d1, (long long)d2, (long long)stride1_bytes, (long long)stride2_bytes, b0, b1, b2);
        return -3;
    }
    return 0;
}

// kernels (defined in the other translation units)
int launch_cast_amax(const void*, void*, int64_t, int, const float*, float*, cudaStream_t);
int launch_cast_amax_2d(const void*, int64_t, void*, int64_t, int64_t, int, const float*, float*, cudaStream_t);
int launch_cast_amax_bgrad(const void*, void*, int64_t, int64_t, int, const float*, float*, float*, void*, cudaStream_t);
int launch_amax_scale_update(float*, float*, float*, const float*, int, int, int, cudaStream_t);
int launch_rmsnorm_quant_fwd(const void*, const void*, void*, float*, int64_t, int64_t, float, int, const float*, float*, cudaStream_t);
int launch_rmsnorm_rsigma(const void*, float*, int64_t, int64_t, float, cudaStream_t);
int launch_rmsnorm_bwd(const void*, const void*, const void*, const float*, const void*, void*, void*, float*, int64_t, int64_t, cudaStream_t);
int rmsnorm_bwd_grid(int64_t);
int launch_rmsnorm_bwd_noalloc(const void*, const void*, const void*, const float*, const void*, void*, void*, int64_t, int64_t, cudaStream_t);
int64_t rmsnorm_bwd_noalloc_max_floats();
int launch_swiglu_quant_fwd(const void*, void*, int64_t, int64_t, int, const float*, float*, cudaStream_t);
int launch_dswiglu_quant_bwd(const void*, const void*, void*, int64_t, int64_t, int, const float*, float*, float*, void*, cudaStream_t);
int launch_mx_quantize(const void*, int64_t, int64_t, int64_t, void*, void*, void*, void*, const float*, const void*, float*, void*, cudaStream_t);
int launch_mx_swiglu_quantize(const void*, int64_t, int64_t, void*, void*, void*, void*, cudaStream_t);
int launch_mx_dswiglu_quantize(const void*, const void*, int64_t, int64_t, void*, void*, void*, void*, float*, void*, cudaStream_t);
int launch_fp8_gemm(const void*, int64_t, int, int, const void*, int64_t, int, int, void*, int64_t, int, int, int, int, const float*, const float*,
                    const void*, const void*, const void*, const void*, int64_t, int, int, cudaStream_t);
int comm_unique_id(void*);
int comm_init(const void*, int, int, void**);
int comm_allreduce_max_f32(void*, float*, int64_t, cudaStream_t);
int comm_allreduce_sum_bf16(void*, void*, int64_t, cudaStream_t);
int comm_destroy(void*);
int launch_nvls_allreduce_sum_bf16(void*, int64_t, void* const*, int64_t, int, int, int, int, cudaStream_t);
int nvls_grid(int64_t, int, int);
int launch_nvls_amax_scale_update(float*, const void*, float*, float*, const float*, int, int, int, void* const*, int64_t, int, int, int, cudaStream_t);
int nvls_timeouts(unsigned int*);
int nvls_timeouts_async(unsigned int*, cudaStream_t);
int nvls_reset();
int launch_bf16_gemm(const void*, int64_t, int, const void*, int64_t, int, void*, int64_t, int, int, int, int, const void*, int, int, int, const int64_t*,
                     const float*, const float*, void*, float*, int64_t, cudaStream_t);
int launch_adamw_mx(void*, const void*, void*, void*, int64_t, int64_t, float, float, float, float, float, float, float, void*, void*, void*, void*, cudaStream_t);
int launch_ce_reduce(const void*, int, int, const float*, const int64_t*, int64_t, float*, float*, float*, cudaStream_t);
int ce_reduce_ws_floats(int);
void gemm_force_bn(int);
void gemm_force_cg(int);
void gemm_debug(int);
void gemm_trace(unsigned long long*);

}  // namespace llmfp8

using namespace llmfp8;
#define S(stream) static_cast<cudaStream_t>(stream)

extern "C" {

int llmfp8_abi_version(void) { return LLMFP8_ABI_VERSION; }
const char* llmfp8_last_error(void) { return g_err; }
int llmfp8_sm_count(void) { return sm_count(); }
uint64_t llmfp8_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int llmfp8_quant_cast_amax(const void* x, void* q, int64_t n, int fmt, const float* scale, float* amax, void* stream) {
    return launch_cast_amax(x, q, n, fmt, scale, amax, S(stream));
}
int llmfp8_quant_cast_amax_2d(const void* x, int64_t ldx, void* q, int64_t rows, int64_t cols, int fmt, const float* scale, float* amax, void* stream) {
    return launch_cast_amax_2d(x, ldx, q, rows, cols, fmt, scale, amax, S(stream));
}
int llmfp8_quant_cast_amax_bgrad(const void* x, void* q, int64_t rows, int64_t cols, int fmt, const float* scale, float* amax, float* colsum_ws,
                                 void* bgrad, void* stream) {
    return launch_cast_amax_bgrad(x, q, rows, cols, fmt, scale, amax, colsum_ws, bgrad, S(stream));
}
int llmfp8_rmsnorm_quant_fwd(const void* x, const void* gamma, void* q, float* rsigma, int64_t M, int64_t H, float eps, int fmt, const float* scale,
                             float* amax, void* stream) {
    return launch_rmsnorm_quant_fwd(x, gamma, q, rsigma, M, H, eps, fmt, scale, amax, S(stream));
}
int llmfp8_rmsnorm_bwd(const void* dy, const void* x, const void* gamma, const float* rsigma, void* dx, void* dgamma, float* ws, int64_t M, int64_t H,
                       void* stream) {
    return launch_rmsnorm_bwd(dy, x, gamma, rsigma, nullptr, dx, dgamma, ws, M, H, S(stream));
}
int llmfp8_rmsnorm_bwd_add(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma, float* ws,
                           int64_t M, int64_t H, void* stream) {
    return launch_rmsnorm_bwd(dy, x, gamma, rsigma, dres, dx, dgamma, ws, M, H, S(stream));
}
int llmfp8_rmsnorm_bwd_ws_rows(int64_t M) { return rmsnorm_bwd_grid(M); }
int llmfp8_rmsnorm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* rsigma, const void* dres, void* dx, void* dgamma, int64_t M,
                               int64_t H, void* stream) {
    return launch_rmsnorm_bwd_noalloc(dy, x, gamma, rsigma, dres, dx, dgamma, M, H, S(stream));
}
int64_t llmfp8_rmsnorm_bwd_noalloc_max_floats(void) { return rmsnorm_bwd_noalloc_max_floats(); }
int llmfp8_swiglu_quant_fwd(const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax, void* stream) {
    return launch_swiglu_quant_fwd(fc1_out, q, M, I, fmt, scale, amax, S(stream));
}
int llmfp8_dswiglu_quant_bwd(const void* da, const void* fc1_out, void* q, int64_t M, int64_t I, int fmt, const float* scale, float* amax,
                             float* dbias_ws, void* dbias, void* stream) {
    return launch_dswiglu_quant_bwd(da, fc1_out, q, M, I, fmt, scale, amax, dbias_ws, dbias, S(stream));
}
int llmfp8_amax_scale_update(float* amax_history, float* scale, float* scale_inv, const float* fmax, int history_len, int num_tensors, int margin,
                             void* stream) {
    return launch_amax_scale_update(amax_history, scale, scale_inv, fmax, history_len, num_tensors, margin, S(stream));
}
int llmfp8_mx_quantize(const void* x, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col, const float* rsigma,
                       const void* gamma, void* stream) {
    return launch_mx_quantize(x, cols, rows, cols, q_row, sf_row, q_col, sf_col, rsigma, gamma, nullptr, nullptr, S(stream));
}
int llmfp8_mx_quantize_ld(const void* x, int64_t ldx, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col,
                          const float* rsigma, const void* gamma, void* stream) {
    return launch_mx_quantize(x, ldx, rows, cols, q_row, sf_row, q_col, sf_col, rsigma, gamma, nullptr, nullptr, S(stream));
}
int llmfp8_mx_quantize_bgrad(const void* x, int64_t ldx, int64_t rows, int64_t cols, void* q_row, void* sf_row, void* q_col, void* sf_col,
                             float* bgrad_ws, void* bgrad_bf16, void* stream) {
    if (!bgrad_ws || !bgrad_bf16) { set_error("mx_quantize_bgrad: bgrad workspace and output are required"); return -1; }
    return launch_mx_quantize(x, ldx, rows, cols, q_row, sf_row, q_col, sf_col, nullptr, nullptr, bgrad_ws, bgrad_bf16, S(stream));
}
int llmfp8_mx_swiglu_quantize(const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col, void* stream) {
    return launch_mx_swiglu_quantize(fc1_out, M, I, q_row, sf_row, q_col, sf_col, S(stream));
}
int llmfp8_mx_dswiglu_quantize(const void* da, const void* fc1_out, int64_t M, int64_t I, void* q_row, void* sf_row, void* q_col, void* sf_col,
                               float* dbias_ws, void* dbias, void* stream) {
    return launch_mx_dswiglu_quantize(da, fc1_out, M, I, q_row, sf_row, q_col, sf_col, dbias_ws, dbias, S(stream));
}
int llmfp8_rmsnorm_rsigma(const void* x, float* rsigma, int64_t M, int64_t H, float eps, void* stream) {
    return launch_rmsnorm_rsigma(x, rsigma, M, H, eps, S(stream));
}
int64_t llmfp8_mx_scale_bytes(int64_t rows, int64_t k) {
    const int64_t kb = (k + 31) / 32;
    return ((rows + 127) / 128 * 128) * ((kb + 3) / 4 * 4);
}
int llmfp8_fp8_gemm(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd, int out_fp32,
                    int M, int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* bias, int accumulate, int max_ctas,
                    void* stream) {
    return launch_fp8_gemm(a, lda, a_fmt, a_mn, b, ldb, b_fmt, b_mn, d, ldd, out_fp32, M, N, K, a_scale_inv, b_scale_inv, nullptr, nullptr, bias,
                           nullptr, 0, accumulate, max_ctas, S(stream));
}
int llmfp8_fp8_gemm_res(const void* a, int64_t lda, int a_fmt, int a_mn, const void* b, int64_t ldb, int b_fmt, int b_mn, void* d, int64_t ldd, int M,
                        int N, int K, const float* a_scale_inv, const float* b_scale_inv, const void* bias, const void* residual, int64_t ldr,
                        int max_ctas, void* stream) {
    if (!residual) {
        set_error("fp8_gemm_res: residual is required (use llmfp8_fp8_gemm otherwise)");
        return -1;
    }
    return launch_fp8_gemm(a, lda, a_fmt, a_mn, b, ldb, b_fmt, b_mn, d, ldd, 0, M, N, K, a_scale_inv, b_scale_inv, nullptr, nullptr, bias, residual, ldr,
                           0, max_ctas, S(stream));
}
int llmfp8_bf16_gemm(const void* a, int64_t lda, int a_mn, const void* b, int64_t ldb, int b_mn, void* d, int64_t ldd, int out_fp32, int M, int N, int K,
                     const void* bias, int accumulate, int max_ctas, void* stream) {
    return launch_bf16_gemm(a, lda, a_mn, b, ldb, b_mn, d, ldd, out_fp32, M, N, K, bias, accumulate, max_ctas, 0, nullptr, nullptr, nullptr, nullptr, nullptr, 0,
                            S(stream));
}
int llmfp8_adamw_mx_step(void* w, const void* grad, void* exp_avg, void* exp_avg_sq, int64_t rows, int64_t cols, float lr, float beta1, float beta2, float eps,
                         float weight_decay, float bias_correction1, float bias_correction2, void* q_row, void* sf_row, void* q_col, void* sf_col, void* stream) {
    return launch_adamw_mx(w, grad, exp_avg, exp_avg_sq, rows, cols, lr, beta1, beta2, eps, weight_decay, bias_correction1, bias_correction2, q_row, sf_row,
                           q_col, sf_col, S(stream));
}
int llmfp8_lm_head_partials(int V) { return 2 * ((V + 255) / 256); }
int llmfp8_lm_head_ce_stats(const void* x, int64_t ldx, const void* w, int64_t ldw, int M, int V, int H, const int64_t* labels, void* partials,
                            float* label_logit, void* stream) {
    return launch_bf16_gemm(x, ldx, 0, w, ldw, 0, nullptr, 0, 0, M, V, H, nullptr, 0, 0, 1, labels, nullptr, nullptr, partials, label_logit, 0, S(stream));
}
int llmfp8_ce_reduce_ws_floats(int M) { return ce_reduce_ws_floats(M); }
int llmfp8_ce_reduce(const void* partials, int n_partials, int M, const float* label_logit, const int64_t* labels, int64_t ignore_index, float* lse,
                     float* block_ws, float* loss_and_count, void* stream) {
    return launch_ce_reduce(partials, n_partials, M, label_logit, labels, ignore_index, lse, block_ws, loss_and_count, S(stream));
}
int llmfp8_lm_head_dlogits(const void* x, int64_t ldx, const void* w, int64_t ldw, int M, int V, int H, const int64_t* labels, int64_t ignore_index,
                           const float* lse, const float* gscale, void* dlogits, int64_t ldd, void* stream) {
    return launch_bf16_gemm(x, ldx, 0, w, ldw, 0, dlogits, ldd, 0, M, V, H, nullptr, 0, 0, 2, labels, lse, gscale, nullptr, nullptr, ignore_index, S(stream));
}
void llmfp8_gemm_force_bn(int bn) { gemm_force_bn(bn); }
void llmfp8_gemm_force_cg(int cg) { gemm_force_cg(cg); }
void llmfp8_gemm_debug(int flags) { gemm_debug(flags); }
void llmfp8_gemm_trace(unsigned long long* device_buf) { gemm_trace(device_buf); }
int llmfp8_mxfp8_gemm(const void* a, int64_t lda, int a_mn, const void* sfa, const void* b, int64_t ldb, int b_mn, const void* sfb, void* d,
                      int64_t ldd, int out_fp32, int M, int N, int K, const void* bias, int accumulate, int max_ctas, void* stream) {
    if (!sfa || !sfb) {
        set_error("mxfp8_gemm: scale tensors are required");
        return -1;
    }
    return launch_fp8_gemm(a, lda, LLMFP8_E4M3, a_mn, b, ldb, LLMFP8_E4M3, b_mn, d, ldd, out_fp32, M, N, K, nullptr, nullptr, sfa, sfb, bias,
                           nullptr, 0, accumulate, max_ctas, S(stream));
}
int llmfp8_mxfp8_gemm_res(const void* a, int64_t lda, int a_mn, const void* sfa, const void* b, int64_t ldb, int b_mn, const void* sfb, void* d,
                          int64_t ldd, int M, int N, int K, const void* bias, const void* residual, int64_t ldr, int max_ctas, void* stream) {
    if (!sfa || !sfb || !residual) {
        set_error("mxfp8_gemm_res: scale tensors and residual are required");
        return -1;
    }
    return launch_fp8_gemm(a, lda, LLMFP8_E4M3, a_mn, b, ldb, LLMFP8_E4M3, b_mn, d, ldd, 0, M, N, K, nullptr, nullptr, sfa, sfb, bias, residual, ldr, 0,
                           max_ctas, S(stream));
}

int llmfp8_nccl_unique_id(void* id128) { return comm_unique_id(id128); }
int llmfp8_nccl_init(const void* id128, int rank, int world, void** comm) { return comm_init(id128, rank, world, comm); }
int llmfp8_nccl_allreduce_max_f32(void* comm, float* buf, int64_t n, void* stream) { return comm_allreduce_max_f32(comm, buf, n, S(stream)); }
int llmfp8_nccl_allreduce_sum_bf16(void* comm, void* buf, int64_t n, void* stream) { return comm_allreduce_sum_bf16(comm, buf, n, S(stream)); }
int llmfp8_nccl_destroy(void* comm) { return comm_destroy(comm); }
int llmfp8_nvls_allreduce_sum_bf16(void* multicast_ptr, int64_t numel, void* const* signal_pads, int64_t signal_pad_bytes, int rank, int world,
                                   int max_ctas, int timeout_ms, void* stream) {
    return launch_nvls_allreduce_sum_bf16(multicast_ptr, numel, signal_pads, signal_pad_bytes, rank, world, max_ctas, timeout_ms, S(stream));
}
int llmfp8_nvls_grid(int64_t numel, int world, int max_ctas) { return nvls_grid(numel, world, max_ctas); }
int llmfp8_nvls_amax_scale_update(float* amax_history, const void* row0_multicast, float* scale, float* scale_inv, const float* fmax, int history_len,
                                  int num_tensors, int margin, void* const* signal_pads, int64_t signal_pad_bytes, int rank, int world, int timeout_ms,
                                  void* stream) {
    return launch_nvls_amax_scale_update(amax_history, row0_multicast, scale, scale_inv, fmax, history_len, num_tensors, margin, signal_pads,
                                         signal_pad_bytes, rank, world, timeout_ms, S(stream));
}
int llmfp8_nvls_timeouts(unsigned int* count) { return nvls_timeouts(count); }
int llmfp8_nvls_timeouts_async(unsigned int* pinned_host_count, void* stream) { return nvls_timeouts_async(pinned_host_count, S(stream)); }
int llmfp8_nvls_reset(void) { return nvls_reset(); }

}  // extern "C"
