// Weight-gradient SUM all-reduce in the NVSwitch fabric (SURVEY.md §8(e) exchange 2): the wgrad GEMMs of every rank
// write their partial dW = dY_local^T X_local into the SAME offset of a symmetric, multicast-mapped gradient arena; this kernel then
// has rank r pull the in-switch sum of its 1/world slice through the multicast address (multimem.ld_reduce: the switch reads the
// slice from all GPUs and adds in fp32) and push the rounded bf16 result back through it (multimem.st: the switch writes all GPUs).
// Per gradient byte each GPU's SMs touch 2/world bytes instead of the ring's 2(world-1)/world reads + writes + adds, so a few
// small CTAs keep up with the rate at which backward produces gradients — the SMs, HBM traffic and power NCCL's 24..32 CTAs take from
// the GEMMs is what keeps the 8-GPU step at 6.87x instead of >= 7x.  Stands in for DDP's bucketed all-reduce (train_multi_gpu.py:355-378).
//
// Ranks meet twice per launch on per-CTA signal slots in peer memory (CAS 0 -> 1 on the peer's slot, CAS 1 -> 0 on one's own: the
// slots are back at 0 when the kernel ends, so consecutive launches on one stream reuse them without epochs):
//   entry (relaxed): every rank's wgrad GEMM for this gradient has completed (its reduce kernel is stream-ordered after it);
//   exit  (release / acquire): every rank's multimem.st has landed before any rank's next kernel reads or overwrites the gradient.
// A rank that waits longer than `timeout_ns` (%globaltimer) gives up, counts it in g_nvls_timeouts (llmfp8_nvls_timeouts) and leaves — the
// result is then wrong and the host must raise, but the GPU is never left spinning.  The failure is STICKY: a rank that gave up may have
// left a 1 on a peer's slot, so the signal pads can no longer be trusted; every later launch on this device sees the non-zero counter on
// entry and returns at once (its peers then time out too and poison themselves) until the host has re-zeroed the pads and called
// llmfp8_nvls_reset.  The host polls the counter once per step without synchronising (llmfp8_nvls_timeouts_async).
//
// The same file holds the other exchange of the path (exchange 1): the delayed-scaling amax MAX all-reduce fused with the scale update it
// feeds (K4) — nvls_amax_scale_update_kernel: meet, one multimem.ld_reduce.max.u32 per fp8 tensor slot (non-negative floats order like
// their bit patterns), meet again (every rank has read every row 0 before any rank rolls its history), then the K4 arithmetic of
// quantize.cu on the reduced value.  One launch instead of an NCCL collective + a kernel, twice per step (LLMFP8_NVLS_AMAX=1).
//
// STATUS: compiled for sm_100a; written after the round-1 GPU budget was spent, NOT yet run on hardware -> opt-in (LLMFP8_NVLS_REDUCE=1).
#include "common.cuh"

namespace llmfp8 {

namespace {

// 128 threads x <= 40 registers = 5 K registers and no shared memory of its own: such a CTA fits on an SM BESIDE a resident GEMM CTA (352 threads x <= 168
// registers = 59 K of the 64 K), so the reduce does not wait for GEMM CTAs to drain and takes no SM away from them — only issue slots.
constexpr int kThreads = 128;
constexpr int kUnroll = 4;
constexpr int kMaxWorld = 32;

__device__ unsigned int g_nvls_timeouts = 0;

__device__ __forceinline__ uint32_t cas_relaxed_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.relaxed.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.release.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
    return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
    uint32_t old;
    asm volatile("atom.acquire.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
    return old;
}

__device__ __forceinline__ long long now_ns() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// thread t < world of each CTA pairs with CTA blockIdx.x of rank t.  Returns false on timeout.
__device__ __forceinline__ bool poisoned() { return *reinterpret_cast<volatile unsigned int*>(&g_nvls_timeouts) != 0u; }

template <bool ORDER>
__device__ __forceinline__ bool meet_ranks(uint32_t* const* pads, int rank, int world, long long timeout_ns) {
    bool ok = true;
    if (static_cast<int>(threadIdx.x) < world) {
        const int peer = threadIdx.x;
        uint32_t* put = pads[peer] + static_cast<size_t>(blockIdx.x) * world + rank;    // my slot on the peer
        uint32_t* get = pads[rank] + static_cast<size_t>(blockIdx.x) * world + peer;    // the peer's slot on me
        const long long t0 = now_ns();
        while ((ORDER ? cas_release_sys(put, 0u, 1u) : cas_relaxed_sys(put, 0u, 1u)) != 0u)
            if (timeout_ns > 0 && now_ns() - t0 > timeout_ns) { ok = false; break; }
        while (ok && (ORDER ? cas_acquire_sys(get, 1u, 0u) : cas_relaxed_sys(get, 1u, 0u)) != 1u)
            if (timeout_ns > 0 && now_ns() - t0 > timeout_ns) { ok = false; break; }
        if (!ok) atomicAdd(&g_nvls_timeouts, 1u);
    }
    return ok;
}

__device__ __forceinline__ uint4 multimem_sum_bf16x8(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(mc)
                 : "memory");
    return v;
}
__device__ __forceinline__ void multimem_store16(void* mc, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// mc: multicast address of the gradient (16-byte aligned); nvec: its length in 16-byte vectors (8 bf16 each)
__global__ void __launch_bounds__(kThreads) nvls_allreduce_sum_bf16_kernel(uint4* mc, int64_t nvec, uint32_t* const* pads, int rank, int world,
                                                                           long long timeout_ns) {
    pdl_wait();
    pdl_launch_dependents();
    if (poisoned()) {                       // an earlier meeting on this device timed out: refuse (uniform over the grid's lifetime or not, no thread meets)
        if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(&g_nvls_timeouts, 1u);
        return;
    }
    // CTA-wide verdict without shared memory (the kernel's footprint beside a GEMM CTA is registers and the 1 KB the system reserves)
    const int alive = __syncthreads_and(meet_ranks<false>(pads, rank, world, timeout_ns) ? 1 : 0);
    if (alive) {
        const int64_t per = (nvec + world - 1) / world;
        const int64_t lo = per * rank;
        const int64_t hi = lo + per < nvec ? lo + per : nvec;
        const int64_t stride = static_cast<int64_t>(gridDim.x) * kThreads;
        for (int64_t i = lo + static_cast<int64_t>(blockIdx.x) * kThreads + threadIdx.x; i < hi; i += stride * kUnroll) {
            uint4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; u++)
                if (i + u * stride < hi) v[u] = multimem_sum_bf16x8(mc + i + u * stride);
#pragma unroll
            for (int u = 0; u < kUnroll; u++)
                if (i + u * stride < hi) multimem_store16(mc + i + u * stride, v[u]);
        }
    }
    __syncthreads();                       // every thread's stores are ordered before the releasing CAS of threads < world
    if (alive) meet_ranks<true>(pads, rank, world, timeout_ns);
}

__device__ __forceinline__ uint32_t multimem_max_u32(const void* mc) {
    uint32_t v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.max.u32 %0, [%1];" : "=r"(v) : "l"(mc) : "memory");
    return v;
}

// hist: this rank's amax history [L, T] row-major (row 0 = the amaxes recorded since the last update), symmetric across ranks;
// row0_mc: multicast address of row 0.  After the launch every rank holds the same scale / scale_inv and a rolled history whose newest
// entry is the maximum over ranks.  The update itself restates amax_scale_update_kernel (quantize.cu; oracle: orc_amax_scale_update)
// with the reduced value in place of the local row 0 — kept apart so that the validated kernel's binary does not change.
__global__ void __launch_bounds__(kThreads) nvls_amax_scale_update_kernel(float* __restrict__ hist, const uint32_t* row0_mc, float* __restrict__ scale,
                                                                          float* __restrict__ scale_inv, const float* __restrict__ fmax, int L, int T,
                                                                          float margin_div, uint32_t* const* pads, int rank, int world,
                                                                          long long timeout_ns) {
    pdl_wait();
    pdl_launch_dependents();
    if (poisoned()) {
        if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(&g_nvls_timeouts, 1u);
        return;
    }
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    int alive = __syncthreads_and(meet_ranks<false>(pads, rank, world, timeout_ns) ? 1 : 0);
    uint32_t m = 0;
    if (alive && t < T) m = multimem_max_u32(row0_mc + t);
    __syncthreads();
    if (alive) alive = __syncthreads_and(meet_ranks<true>(pads, rank, world, timeout_ns) ? 1 : 0);     // nobody rolls before everybody has read
    if (!alive || t >= T) return;                      // timed out: state untouched, the host sees llmfp8_nvls_timeouts != 0
    const float first = __uint_as_float(m);
    float am = 0.f;
    bool nonfinite = false;
    for (int l = 0; l < L; l++) {
        const float h = l == 0 ? first : hist[l * T + t];
        nonfinite |= !isfinite(h);
        am = (h > am) ? h : am;
        if (l + 1 < L) hist[l * T + t] = hist[(l + 1) * T + t];
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

}  // namespace

int nvls_grid(int64_t numel, int world, int max_ctas) {
    const int64_t nvec = numel / 8;
    const int64_t per = (nvec + world - 1) / world;
    int64_t ctas = (per + static_cast<int64_t>(kThreads) * kUnroll - 1) / (static_cast<int64_t>(kThreads) * kUnroll);
    const int cap = max_ctas > 0 ? max_ctas : 32;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    return static_cast<int>(ctas);
}

int launch_nvls_allreduce_sum_bf16(void* multicast_ptr, int64_t numel, void* const* signal_pads, int64_t pad_bytes, int rank, int world, int max_ctas,
                                   int timeout_ms, cudaStream_t stream) {
    LLMFP8_REQUIRE(world >= 2 && world <= kMaxWorld && rank >= 0 && rank < world, "nvls_allreduce_sum_bf16: bad rank %d of %d", rank, world);
    LLMFP8_REQUIRE(numel >= 0 && numel % 8 == 0, "nvls_allreduce_sum_bf16: numel %lld must be a multiple of 8 (16-byte vectors)", (long long)numel);
    if (numel == 0) return 0;
    LLMFP8_REQUIRE(multicast_ptr != nullptr && (reinterpret_cast<uintptr_t>(multicast_ptr) & 15) == 0,
                   "nvls_allreduce_sum_bf16: the multicast address must be non-null and 16-byte aligned");
    LLMFP8_REQUIRE(signal_pads != nullptr, "nvls_allreduce_sum_bf16: null signal pads");
    const int ctas = nvls_grid(numel, world, max_ctas);
    LLMFP8_REQUIRE(static_cast<int64_t>(ctas) * world * 4 <= pad_bytes, "nvls_allreduce_sum_bf16: %d CTAs x %d ranks need %lld bytes of signal pad, have %lld",
                   ctas, world, (long long)ctas * world * 4, (long long)pad_bytes);
    // 0 -> 10 s (far beyond any rank skew of a healthy step), < 0 -> wait for ever (what NCCL does)
    const long long timeout_ns = timeout_ms == 0 ? 10000000000LL : (timeout_ms < 0 ? 0LL : static_cast<long long>(timeout_ms) * 1000000LL);
    LLMFP8_LAUNCH(nvls_allreduce_sum_bf16_kernel, dim3(ctas), dim3(kThreads), 0, stream, static_cast<uint4*>(multicast_ptr), numel / 8,
                  reinterpret_cast<uint32_t* const*>(signal_pads), rank, world, timeout_ns);
    LLMFP8_LAUNCH_CHECK("nvls_allreduce_sum_bf16");
    return 0;
}

int launch_nvls_amax_scale_update(float* hist, const void* row0_multicast, float* scale, float* scale_inv, const float* fmax, int L, int T, int margin,
                                  void* const* signal_pads, int64_t pad_bytes, int rank, int world, int timeout_ms, cudaStream_t stream) {
    LLMFP8_REQUIRE(L >= 1 && T >= 0, "nvls_amax_scale_update: bad sizes L=%d T=%d", L, T);
    LLMFP8_REQUIRE(margin >= 0 && margin < 120, "nvls_amax_scale_update: bad margin %d", margin);
    LLMFP8_REQUIRE(world >= 2 && world <= kMaxWorld && rank >= 0 && rank < world, "nvls_amax_scale_update: bad rank %d of %d", rank, world);
    if (T == 0) return 0;
    LLMFP8_REQUIRE(hist && scale && scale_inv && fmax, "nvls_amax_scale_update: null state pointer");
    LLMFP8_REQUIRE(row0_multicast != nullptr && (reinterpret_cast<uintptr_t>(row0_multicast) & 3) == 0,
                   "nvls_amax_scale_update: the multicast address of row 0 must be non-null and 4-byte aligned");
    LLMFP8_REQUIRE(signal_pads != nullptr, "nvls_amax_scale_update: null signal pads");
    const int ctas = (T + kThreads - 1) / kThreads;
    LLMFP8_REQUIRE(static_cast<int64_t>(ctas) * world * 4 <= pad_bytes, "nvls_amax_scale_update: %d CTAs x %d ranks need %lld bytes of signal pad, have %lld",
                   ctas, world, (long long)ctas * world * 4, (long long)pad_bytes);
    const long long timeout_ns = timeout_ms == 0 ? 10000000000LL : (timeout_ms < 0 ? 0LL : static_cast<long long>(timeout_ms) * 1000000LL);
    LLMFP8_LAUNCH(nvls_amax_scale_update_kernel, dim3(ctas), dim3(kThreads), 0, stream, hist, static_cast<const uint32_t*>(row0_multicast), scale, scale_inv,
                  fmax, L, T, ldexpf(1.0f, margin), reinterpret_cast<uint32_t* const*>(signal_pads), rank, world, timeout_ns);
    LLMFP8_LAUNCH_CHECK("nvls_amax_scale_update");
    return 0;
}

int nvls_timeouts(unsigned int* count) {
    LLMFP8_REQUIRE(count != nullptr, "nvls_timeouts: null output");
    cudaError_t e = cudaMemcpyFromSymbol(count, g_nvls_timeouts, sizeof(unsigned int));     // synchronises: diagnostics only
    if (e != cudaSuccess) {
        set_error("nvls_timeouts: %s", cudaGetErrorString(e));
        return static_cast<int>(e);
    }
    return 0;
}

int nvls_timeouts_async(unsigned int* pinned_host, cudaStream_t stream) {
    LLMFP8_REQUIRE(pinned_host != nullptr, "nvls_timeouts_async: null output");
    cudaError_t e = cudaMemcpyFromSymbolAsync(pinned_host, g_nvls_timeouts, sizeof(unsigned int), 0, cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) {
        set_error("nvls_timeouts_async: %s", cudaGetErrorString(e));
        return static_cast<int>(e);
    }
    return 0;
}

int nvls_reset() {
    const unsigned int zero = 0;
    cudaError_t e = cudaMemcpyToSymbol(g_nvls_timeouts, &zero, sizeof(unsigned int));      // synchronises; the caller has re-zeroed the signal pads
    if (e != cudaSuccess) {
        set_error("nvls_reset: %s", cudaGetErrorString(e));
        return static_cast<int>(e);
    }
    return 0;
}

}  // namespace llmfp8
