// The two exchanges of the data-parallel path behind the C ABI, for hosts that do not bring torch.distributed:
//   amax MAX all-reduce (fp32, one flat pool per direction per step) and weight-gradient SUM all-reduce (bf16, in place).
// Replaces what the reference gets from DDP / accelerate (train_multi_gpu.py:355-378) and TE's amax reduction inside fp8_autocast.
// NCCL is resolved at run time (dlopen "libnccl.so.2": the copy PyTorch already mapped if there is one), so the library neither
// links against NCCL nor needs its headers; only the handful of entry points and enum values below, stable since NCCL 2.10.
#include <dlfcn.h>

#include <cstring>
#include <mutex>

#include "common.cuh"

namespace llmfp8 {

namespace {

struct NcclUniqueId { char internal[128]; };
typedef void* NcclComm;
typedef int NcclResult;
constexpr int kNcclFloat32 = 7, kNcclBfloat16 = 9;   // ncclDataType_t
constexpr int kNcclSum = 0, kNcclMax = 2;            // ncclRedOp_t

struct NcclApi {
    NcclResult (*GetUniqueId)(NcclUniqueId*) = nullptr;
    NcclResult (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    NcclResult (*AllReduce)(const void*, void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    NcclResult (*CommDestroy)(NcclComm) = nullptr;
    const char* (*GetErrorString)(NcclResult) = nullptr;
    bool ok = false;
};

NcclApi& api() {
    static NcclApi a;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);   // already in the process (PyTorch's)?
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) return;
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(dlsym(h, "ncclAllReduce"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        a.ok = a.GetUniqueId && a.CommInitRank && a.AllReduce && a.CommDestroy;
    });
    return a;
}

int nccl_fail(const char* what, NcclResult r) {
    NcclApi& a = api();
    set_error("%s: NCCL error %d (%s)", what, r, a.GetErrorString ? a.GetErrorString(r) : "?");
    return 1000 + r;      // > 0 like a cudaError_t, offset so the two cannot be confused
}

bool have_api(const char* what) {
    if (api().ok) return true;
    set_error("%s: libnccl.so.2 could not be loaded (%s)", what, dlerror() ? dlerror() : "symbols missing");
    return false;
}

}  // namespace

int comm_unique_id(void* id128) {
    LLMFP8_REQUIRE(id128 != nullptr, "nccl_unique_id: null output");
    if (!have_api("nccl_unique_id")) return -4;
    NcclUniqueId id;
    NcclResult r = api().GetUniqueId(&id);
    if (r != 0) return nccl_fail("nccl_unique_id", r);
    std::memcpy(id128, &id, sizeof id);
    return 0;
}

int comm_init(const void* id128, int rank, int world, void** comm) {
    LLMFP8_REQUIRE(id128 && comm, "nccl_init: null argument");
    LLMFP8_REQUIRE(world >= 1 && rank >= 0 && rank < world, "nccl_init: bad rank %d of %d", rank, world);
    if (!have_api("nccl_init")) return -4;
    NcclUniqueId id;
    std::memcpy(&id, id128, sizeof id);
    NcclComm c = nullptr;
    NcclResult r = api().CommInitRank(&c, world, id, rank);     // binds to the calling thread's current CUDA device
    if (r != 0) return nccl_fail("nccl_init", r);
    *comm = c;
    return 0;
}

static int comm_allreduce(const char* what, void* comm, void* buf, int64_t n, int dtype, int op, cudaStream_t stream) {
    LLMFP8_REQUIRE(comm != nullptr, "%s: null communicator", what);
    LLMFP8_REQUIRE(n >= 0 && (n == 0 || buf != nullptr), "%s: bad buffer", what);
    if (n == 0) return 0;
    if (!have_api(what)) return -4;
    NcclResult r = api().AllReduce(buf, buf, static_cast<size_t>(n), dtype, op, static_cast<NcclComm>(comm), stream);
    if (r != 0) return nccl_fail(what, r);
    return 0;
}

int comm_allreduce_max_f32(void* comm, float* buf, int64_t n, cudaStream_t stream) {
    return comm_allreduce("nccl_allreduce_max_f32", comm, buf, n, kNcclFloat32, kNcclMax, stream);
}

int comm_allreduce_sum_bf16(void* comm, void* buf, int64_t n, cudaStream_t stream) {
    return comm_allreduce("nccl_allreduce_sum_bf16", comm, buf, n, kNcclBfloat16, kNcclSum, stream);
}

int comm_destroy(void* comm) {
    if (!comm) return 0;
    if (!have_api("nccl_destroy")) return -4;
    NcclResult r = api().CommDestroy(static_cast<NcclComm>(comm));
    if (r != 0) return nccl_fail("nccl_destroy", r);
    return 0;
}

}  // namespace llmfp8
