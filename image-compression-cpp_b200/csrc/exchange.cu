// The path's only exchange inside the C ABI: every rank receives all rewards (all-gather) and the
// global {sum, count} (all-reduce), both in ONE NCCL group on the caller's stream.  The reference is single-GPU
// (include/ema_baseline.hpp:13 takes the mean of the whole batch); this is what a sharded caller needs to keep that
// contract.  NCCL is bound at run time from the libnccl.so.2 the process already holds (torch's, or the system's), so the
// library has no link-time dependency on it and a single-GPU user never loads it.
#include "icr_internal.h"
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>

namespace icr {

struct NcclId { char internal[128]; };       // ncclUniqueId
typedef void* Comm;                          // ncclComm_t
enum { NCCL_FLOAT64 = 8, NCCL_SUM = 0 };     // ncclDataType_t / ncclRedOp_t values (nccl.h, stable since 2.0)

struct Nccl {
    void* so = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(Comm*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*CommCount)(Comm, int*) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, Comm, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    std::string why;
};

static Nccl* nccl() {
    static Nccl n;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* env = getenv("ICR_NCCL_LIB");
        if (env && *env) n.so = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
        if (!n.so) n.so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);     // the copy the process already uses (torch's)
        if (!n.so) n.so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!n.so) { n.why = std::string("libnccl.so.2 could not be loaded (") + dlerror() + "); set ICR_NCCL_LIB"; return; }
        auto sym = [&](const char* name) {
            void* p = dlsym(n.so, name);
            if (!p && n.why.empty()) n.why = std::string("libnccl.so.2 has no symbol ") + name;
            return p;
        };
        n.GetErrorString = (decltype(n.GetErrorString))sym("ncclGetErrorString");
        n.GetUniqueId = (decltype(n.GetUniqueId))sym("ncclGetUniqueId");
        n.CommInitRank = (decltype(n.CommInitRank))sym("ncclCommInitRank");
        n.CommDestroy = (decltype(n.CommDestroy))sym("ncclCommDestroy");
        n.CommCount = (decltype(n.CommCount))sym("ncclCommCount");
        n.GroupStart = (decltype(n.GroupStart))sym("ncclGroupStart");
        n.GroupEnd = (decltype(n.GroupEnd))sym("ncclGroupEnd");
        n.AllGather = (decltype(n.AllGather))sym("ncclAllGather");
        n.AllReduce = (decltype(n.AllReduce))sym("ncclAllReduce");
    });
    return &n;
}

int api_fail(int code, const std::string& msg);      // api.cu: sets icr_last_error() of the calling thread

static int nccl_fail(Nccl* n, int rc, const char* what) {
    return api_fail(ICR_ECUDA, std::string("NCCL error in ") + what + ": " + (n->GetErrorString ? n->GetErrorString(rc) : "?"));
}

}  // namespace icr

using namespace icr;

#define ICR_NCCL(n, x, what) do { int r_ = (x); if (r_ != 0) return nccl_fail(n, r_, what); } while (0)

extern "C" {

int icr_nccl_unique_id(unsigned char* id128) {
    if (!id128) return api_fail(ICR_EINVAL, "null ptr");
    Nccl* n = nccl();
    if (!n->why.empty()) return api_fail(ICR_ECUDA, n->why);
    NcclId id;
    ICR_NCCL(n, n->GetUniqueId(&id), "ncclGetUniqueId");
    memcpy(id128, id.internal, 128);
    return ICR_OK;
}

int icr_nccl_comm_create(icr_nccl_comm* comm, const unsigned char* id128, int ranks, int rank) {
    if (!comm || !id128) return api_fail(ICR_EINVAL, "null ptr");
    if (ranks <= 0 || rank < 0 || rank >= ranks) return api_fail(ICR_EINVAL, "rank must lie in [0, ranks)");
    Nccl* n = nccl();
    if (!n->why.empty()) return api_fail(ICR_ECUDA, n->why);
    NcclId id;
    memcpy(id.internal, id128, 128);
    Comm c = nullptr;
    ICR_NCCL(n, n->CommInitRank(&c, ranks, id, rank), "ncclCommInitRank");
    *comm = c;
    return ICR_OK;
}

int icr_nccl_comm_destroy(icr_nccl_comm comm) {
    if (!comm) return ICR_OK;
    Nccl* n = nccl();
    if (!n->why.empty()) return api_fail(ICR_ECUDA, n->why);
    ICR_NCCL(n, n->CommDestroy((Comm)comm), "ncclCommDestroy");
    return ICR_OK;
}

int icr_nccl_comm_ranks(icr_nccl_comm comm, int* ranks) {
    if (!comm || !ranks) return api_fail(ICR_EINVAL, "null ptr");
    Nccl* n = nccl();
    if (!n->why.empty()) return api_fail(ICR_ECUDA, n->why);
    ICR_NCCL(n, n->CommCount((Comm)comm, ranks), "ncclCommCount");
    return ICR_OK;
}

int icr_exchange_rewards(icr_nccl_comm comm, const double* rewards_local, int n_local, double* gathered, double* sum_count,
                         icr_stream stream) {
    if (!comm || !rewards_local) return api_fail(ICR_EINVAL, "null ptr");
    if (!gathered && !sum_count) return api_fail(ICR_EINVAL, "nothing to exchange: gathered and sum_count are both null");
    if (n_local <= 0) return api_fail(ICR_EINVAL, "n_local must be positive");
    Nccl* n = nccl();
    if (!n->why.empty()) return api_fail(ICR_ECUDA, n->why);
    cudaStream_t st = (cudaStream_t)stream;
    if (sum_count) {
        cudaError_t e = launch_sum_count(rewards_local, n_local, sum_count, st);       // this rank's {sum, count}, reduced in place below
        if (e != cudaSuccess) return api_fail(ICR_ECUDA, std::string("CUDA error in sum_count: ") + cudaGetErrorString(e));
    }
    ICR_NCCL(n, n->GroupStart(), "ncclGroupStart");
    int rc_g = 0, rc_r = 0;
    if (gathered) rc_g = n->AllGather(rewards_local, gathered, (size_t)n_local, NCCL_FLOAT64, (Comm)comm, st);
    if (sum_count) rc_r = n->AllReduce(sum_count, sum_count, 2, NCCL_FLOAT64, NCCL_SUM, (Comm)comm, st);
    const int rc_e = n->GroupEnd();
    if (rc_g) return nccl_fail(n, rc_g, "ncclAllGather");
    if (rc_r) return nccl_fail(n, rc_r, "ncclAllReduce");
    if (rc_e) return nccl_fail(n, rc_e, "ncclGroupEnd");
    return ICR_OK;
}

}  // extern "C"
