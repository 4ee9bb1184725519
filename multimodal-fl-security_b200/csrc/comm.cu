// Communicator of the C-ABI: the one collective the sharded path needs -- an all-reduce of the
// rank's partial n x n FP64 Gram matrix (Krum / Multi-Krum, krum.py:73-99 computed per column
// shard) or of its 2n+1 FP64 partial row sums and n fp32 maxima (FLTrust, clipping, norm
// bounding: fltrust.py:170-176, differential_privacy.py:225-230) -- over NCCL (NVLink 5 /
// NVSwitch on one 8 x B200 box).  One process per GPU; the 128-byte NCCL unique id is created
// by rank 0 (flagg_comm_unique_id) and handed to the other ranks by whatever the host program
// already has (torch.distributed, MPI, a file).
//
// NCCL is resolved at run time from the process (torch loads libnccl.so.2; RTLD_NOLOAD finds
// that copy) or from the loader path, so libflagg.so has no link-time dependency on it and the
// single-GPU path never touches it.
#include <dlfcn.h>

#include <mutex>

#include "common.cuh"

namespace {

// the slice of nccl.h this file needs (stable since NCCL 2.x)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7, ncclFloat64 = 8 };
enum { ncclSum = 0, ncclMax = 2 };

struct NcclApi {
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    bool ok = false;
    std::string why;
};

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { api.why = dlerror() ? dlerror() : "libnccl.so.2 not found"; return; }
        auto sym = [&](const char* name) -> void* {
            void* p = dlsym(h, name);
            if (!p && api.why.empty()) api.why = std::string("missing symbol ") + name;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
        api.ok = api.why.empty();
    });
    return api;
}

#define FLAGG_NCCL(call)                                                                      \
    do {                                                                                      \
        int r__ = (call);                                                                     \
        if (r__ != ncclSuccess) {                                                             \
            flagg::set_error("%s:%d: %s -> NCCL: %s", __FILE__, __LINE__, #call,              \
                             nccl().GetErrorString ? nccl().GetErrorString(r__) : "error");   \
            return FLAGG_ECUDA;                                                               \
        }                                                                                     \
    } while (0)

}  // namespace

struct flagg_comm {
    ncclComm_t comm = nullptr;
    int nranks = 0, rank = 0, device = 0;
};

using namespace flagg;

extern "C" {

int flagg_comm_unique_id(void* id128) {
    FLAGG_REQUIRE(id128, "flagg_comm_unique_id: NULL argument");
    FLAGG_REQUIRE(nccl().ok, "flagg_comm_unique_id: NCCL unavailable (%s)", nccl().why.c_str());
    ncclUniqueId id;
    FLAGG_NCCL(nccl().GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return FLAGG_OK;
}

int flagg_comm_create(flagg_comm** out, int nranks, int rank, const void* id128) {
    FLAGG_REQUIRE(out && id128, "flagg_comm_create: NULL argument");
    FLAGG_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "flagg_comm_create: bad rank %d of %d", rank, nranks);
    FLAGG_REQUIRE(nccl().ok, "flagg_comm_create: NCCL unavailable (%s)", nccl().why.c_str());
    flagg_comm* c = new flagg_comm;
    c->nranks = nranks;
    c->rank = rank;
    cudaError_t e = cudaGetDevice(&c->device);
    if (e != cudaSuccess) { delete c; set_error("flagg_comm_create: %s", cudaGetErrorString(e)); return FLAGG_ECUDA; }
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    int r = nccl().CommInitRank(&c->comm, nranks, id, rank);
    if (r != ncclSuccess) {
        set_error("flagg_comm_create: ncclCommInitRank -> %s", nccl().GetErrorString(r));
        delete c;
        return FLAGG_ECUDA;
    }
    *out = c;
    return FLAGG_OK;
}

int flagg_comm_destroy(flagg_comm* c) {
    if (!c) return FLAGG_OK;
    if (c->comm && nccl().ok) nccl().CommDestroy(c->comm);
    delete c;
    return FLAGG_OK;
}

int flagg_comm_size(const flagg_comm* c) { return c ? c->nranks : 0; }
int flagg_comm_rank(const flagg_comm* c) { return c ? c->rank : -1; }

int flagg_allreduce_f64(flagg_comm* c, double* buf, size_t count, int op, void* stream) {
    FLAGG_REQUIRE(c && c->comm && (buf || count == 0), "flagg_allreduce_f64: NULL argument");
    FLAGG_REQUIRE(op == FLAGG_SUM || op == FLAGG_MAX, "flagg_allreduce_f64: bad op %d", op);
    if (count == 0) return FLAGG_OK;
    FLAGG_NCCL(nccl().AllReduce(buf, buf, count, ncclFloat64, op == FLAGG_SUM ? ncclSum : ncclMax, c->comm,
                                static_cast<cudaStream_t>(stream)));
    return FLAGG_OK;
}

int flagg_allreduce_f32(flagg_comm* c, float* buf, size_t count, int op, void* stream) {
    FLAGG_REQUIRE(c && c->comm && (buf || count == 0), "flagg_allreduce_f32: NULL argument");
    FLAGG_REQUIRE(op == FLAGG_SUM || op == FLAGG_MAX, "flagg_allreduce_f32: bad op %d", op);
    if (count == 0) return FLAGG_OK;
    FLAGG_NCCL(nccl().AllReduce(buf, buf, count, ncclFloat32, op == FLAGG_SUM ? ncclSum : ncclMax, c->comm,
                                static_cast<cudaStream_t>(stream)));
    return FLAGG_OK;
}

}  // extern "C"
