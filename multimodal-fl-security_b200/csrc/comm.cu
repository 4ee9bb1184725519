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
#include <stdlib.h>

#include <algorithm>

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
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
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
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
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

// Ingest for the parameter-sharded path: the server process that received the clients'
// updates (robust_server.py:99-108) holds whole client vectors on ONE GPU; every rank needs the [n, hi - lo)
// column shard of all rows.  Rank `root` sends, for every row and every peer, the pieces of that peer's column
// range (one piece per tensor the range touches) with ncclSend; the peer posts the matching ncclRecv's into one
// dense matrix dst[n, ld].  Both sides walk rows, peers and pieces in the same order, in NCCL groups of at most
// SC_GROUP_ROWS rows, so the operations pair up; over NVLink / NVSwitch the root's egress is the limit.
// The root's own shard is copied device-to-device when `dst` is given (NULL: the caller keeps views of its rows).
constexpr int SC_GROUP_ROWS = 64;

int flagg_comm_scatter_columns(flagg_comm* c, int root, int n, int P, const int64_t* seg_numel,
                               const void* const* root_seg_ptrs, const int64_t* lo, const int64_t* hi,
                               float* dst, int64_t ld, void* stream) {
    FLAGG_REQUIRE(c && c->comm && seg_numel && lo && hi, "flagg_comm_scatter_columns: NULL argument");
    FLAGG_REQUIRE(n >= 1 && P >= 1 && root >= 0 && root < c->nranks, "flagg_comm_scatter_columns: bad n / P / root");
    const bool is_root = c->rank == root;
    FLAGG_REQUIRE(!is_root || root_seg_ptrs, "flagg_comm_scatter_columns: the root needs the row pointers");
    FLAGG_REQUIRE(is_root || dst, "flagg_comm_scatter_columns: non-root ranks need a destination");
    std::vector<int64_t> start(P + 1, 0);
    for (int p = 0; p < P; ++p) {
        FLAGG_REQUIRE(seg_numel[p] >= 0, "flagg_comm_scatter_columns: negative tensor size");
        start[p + 1] = start[p] + seg_numel[p];
    }
    for (int r = 0; r < c->nranks; ++r)
        FLAGG_REQUIRE(lo[r] >= 0 && lo[r] <= hi[r] && hi[r] <= start[P], "flagg_comm_scatter_columns: bad column range of rank %d", r);
    FLAGG_REQUIRE(ld >= hi[c->rank] - lo[c->rank] || !dst, "flagg_comm_scatter_columns: ld smaller than the shard");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // pieces of column range [a, b): (tensor p, offset inside it, length, offset inside the shard)
    auto for_pieces = [&](int64_t a, int64_t b, auto&& fn) -> int {
        for (int p = 0; p < P; ++p) {
            const int64_t x = std::max(a, start[p]), y = std::min(b, start[p + 1]);
            if (y > x) {
                const int rc = fn(p, x - start[p], y - x, x - a);
                if (rc != FLAGG_OK) return rc;
            }
        }
        return FLAGG_OK;
    };
    int group_rows = SC_GROUP_ROWS;
    if (const char* e = getenv("FLAGG_SCATTER_GROUP_ROWS")) group_rows = std::max(1, atoi(e));
    for (int i0 = 0; i0 < n; i0 += group_rows) {
        const int i1 = std::min(n, i0 + group_rows);
        FLAGG_NCCL(nccl().GroupStart());
        int rc = FLAGG_OK;
        if (is_root) {      // peer-major inside a group: each peer's operations are consecutive, in row order
            for (int r = 0; r < c->nranks && rc == FLAGG_OK; ++r) {
                if (r == root) continue;
                for (int i = i0; i < i1 && rc == FLAGG_OK; ++i)
                    rc = for_pieces(lo[r], hi[r], [&](int p, int64_t off, int64_t len, int64_t) -> int {
                        const float* src = static_cast<const float*>(root_seg_ptrs[size_t(i) * P + p]) + off;
                        FLAGG_NCCL(nccl().Send(src, size_t(len), ncclFloat32, r, c->comm, st));
                        return FLAGG_OK;
                    });
            }
        }
        for (int i = i0; i < i1 && rc == FLAGG_OK; ++i) {
            if (is_root) {
                break;
            } else {
                rc = for_pieces(lo[c->rank], hi[c->rank], [&](int, int64_t, int64_t len, int64_t at) -> int {
                    FLAGG_NCCL(nccl().Recv(dst + int64_t(i) * ld + at, size_t(len), ncclFloat32, root, c->comm, st));
                    return FLAGG_OK;
                });
            }
        }
        const int ge = nccl().GroupEnd();
        if (rc != FLAGG_OK) return rc;
        FLAGG_NCCL(ge);
    }
    if (is_root && dst) {
        for (int i = 0; i < n; ++i) {
            const int rc = for_pieces(lo[root], hi[root], [&](int p, int64_t off, int64_t len, int64_t at) -> int {
                FLAGG_CUDA(cudaMemcpyAsync(dst + int64_t(i) * ld + at, static_cast<const float*>(root_seg_ptrs[size_t(i) * P + p]) + off,
                                           sizeof(float) * size_t(len), cudaMemcpyDeviceToDevice, st));
                return FLAGG_OK;
            });
            if (rc != FLAGG_OK) return rc;
        }
    }
    return FLAGG_OK;
}

}  // extern "C"
