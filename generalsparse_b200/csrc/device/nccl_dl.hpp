// NCCL reached through dlopen: the library must load on a CPU-only box and on boxes without NCCL; only the optional all-gather of
// C (north star (3): "NCCL over NVLink is used only for an optional C all-gather") needs it.  The handful of declarations below are
// NCCL 2.x's stable C ABI (nccl.h: ncclUniqueId is 128 opaque bytes, ncclUint8 = 1, ncclSuccess = 0).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stddef.h>
#include <stdint.h>

#include <string>

namespace gsd {
namespace nccl {

struct unique_id {
    char internal[128];
};
typedef struct ncclComm *comm_t;
constexpr int SUCCESS = 0, UINT8 = 1;

struct api {
    int (*GetUniqueId)(unique_id *) = nullptr;
    int (*CommInitRank)(comm_t *, int, unique_id, int) = nullptr;
    int (*CommInitAll)(comm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(comm_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    int (*Broadcast)(const void *, void *, size_t, int, int, comm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, comm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int *) = nullptr;
    std::string why; // why loading failed
    bool ok = false;
};

inline const api &load()
{
    static const api a = [] {
        api x;
        void *h = nullptr;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
        }
        if (!h) {
            const char *e = dlerror();
            x.why = std::string("NCCL: libnccl.so.2 not found (") + (e ? e : "dlopen failed") + ")";
            return x;
        }
        bool all = true;
        auto sym = [&](const char *n) {
            void *p = dlsym(h, n);
            if (!p) all = false, x.why = std::string("NCCL: symbol missing: ") + n;
            return p;
        };
        x.GetUniqueId = (int (*)(unique_id *))sym("ncclGetUniqueId");
        x.CommInitRank = (int (*)(comm_t *, int, unique_id, int))sym("ncclCommInitRank");
        x.CommInitAll = (int (*)(comm_t *, int, const int *))sym("ncclCommInitAll");
        x.CommDestroy = (int (*)(comm_t))sym("ncclCommDestroy");
        x.GroupStart = (int (*)())sym("ncclGroupStart");
        x.GroupEnd = (int (*)())sym("ncclGroupEnd");
        x.Broadcast = (int (*)(const void *, void *, size_t, int, int, comm_t, cudaStream_t))sym("ncclBroadcast");
        x.AllGather = (int (*)(const void *, void *, size_t, int, comm_t, cudaStream_t))sym("ncclAllGather");
        x.GetErrorString = (const char *(*)(int))sym("ncclGetErrorString");
        x.GetVersion = (int (*)(int *))sym("ncclGetVersion");
        x.ok = all;
        return x;
    }();
    return a;
}

// Ragged all-gather of row blocks, in place: rank g's rows [cuts[g], cuts[g+1]) already sit at their offset of `full` on rank g; after
// the call every rank holds all rows.  One grouped call of n broadcasts (equal blocks would be one ncclAllGather; nnz-balanced blocks are
// never equal).  Called once per communicator of a group: the caller brackets several communicators with GroupStart / GroupEnd itself.
inline int allgather_rows(const api &a, comm_t comm, cudaStream_t s, void *full, const uint64_t *cuts, int n, uint64_t row_bytes)
{
    for (int g = 0; g < n; g++) {
        const uint64_t bytes = (cuts[g + 1] - cuts[g]) * row_bytes;
        if (!bytes) continue;
        unsigned char *blk = (unsigned char *)full + cuts[g] * row_bytes;
        const int r = a.Broadcast(blk, blk, bytes, UINT8, g, comm, s);
        if (r != SUCCESS) return r;
    }
    return SUCCESS;
}

} // namespace nccl
} // namespace gsd
