// Run-time binding to NCCL (no link-time dependency).  The host process may
// already have an NCCL loaded (torch bundles 2.28.x); dlopen by SONAME returns
// that copy, otherwise the system libnccl.so.2 is used.  Only the handful of
// entry points the position exchange needs are bound.  Types are restated here
// (they are ABI-stable across NCCL 2.x) so the build does not need nccl.h.
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <string>

namespace nb {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;  // 0 == ncclSuccess
enum { ncclUint8_ = 1, ncclInt64_ = 4, ncclFloat64_ = 8 };
enum { ncclSum_ = 0 };

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    void* handle = nullptr;
    std::string error;

    bool ok() const { return handle != nullptr; }
};

inline NcclApi& nccl()
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    const char* names[4] = {getenv("NBODY_NCCL_LIB"), "libnccl.so.2", "libnccl.so", nullptr};
    void* h = nullptr;
    for (int i = 0; i < 3 && !h; ++i)
        if (names[i] && names[i][0]) h = dlopen(names[i], RTLD_NOW | RTLD_LOCAL);
    if (!h) {
        const char* e = dlerror();
        api.error = std::string("cannot load libnccl.so.2: ") + (e ? e : "?");
        return api;
    }
    bool all = true;
    auto bind = [&](auto& fn, const char* sym) {
        void* p = dlsym(h, sym);
        if (!p) {
            all = false;
            api.error = std::string("NCCL symbol missing: ") + sym;
        }
        memcpy(&fn, &p, sizeof(p));
    };
    bind(api.GetUniqueId, "ncclGetUniqueId");
    bind(api.CommInitRank, "ncclCommInitRank");
    bind(api.CommInitAll, "ncclCommInitAll");
    bind(api.CommDestroy, "ncclCommDestroy");
    bind(api.AllGather, "ncclAllGather");
    bind(api.AllReduce, "ncclAllReduce");
    bind(api.GroupStart, "ncclGroupStart");
    bind(api.GroupEnd, "ncclGroupEnd");
    bind(api.GetErrorString, "ncclGetErrorString");
    bind(api.GetVersion, "ncclGetVersion");
    if (all) api.handle = h;
    return api;
}

}  // namespace nb
