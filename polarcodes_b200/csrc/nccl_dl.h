// NCCL, bound at run time.  The decode path has no data-path collective (frames are independent,
// SURVEY.md section 8(e)); NCCL only sums the three error counters of a Monte-Carlo point across
// the ranks of one box (the reference's `er` counter, PolarCodes.cpp:129,165,174,183,193).  The
// library is therefore not a link-time dependency: libnccl.so.2 is opened when the first
// communicator is created (the copy already in the process — e.g. the one torch loaded — wins),
// and a single-GPU user never needs it.  Only the five entry points below are used; their ABI has
// been stable across NCCL 2.x.
#pragma once
#include <dlfcn.h>
#include <cuda_runtime.h>

namespace pcb {

struct NcclApi {
    struct UniqueId { char internal[128]; };          // ncclUniqueId
    typedef void* Comm;                                 // ncclComm_t
    enum { kUint64 = 5 /* ncclUint64 */, kSum = 0 /* ncclSum */ };
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*CommAbort)(Comm) = nullptr;                   // optional: teardown that does not wait for the other ranks
    const char* (*GetErrorString)(int) = nullptr;
    void* so = nullptr;

    // returns nullptr on success, else what is missing
    const char* load() {
        if (so) return nullptr;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            so = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);     // already in the process?
            if (so) break;
        }
        for (const char* n : names) {
            if (so) break;
            so = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        }
        if (!so) return "libnccl.so.2 not found";
        GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(dlsym(so, "ncclGetUniqueId"));
        CommInitRank = reinterpret_cast<decltype(CommInitRank)>(dlsym(so, "ncclCommInitRank"));
        AllReduce = reinterpret_cast<decltype(AllReduce)>(dlsym(so, "ncclAllReduce"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(so, "ncclCommDestroy"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(so, "ncclGetErrorString"));
        CommAbort = reinterpret_cast<decltype(CommAbort)>(dlsym(so, "ncclCommAbort"));
        if (!GetUniqueId || !CommInitRank || !AllReduce || !CommDestroy || !GetErrorString) { so = nullptr; return "libnccl lacks an entry point"; }
        return nullptr;
    }
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    return api;
}

}  // namespace pcb
