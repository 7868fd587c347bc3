// NCCL bound at first use instead of at link time.
//
// libonilda.so is loaded into processes that also hold PyTorch (bench.py, the tests), and PyTorch brings its own
// libnccl.so.2 (2.28) while the system one is 2.27: a DT_NEEDED entry would pull the system library in as soon as
// libonilda.so is loaded, and a later `import torch` then fails on symbols only the newer one has.  dlopen of the
// soname at the first communicator call picks up whichever libnccl.so.2 the process already holds (PyTorch's, when
// it is there) and the system one otherwise (the `lda` program).  Single-GPU runs never load it.
#pragma once

#include <dlfcn.h>
#include <stdlib.h>
#include <nccl.h>

namespace onilda {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*);
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    const char* (*GetErrorString)(ncclResult_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    ncclResult_t (*ReduceScatter)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    bool ok;
};

// nullptr (with onilda_fail's message set) when no libnccl.so.2 can be found
inline const NcclApi* nccl_api()
{
    static NcclApi api = [] {
        NcclApi a = {};
        const char* named = getenv("ONILDA_NCCL_LIB");      // an explicit library, if the caller wants one
        void* h = named ? dlopen(named, RTLD_NOW | RTLD_LOCAL) : nullptr;
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
        if (!h) return a;
        bool all = true;
        auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) all = false; return p; };
        a.GetUniqueId = (decltype(a.GetUniqueId))sym("ncclGetUniqueId");
        a.CommInitRank = (decltype(a.CommInitRank))sym("ncclCommInitRank");
        a.CommInitAll = (decltype(a.CommInitAll))sym("ncclCommInitAll");
        a.CommDestroy = (decltype(a.CommDestroy))sym("ncclCommDestroy");
        a.GetErrorString = (decltype(a.GetErrorString))sym("ncclGetErrorString");
        a.GroupStart = (decltype(a.GroupStart))sym("ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))sym("ncclGroupEnd");
        a.ReduceScatter = (decltype(a.ReduceScatter))sym("ncclReduceScatter");
        a.AllReduce = (decltype(a.AllReduce))sym("ncclAllReduce");
        a.AllGather = (decltype(a.AllGather))sym("ncclAllGather");
        a.ok = all;
        return a;
    }();
    if (!api.ok) {
        onilda_fail(ONILDA_ECUDA, "NCCL (libnccl.so.2) cannot be loaded: %s", dlerror() ? dlerror() : "symbols missing");
        return nullptr;
    }
    return &api;
}

}  // namespace onilda
