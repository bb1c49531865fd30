// nccl_dl.h -- NCCL resolved at run time (dlopen), so that libmcsim_cuda.so has no link-time dependency on it: a
// single-device handle never loads it, and a process that already carries an NCCL (PyTorch bundles its own) shares that
// copy instead of loading a second one.  Types and enums come from <nccl.h>; only the seven entry points used by
// mcsim_group.cuh are looked up.
#ifndef MCSIM_NCCL_DL_H
#define MCSIM_NCCL_DL_H

#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>

#include <mutex>
#include <string>

struct NcclApi {
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    void* handle = nullptr;
    std::string error;
    bool ok() const { return handle != nullptr; }
};

// Loads once.  Order: a copy already in the process, MCSIM_NCCL_LIB (a path), libnccl.so.2 on the loader path.
static inline const NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!h) { const char* env = getenv("MCSIM_NCCL_LIB"); if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL); }
        if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { const char* e = dlerror(); api.error = std::string("cannot load libnccl.so.2: ") + (e ? e : "?"); return; }
        bool all = true;
        auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) { all = false; api.error = std::string("libnccl.so.2 lacks ") + name; } return p; };
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
        if (all) api.handle = h;
    });
    return api;
}

#endif
