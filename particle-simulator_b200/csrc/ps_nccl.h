// ps_nccl.h — NCCL entry points resolved at run time.
//
// libparticles_cuda does not link NCCL: a single-GPU user needs none, and a multi-GPU user normally
// already has one NCCL in the process (PyTorch bundles 2.28.9; the system has 2.27.3) — binding at
// link time could put two different NCCLs behind the same symbols.  The slab path therefore looks
// the functions up once: first in whatever libnccl the process has loaded, then by dlopen.
// Only the point-to-point subset is used (ncclSend / ncclRecv inside ncclGroupStart / End).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <mutex>
#include <string>

namespace ps {

struct NcclUniqueId { char internal[128]; };   // NCCL_UNIQUE_ID_BYTES
using NcclComm = void*;
constexpr int kNcclUint8 = 1;                   // ncclUint8
constexpr int kNcclSuccess = 0;

struct NcclApi {
    int (*GetUniqueId)(NcclUniqueId*) = nullptr;
    int (*CommInitRank)(NcclComm*, int, NcclUniqueId, int) = nullptr;
    int (*CommDestroy)(NcclComm) = nullptr;
    int (*Send)(const void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, NcclComm, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string error;
    bool ok = false;
};

inline const NcclApi& nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        void* h = nullptr;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL); if (h) break; }   // already in the process (torch)
        if (!h) for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) { api.error = std::string("cannot load libnccl: ") + (dlerror() ? dlerror() : "not found"); return; }
        auto sym = [&](const char* name) -> void* {
            void* p = dlsym(h, name);
            if (!p && api.error.empty()) api.error = std::string("libnccl lacks ") + name;
            return p;
        };
        api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
        api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
        api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
        api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
        api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
        api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
        api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
        api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
        api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
        api.ok = api.error.empty();
    });
    return api;
}

}  // namespace ps
