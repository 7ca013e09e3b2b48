// NCCL entry points of libkvarq_b200.so, bound at run time.
//
// The library is loaded into processes that may already hold an NCCL (PyTorch ships its own
// libnccl.so.2) and into processes on single-GPU boxes that need none.  So the handful of NCCL
// calls the engine makes are resolved on first use: the copy already mapped into the process if
// there is one, else $KVQ_NCCL_LIB, else the system's libnccl.so.2.  Types come from <nccl.h>.
#pragma once
#include <dlfcn.h>
#include <nccl.h>
#include <stdlib.h>
#include <string>

namespace kvq_nccl {

struct Api {
    void *handle = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    std::string error;
};

inline Api &api()
{
    static Api a;
    static bool tried = false;
    if (tried) return a;
    tried = true;
    const char *names[4] = {nullptr, nullptr, nullptr, nullptr};
    int n = 0;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // already in the process (e.g. through torch)
    if (!h) {
        if (const char *env = getenv("KVQ_NCCL_LIB")) names[n++] = env;
        names[n++] = "libnccl.so.2";
        names[n++] = "libnccl.so";
        for (int i = 0; i < n && !h; i++) h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) { a.error = std::string("cannot load NCCL: ") + (dlerror() ? dlerror() : "libnccl.so.2 not found"); return a; }
    a.handle = h;
#define KVQ_NCCL_SYM(f) a.f = reinterpret_cast<decltype(a.f)>(dlsym(h, "nccl" #f)); if (!a.f) a.error = "NCCL symbol nccl" #f " missing";
    KVQ_NCCL_SYM(GetUniqueId) KVQ_NCCL_SYM(CommInitRank) KVQ_NCCL_SYM(CommDestroy) KVQ_NCCL_SYM(AllReduce)
    KVQ_NCCL_SYM(GroupStart) KVQ_NCCL_SYM(GroupEnd) KVQ_NCCL_SYM(GetErrorString) KVQ_NCCL_SYM(GetVersion)
#undef KVQ_NCCL_SYM
    return a;
}

} // namespace kvq_nccl
