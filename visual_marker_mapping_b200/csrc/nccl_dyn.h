// nccl_dyn.h — NCCL bound at run time (dlopen) so that libvmm_ba.so has no link-time dependency on
// a particular libnccl: inside a torchrun rank the already-loaded torch-bundled libnccl.so.2 is
// reused (same soname), in a plain C++ host the system one is opened.  Only the handful of entry
// points the engine needs.
#pragma once
#include <dlfcn.h>
#include <stddef.h>
#include <cuda_runtime.h>

namespace vmm {

typedef struct ncclComm* nccl_comm_t;
typedef struct { char internal[128]; } nccl_unique_id;
// values of ncclDataType_t / ncclRedOp_t (stable across NCCL 2.x)
enum { NCCL_INT32 = 2, NCCL_FLOAT64 = 8 };
enum { NCCL_SUM = 0, NCCL_MAX = 2 };

struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(nccl_unique_id*) = nullptr;
    int (*CommInitRank)(nccl_comm_t*, int, nccl_unique_id, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, nccl_comm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    bool ok = false;
    const char* error = "";

    bool load() {
        if (ok) return true;
        const char* names[] = { "libnccl.so.2", "libnccl.so" };
        for (const char* n : names) {
            handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
            if (handle) break;
        }
        if (!handle) { error = "libnccl.so.2 not found (dlopen)"; return false; }
        GetUniqueId = (decltype(GetUniqueId))dlsym(handle, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(handle, "ncclCommInitRank");
        CommDestroy = (decltype(CommDestroy))dlsym(handle, "ncclCommDestroy");
        AllReduce = (decltype(AllReduce))dlsym(handle, "ncclAllReduce");
        AllGather = (decltype(AllGather))dlsym(handle, "ncclAllGather");
        GetErrorString = (decltype(GetErrorString))dlsym(handle, "ncclGetErrorString");
        GetVersion = (decltype(GetVersion))dlsym(handle, "ncclGetVersion");
        if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllReduce || !AllGather || !GetErrorString) {
            error = "libnccl is missing required symbols";
            return false;
        }
        ok = true;
        return true;
    }
};

inline NcclApi& nccl_api() {
    static NcclApi api;
    return api;
}

}  // namespace vmm
