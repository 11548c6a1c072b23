// ba_no_alloc.cuh — ONE entry point that allocates nothing: vmm_ba_cost_no_alloc (include/vmm_ba.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the robust cost of the problem
// — what vmm_ba_evaluate returns in *cost, the two-stage k_cost / k_sum_partials reduction of the engine — on DEVICE
// buffers the caller owns, with the scratch the engine normally allocates (prepared pose blocks, per-block cost
// partials) in a file-scope __device__ array of fixed size.  Every GPU that loads the library holds that array
// (kNoAllocScratchBytes, 0.6 MB) until the process exits.
//
// This header is included by ba_engine.cu (the library is one translation unit), so the array, the
// cudaGetSymbolAddress call and the launches that use its address are in the same .cu file.  In host code the
// array's NAME is a host-side stand-in, not a device address: the address handed to the kernels comes from
// cudaGetSymbolAddress, taken once per device with that device current.
//
// All calls share the array and it keeps the last call's contents: every call rewrites every word it reads (prepared
// blocks of all cameras and markers, one partial per block) before reading it, so nothing has to start at zero; two
// calls that overlap on the GPU (different streams, no dependency) race on it.
#pragma once
#include <mutex>

namespace vmm {

constexpr int64_t kNoAllocMaxCameras = 4096;
constexpr int64_t kNoAllocMaxMarkers = 1024;
constexpr int64_t kNoAllocMaxTagObs = 131072;
constexpr int kNoAllocThreads = 128;
constexpr int64_t kNoAllocMaxBlocks = kNoAllocMaxTagObs / kNoAllocThreads;   // 1024 cost partials
constexpr size_t kNoAllocCamOff = 0;                                                       // doubles
constexpr size_t kNoAllocMarkOff = kNoAllocCamOff + (size_t)kNoAllocMaxCameras * CAM_PREP;
constexpr size_t kNoAllocPartOff = kNoAllocMarkOff + (size_t)kNoAllocMaxMarkers * MARK_PREP;
constexpr size_t kNoAllocScratchDoubles = kNoAllocPartOff + (size_t)kNoAllocMaxBlocks;
constexpr size_t kNoAllocScratchBytes = kNoAllocScratchDoubles * sizeof(double);           // 499,712 B

static __device__ __align__(256) double g_no_alloc_scratch[kNoAllocScratchDoubles];

}  // namespace vmm

namespace {

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_no_alloc_mutex;
double* g_no_alloc_ptr[kNoAllocMaxDevices] = { nullptr };   // device address of g_no_alloc_scratch per device

// device address of the scratch array on `device` (made current by the caller); nullptr + error text on failure
double* no_alloc_scratch(int device) {
    std::lock_guard<std::mutex> lock(g_no_alloc_mutex);
    if (!g_no_alloc_ptr[device]) {
        double* p = nullptr;
        const cudaError_t err = cudaGetSymbolAddress((void**)&p, vmm::g_no_alloc_scratch);
        if (err != cudaSuccess || !p) {
            fail_with(nullptr, VMM_BA_ERR_CUDA, "vmm_ba_cost_no_alloc: cudaGetSymbolAddress(scratch) on device %d: %s", device, cudaGetErrorString(err));
            return nullptr;
        }
        g_no_alloc_ptr[device] = p;
    }
    return g_no_alloc_ptr[device];
}

}  // namespace

extern "C" int vmm_ba_cost_no_alloc(int device, void* cuda_stream, double fx, double fy, double cx, double cy, const double dist[5],
                                    int64_t num_cameras, const double* d_cam_q, const double* d_cam_t, int64_t num_markers,
                                    const double* d_tag_q, const double* d_tag_t, const double* d_tag_wh, int64_t num_tag_obs,
                                    const int32_t* d_to_cam, const int32_t* d_to_tag, const double* d_to_uv, int loss_kind,
                                    double loss_scale, double* d_cost) {
    using namespace vmm;
    if (device < 0 || device >= kNoAllocMaxDevices) return fail_with(nullptr, VMM_BA_ERR_INVALID_ARGUMENT, "vmm_ba_cost_no_alloc: device %d out of range", device);
    if (!dist || !d_cost || num_cameras < 0 || num_markers < 0 || num_tag_obs < 0)
        return fail_with(nullptr, VMM_BA_ERR_INVALID_ARGUMENT, "vmm_ba_cost_no_alloc: null pointer or negative size");
    if (num_cameras > kNoAllocMaxCameras || num_markers > kNoAllocMaxMarkers || num_tag_obs > kNoAllocMaxTagObs)
        return fail_with(nullptr, VMM_BA_ERR_INVALID_ARGUMENT,
                         "vmm_ba_cost_no_alloc: %lld cameras / %lld markers / %lld tag observations exceed the static scratch (at most %lld / %lld / %lld); "
                         "use vmm_ba_evaluate for larger problems",
                         (long long)num_cameras, (long long)num_markers, (long long)num_tag_obs, (long long)kNoAllocMaxCameras,
                         (long long)kNoAllocMaxMarkers, (long long)kNoAllocMaxTagObs);
    if (loss_kind < 0 || loss_kind > 2 || !(loss_scale > 0)) return fail_with(nullptr, VMM_BA_ERR_INVALID_ARGUMENT, "vmm_ba_cost_no_alloc: bad loss");
    if ((num_cameras > 0 && (!d_cam_q || !d_cam_t)) || (num_markers > 0 && (!d_tag_q || !d_tag_t || !d_tag_wh)) ||
        (num_tag_obs > 0 && (!d_to_cam || !d_to_tag || !d_to_uv)))
        return fail_with(nullptr, VMM_BA_ERR_INVALID_ARGUMENT, "vmm_ba_cost_no_alloc: null input array");
    cudaError_t err = cudaSetDevice(device);
    if (err != cudaSuccess) return fail_with(nullptr, VMM_BA_ERR_CUDA, "vmm_ba_cost_no_alloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(err));
    double* scratch = no_alloc_scratch(device);
    if (!scratch) return VMM_BA_ERR_CUDA;
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    double* camP = scratch + kNoAllocCamOff;
    double* mkP = scratch + kNoAllocMarkOff;
    double* partials = scratch + kNoAllocPartOff;
    const Intrinsics in{ fx, fy, cx, cy, dist[0], dist[1], dist[2], dist[3], dist[4] };
    if (num_cameras > 0) k_prep_cameras<<<nblocks(num_cameras, kThreads), kThreads, 0, stream>>>((int)num_cameras, d_cam_q, d_cam_t, camP);
    if (num_markers > 0) k_prep_markers<<<nblocks(num_markers, kThreads), kThreads, 0, stream>>>((int)num_markers, d_tag_q, d_tag_t, d_tag_wh, mkP);
    const int blocks = num_tag_obs > 0 ? nblocks(num_tag_obs, kNoAllocThreads) : 0;
    if (blocks > 0)
        k_cost<<<blocks, kNoAllocThreads, 0, stream>>>(in, num_tag_obs, reinterpret_cast<const double2*>(d_to_uv), d_to_cam, d_to_tag, camP, mkP, loss_kind,
                                                        loss_scale, partials);
    k_sum_partials<<<1, 256, 0, stream>>>(blocks, partials, d_cost);   // blocks == 0: writes 0
    err = cudaGetLastError();
    if (err != cudaSuccess) return fail_with(nullptr, VMM_BA_ERR_CUDA, "vmm_ba_cost_no_alloc: launch failed: %s", cudaGetErrorString(err));
    return VMM_BA_OK;
}
