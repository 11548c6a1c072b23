// ba_stage.cuh — device-side staging of a problem (vmm_ba_set_problem).
//
// The caller's tag observations (any order) are uploaded once as they are; both device orders — camera
// order for the linearisation / cost passes and marker order for the E-block streams — are produced on
// the GPU: two stable radix sorts of 32-bit keys (cub::DeviceRadixSort, sort plumbing of the set-up, not
// part of the solve) and two gather kernels.  Replaces the std::map walk of the reference's assembly loop
// (src/TagReconstructor.cpp:699-724), which is rebuilt from scratch before every solve.
#pragma once
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>
#include <stdint.h>

namespace vmm {

__global__ void k_iota(int64_t n, int* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (int)i;
}

// camera order: sorted local position s -> caller index k = perm[s]; one thread per corner (4 per tag observation)
__global__ void k_stage_camera_order(int64_t ntol, const int* __restrict__ perm, const int* __restrict__ raw_cam,
                                     const int* __restrict__ raw_tag, const double2* __restrict__ raw_uv, int cam0,
                                     int* __restrict__ ocam, int* __restrict__ otag, double2* __restrict__ uv) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 4 * ntol) return;
    const int64_t s = idx >> 2;
    const int c = (int)(idx & 3);
    const int64_t k = perm[s];
    uv[idx] = raw_uv[4 * k + c];
    if (c == 0) {
        ocam[s] = raw_cam[k] - cam0;
        otag[s] = raw_tag[k];
    }
}

// marker order: position pos -> camera-order position s = mperm[pos]; corners go to four planes of stride `stride`
__global__ void k_stage_marker_order(int64_t ntol, const int* __restrict__ mperm, const int* __restrict__ ocam,
                                     const double2* __restrict__ uv, size_t stride, int* __restrict__ mo_cam,
                                     double2* __restrict__ uvm, int* __restrict__ cpos) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 4 * ntol) return;
    const int64_t pos = idx % ntol;
    const int c = (int)(idx / ntol);
    const int64_t s = mperm[pos];
    uvm[(size_t)c * stride + pos] = uv[4 * s + c];
    if (c == 0) {
        mo_cam[pos] = ocam[s];
        cpos[s] = (int)pos;   // camera-order position -> marker-order position (the E block of that observation)
    }
}

inline int key_bits(int64_t n) {
    int b = 1;
    while (b < 31 && ((int64_t)1 << b) < n) ++b;
    return b;
}

// stable sort of (keys, values) pairs; temp storage is (re)allocated by the caller through `alloc_temp`
template <typename Alloc>
cudaError_t stable_sort_pairs(const int* keys_in, int* keys_out, const int* vals_in, int* vals_out, int64_t n, int bits,
                              cudaStream_t stream, Alloc&& alloc_temp) {
    size_t bytes = 0;
    cudaError_t err = cub::DeviceRadixSort::SortPairs(nullptr, bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, bits, stream);
    if (err != cudaSuccess) return err;
    void* tmp = alloc_temp(bytes);
    if (!tmp && bytes) return cudaErrorMemoryAllocation;
    return cub::DeviceRadixSort::SortPairs(tmp, bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, bits, stream);
}

}  // namespace vmm
