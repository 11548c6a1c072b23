// ba_cov.cuh — marker-translation covariance (vmm_ba_marker_translation_covariance).
//
// Replaces the ceres::Covariance block of TagReconstructor::doBundleAdjustment (src/TagReconstructor.cpp:744-783):
// the 3x3 blocks of (J^T J)^-1 that belong to every marker's translation, at the current poses.  With the cameras
// eliminated,  (H^-1)_mm = (C - E^T B^-1 E)^-1 =: R^-1,  R dense of order 6 x markers (the reference prints this
// once, for maps of tens to a few hundred markers).  R is assembled camera by camera from the E blocks of the
// linearisation and inverted in place by the blocked sweep operator (k_coarse_sweep_invert, ba_coarse_inv.cuh — hand-written,
// no library call); the wanted blocks are the translation 3x3 blocks on its diagonal.
#pragma once
#include "ba_kernels.cuh"
#include "ba_schur.cuh"

namespace vmm {

// R -= E_a^T B_i^-1 E_b for every ordered pair (a, b) of the tag observations of camera i.  One CTA per camera,
// threads stride over the pairs.  cam_ptr: CSR of the camera order; cpos[s] = marker-order position of
// camera-order observation s; R column-major == row-major (symmetric), leading dimension n.
__global__ void __launch_bounds__(128)
k_cov_eliminate_cameras(const int* __restrict__ cam_ptr, const int* __restrict__ otag, const int* __restrict__ cpos,
                        const int* __restrict__ mptr, const int* __restrict__ mtile, const unsigned char* __restrict__ tag_active,
                        const double* __restrict__ Binv, const unsigned char* __restrict__ Et, int n, double* __restrict__ R) {
    const int i = blockIdx.x;
    const int s0 = cam_ptr[i], cnt = cam_ptr[i + 1] - s0;
    __shared__ double s_binv[36];
    if (threadIdx.x < 36) s_binv[threadIdx.x] = Binv[(size_t)i * 36 + threadIdx.x];
    __syncthreads();
    for (int pr = threadIdx.x; pr < cnt * cnt; pr += blockDim.x) {
        const int a = pr / cnt, b = pr % cnt;
        const int ja = otag[s0 + a], jb = otag[s0 + b];
        if (!tag_active[ja] || !tag_active[jb]) continue;
        double Ea[36], Eb[36];
        {
            const int pa = cpos[s0 + a], pb = cpos[s0 + b];
#pragma unroll
            for (int p = 0; p < 18; ++p) {
                const double2 va = *e_entry(Et, mtile[ja], pa - mptr[ja], p);
                const double2 vb = *e_entry(Et, mtile[jb], pb - mptr[jb], p);
                Ea[2 * p] = va.x; Ea[2 * p + 1] = va.y;
                Eb[2 * p] = vb.x; Eb[2 * p + 1] = vb.y;
            }
        }
        // T = B^-1 E_b (camera dof x marker dof), then E_a^T T
        double T[36];
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) s += s_binv[r * 6 + k] * Eb[k * 6 + c];
                T[r * 6 + c] = s;
            }
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) s += Ea[k * 6 + r] * T[k * 6 + c];
                atomicAdd(R + (size_t)(6 * ja + r) * n + 6 * jb + c, -s);
            }
    }
}

// R_jj += C_j; markers without columns (the constant origin tag, unobserved tags) get an identity block.
__global__ void k_cov_add_marker_blocks(int nm, const double* __restrict__ C, const unsigned char* __restrict__ tag_active, int n,
                                        double* __restrict__ R) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nm * 36) return;
    const int j = idx / 36, r = (idx % 36) / 6, c = idx % 6;
    double* dst = R + (size_t)(6 * j + r) * n + 6 * j + c;
    if (tag_active[j]) *dst += C[(size_t)j * 36 + r * 6 + c];
    else *dst = (r == c) ? 1.0 : 0.0;
}

// cov_j = the translation 3x3 block of the inverted matrix (row-major, leading dimension n); zeros for markers without columns.
__global__ void k_cov_diagonal_blocks(int nm, int n, const double* __restrict__ Rinv, const unsigned char* __restrict__ tag_active,
                                      double* __restrict__ cov) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nm) return;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) cov[(size_t)j * 9 + r * 3 + c] = tag_active[j] ? Rinv[(size_t)(6 * j + r) * n + 6 * j + c] : 0.0;
}

}  // namespace vmm
