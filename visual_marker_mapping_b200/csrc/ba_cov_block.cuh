// ba_cov_block.cuh — marker-translation covariance at site scale (src/TagReconstructor.cpp:744-783 on the final, full
// problem): the diagonal blocks of R^-1, R = C - E^T B^-1 E (cameras eliminated, order 6 x markers), WITHOUT forming R
// densely.
//
// Two markers are coupled in R only when one camera sees both.  Breadth-first levels of that co-observation graph
// (host, once per call) therefore make R block tridiagonal: level k couples to levels k - 1 and k + 1 only.  With diagonal
// blocks D_k and sub-diagonal blocks L_k = R[level k+1, level k] the diagonal blocks of the inverse follow from the
// classical two-sweep recursion, exactly (no truncation, no probing):
//     forward   S_0 = D_0,            S_k = D_k - L_{k-1} S_{k-1}^-1 L_{k-1}^T        T_{k-1} = L_{k-1} S_{k-1}^-1 kept
//     backward  G_last = S_last^-1,   G_k = S_k^-1 + T_k^T G_{k+1} T_k
// and cov_j is the translation 3x3 block of G_{level(j)}.  Every step is a dense fp64 operation on matrices of a few
// thousand rows: the in-place sweep inversion of ba_coarse_inv.cuh and the tiled product below — hand-written, no
// library.  C3 (20k markers, ~30 levels of <= 1300 markers): ~3e13 multiply-adds instead of the 1.7e15 of a dense
// inversion of the 120k x 120k matrix (which would also need 115 GB).
#pragma once
#include "ba_coarse_inv.cuh"
#include "ba_cov.cuh"

namespace vmm {

struct CovPlace {
    int block;   // level block of the marker (-1: no columns)
    int off;     // first row of the marker inside its block
};

// Like k_cov_eliminate_cameras, into the block-tridiagonal storage: diag[k] (order ld[k], row-major) and sub[k] (rows of
// block k + 1, columns of block k, leading dimension ld[k]).  A pair of markers more than one level apart raises *bad.
__global__ void __launch_bounds__(128)
k_covb_eliminate_cameras(const int* __restrict__ cam_ptr, const int* __restrict__ otag, const int* __restrict__ cpos,
                         const int* __restrict__ mptr, const int* __restrict__ mtile, const CovPlace* __restrict__ place,
                         const double* __restrict__ Binv, const unsigned char* __restrict__ Et, double* const* __restrict__ diag,
                         double* const* __restrict__ sub, const int* __restrict__ ld, int* __restrict__ bad) {
    const int i = blockIdx.x;
    const int s0 = cam_ptr[i], cnt = cam_ptr[i + 1] - s0;
    __shared__ double s_binv[36];
    if (threadIdx.x < 36) s_binv[threadIdx.x] = Binv[(size_t)i * 36 + threadIdx.x];
    __syncthreads();
    for (int pr = threadIdx.x; pr < cnt * cnt; pr += blockDim.x) {
        const int a = pr / cnt, b = pr % cnt;
        const int ja = otag[s0 + a], jb = otag[s0 + b];
        const CovPlace pa = place[ja], pb = place[jb];
        if (pa.block < 0 || pb.block < 0 || pa.block < pb.block) continue;   // the upper blocks are the transposes of the lower ones
        if (pa.block > pb.block + 1) { *bad = 1; continue; }
        double Ea[36], Eb[36];
        {
            const int qa = cpos[s0 + a], qb = cpos[s0 + b];
#pragma unroll
            for (int p = 0; p < 18; ++p) {
                const double2 va = *e_entry(Et, mtile[ja], qa - mptr[ja], p);
                const double2 vb = *e_entry(Et, mtile[jb], qb - mptr[jb], p);
                Ea[2 * p] = va.x; Ea[2 * p + 1] = va.y;
                Eb[2 * p] = vb.x; Eb[2 * p + 1] = vb.y;
            }
        }
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
        double* dst = (pa.block == pb.block ? diag[pa.block] : sub[pb.block]) + (size_t)pa.off * ld[pb.block] + pb.off;
        const int l = ld[pb.block];
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) s += Ea[k * 6 + r] * T[k * 6 + c];
                atomicAdd(dst + (size_t)r * l + c, -s);
            }
    }
}

// D_jj += C_j for the markers with columns
__global__ void k_covb_add_marker_blocks(int nm, const double* __restrict__ C, const CovPlace* __restrict__ place, double* const* __restrict__ diag,
                                         const int* __restrict__ ld) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= nm * 36) return;
    const int j = idx / 36, r = (idx % 36) / 6, c = idx % 6;
    const CovPlace p = place[j];
    if (p.block < 0) return;
    diag[p.block][(size_t)(p.off + r) * ld[p.block] + p.off + c] += C[(size_t)j * 36 + r * 6 + c];
}

// cov_j = translation block of G_{block(j)}; called once per block while its G is alive
__global__ void k_covb_collect(int nm, int block, const CovPlace* __restrict__ place, const double* __restrict__ G, int ld, double* __restrict__ cov) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nm || place[j].block != block) return;
    const int o = place[j].off;
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) cov[(size_t)j * 9 + r * 3 + c] = G[(size_t)(o + r) * ld + o + c];
}

constexpr size_t kGemmSmemBytes = sizeof(double) * 2 * kInvNB * kInvNB;

// C (m x n, row-major, ldc) = beta * C + alpha * op(X) * op(Y), all dimensions multiples of 64.
//   X: m x k as stored (tx = 0, ldx) or stored k x m (tx = 1);  Y: k x n as stored (ty = 0) or stored n x k (ty = 1).
// One CTA per 64 x 64 tile of C, 256 threads, 4 x 4 register tiles out of k-major shared-memory panels.
__global__ void __launch_bounds__(kInvThreads)
k_dgemm64(int m, int n, int k, double alpha, const double* __restrict__ X, int ldx, int tx, const double* __restrict__ Y, int ldy, int ty,
          double beta, double* __restrict__ Cm, int ldc) {
    extern __shared__ __align__(16) unsigned char gemm_smem[];   // 2 x 64 x 64 doubles (kGemmSmemBytes)
    double* At = reinterpret_cast<double*>(gemm_smem);
    double* Bt = At + kInvNB * kInvNB;
    const int tiles_n = n / kInvNB;
    const int ti = threadIdx.x / 16, tj = threadIdx.x % 16;
    for (int tile = blockIdx.x; tile < (m / kInvNB) * tiles_n; tile += gridDim.x) {
        const int I = tile / tiles_n, J = tile % tiles_n;
        double acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
        for (int K = 0; K < k; K += kInvNB) {
            __syncthreads();
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
                const int r = e / kInvNB, c = e % kInvNB;   // coalesced over the stored rows
                // At[kk][i] = op(X)[I*64 + i][K + kk]
                if (tx) At[r * kInvNB + c] = X[(size_t)(K + r) * ldx + I * kInvNB + c];
                else At[c * kInvNB + r] = X[(size_t)(I * kInvNB + r) * ldx + K + c];
                // Bt[kk][j] = op(Y)[K + kk][J*64 + j]
                if (ty) Bt[c * kInvNB + r] = Y[(size_t)(J * kInvNB + r) * ldy + K + c];
                else Bt[r * kInvNB + c] = Y[(size_t)(K + r) * ldy + J * kInvNB + c];
            }
            __syncthreads();
#pragma unroll 8
            for (int kk = 0; kk < kInvNB; ++kk) {
                const double2 a01 = *reinterpret_cast<const double2*>(At + kk * kInvNB + ti * 4);
                const double2 a23 = *reinterpret_cast<const double2*>(At + kk * kInvNB + ti * 4 + 2);
                const double2 b01 = *reinterpret_cast<const double2*>(Bt + kk * kInvNB + tj * 4);
                const double2 b23 = *reinterpret_cast<const double2*>(Bt + kk * kInvNB + tj * 4 + 2);
                const double av[4] = { a01.x, a01.y, a23.x, a23.y }, bv[4] = { b01.x, b01.y, b23.x, b23.y };
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
            }
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                double* dst = Cm + (size_t)(I * kInvNB + ti * 4 + a) * ldc + J * kInvNB + tj * 4 + b;
                *dst = (beta == 0.0 ? 0.0 : beta * *dst) + alpha * acc[a][b];
            }
    }
}

}  // namespace vmm
