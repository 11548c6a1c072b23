// ba_coarse.cuh — the coarse level of the two-level preconditioner of the reduced camera system.
//
// Block-Jacobi PCG on S = (B + L_c) - Es Es^T needs O(site diameter / camera footprint) iterations:
// the smooth modes of a large marker map (whole regions drifting rigidly against each other) are
// invisible to a per-camera preconditioner.  Cameras are grouped into spatial clusters
// (vmm_ba_set_problem, k-d splits of the camera centres); cluster c contributes the six rigid motions
// of its part of the world to a coarse space,
//     P_i y_c = [ -R_i (y_tau + xbar_c x y_theta) ;  -1/2 R_i y_theta ]        (camera i in cluster c)
// in the solver's tangent coordinates (translation, then the half-angle rotation increment of
// ceres::QuaternionParameterization, src/TagReconstructor.cpp:661): the pose change of a camera that
// rides along when the world around it is moved by (y_tau, y_theta) about the cluster centroid.  The
// preconditioner is additive,  M^-1 r = blockdiag(S)^-1 r + P (P^T S P)^-1 P^T r,  with the dense
// coarse matrix A_c = P^T S P (6 n_clusters squared, a few thousand) assembled and inverted once per
// LM iteration and applied as one dense matrix-vector product per PCG iteration.
#pragma once
#include "ba_kernels.cuh"
#include "ba_schur.cuh"

namespace vmm {

// P_i y for one camera: R row-major (camP), xbar the cluster centroid
__device__ __forceinline__ void coarse_prolong(const double* __restrict__ R, const double* __restrict__ xb,
                                               const double (&y)[6], double (&out)[6]) {
    const double ux = y[0] + (xb[1] * y[5] - xb[2] * y[4]);
    const double uy = y[1] + (xb[2] * y[3] - xb[0] * y[5]);
    const double uz = y[2] + (xb[0] * y[4] - xb[1] * y[3]);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        out[a] = -(R[3 * a] * ux + R[3 * a + 1] * uy + R[3 * a + 2] * uz);
        out[3 + a] = -0.5 * (R[3 * a] * y[3] + R[3 * a + 1] * y[4] + R[3 * a + 2] * y[5]);
    }
}
// the 6x6 matrix P_i (row-major): columns = coarse dofs
__device__ __forceinline__ void coarse_basis(const double* __restrict__ R, const double* __restrict__ xb, double (&P)[36]) {
#pragma unroll
    for (int c = 0; c < 6; ++c) {
        double y[6] = { 0, 0, 0, 0, 0, 0 }, o[6];
        y[c] = 1.0;
        coarse_prolong(R, xb, y, o);
#pragma unroll
        for (int a = 0; a < 6; ++a) P[a * 6 + c] = o[a];
    }
}

// P_i^T r, formed from the same basis as the prolongation (exact transpose by construction)
__device__ __forceinline__ void coarse_restrict(const double* __restrict__ R, const double* __restrict__ xb,
                                                const double (&r)[6], double (&out)[6]) {
    double P[36];
    coarse_basis(R, xb, P);
#pragma unroll
    for (int c = 0; c < 6; ++c) {
        double s = 0.0;
#pragma unroll
        for (int a = 0; a < 6; ++a) s += P[a * 6 + c] * r[a];
        out[c] = s;
    }
}

// v += P_i yc[cluster(i)]  (declared in ba_kernels.cuh for k_pcg_direction)
__device__ __forceinline__ void coarse_add(const CoarseView& cv, int i, double (&v)[6]) {
    const int c = cv.cl_of_cam[i];
    if (c < 0) return;
    double y[6], o[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) y[a] = cv.yc[(size_t)c * 6 + a];
    coarse_prolong(cv.camP + (size_t)i * CAM_PREP, cv.xbar + 3 * (size_t)c, y, o);
#pragma unroll
    for (int a = 0; a < 6; ++a) v[a] += o[a];
}

// row a of P_i yc[cluster(i)] (k_pcg_direction, one thread per block row): the same arithmetic as coarse_prolong
__device__ __forceinline__ double coarse_row(const CoarseView& cv, int i, int a) {
    const int c = cv.cl_of_cam[i];
    if (c < 0) return 0.0;
    const double* __restrict__ y = cv.yc + (size_t)c * 6;
    const double* __restrict__ xb = cv.xbar + 3 * (size_t)c;
    const int k = a < 3 ? a : a - 3;
    const double* __restrict__ R = cv.camP + (size_t)i * CAM_PREP + 3 * k;
    if (a >= 3) return -0.5 * (R[0] * y[3] + R[1] * y[4] + R[2] * y[5]);
    const double ux = y[0] + (xb[1] * y[5] - xb[2] * y[4]);
    const double uy = y[1] + (xb[2] * y[3] - xb[0] * y[5]);
    const double uz = y[2] + (xb[0] * y[4] - xb[1] * y[3]);
    return -(R[0] * ux + R[1] * uy + R[2] * uz);
}

constexpr int kCoarseThreads = 128;

// A_c[c,c] += sum_{i in c} P_i^T (B_i + diag(lam_i)) P_i.  One CTA per cluster.
__global__ void __launch_bounds__(kCoarseThreads)
k_coarse_diag(const int* __restrict__ cl_ptr, const int* __restrict__ cl_cams, const double* __restrict__ xbar,
              const double* __restrict__ camP, const double* __restrict__ B, const double* __restrict__ lam, int nC,
              double* __restrict__ Ac) {
    const int c = blockIdx.x;
    __shared__ double s_acc[36];
    if (threadIdx.x < 36) s_acc[threadIdx.x] = 0.0;
    __syncthreads();
    double acc[36];
#pragma unroll
    for (int k = 0; k < 36; ++k) acc[k] = 0.0;
    const double xb[3] = { xbar[3 * c], xbar[3 * c + 1], xbar[3 * c + 2] };
    for (int q = cl_ptr[c] + threadIdx.x; q < cl_ptr[c + 1]; q += kCoarseThreads) {
        const int i = cl_cams[q];
        double P[36], T[36];
        coarse_basis(camP + (size_t)i * CAM_PREP, xb, P);
        const double* Bi = B + (size_t)i * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int col = 0; col < 6; ++col) {
                double s = lam[(size_t)i * 6 + a] * P[a * 6 + col];
#pragma unroll
                for (int b = 0; b < 6; ++b) s += Bi[a * 6 + b] * P[b * 6 + col];
                T[a * 6 + col] = s;
            }
#pragma unroll
        for (int r = 0; r < 6; ++r)
#pragma unroll
            for (int col = 0; col < 6; ++col) {
                double s = 0.0;
#pragma unroll
                for (int a = 0; a < 6; ++a) s += P[a * 6 + r] * T[a * 6 + col];
                acc[r * 6 + col] += s;
            }
    }
#pragma unroll
    for (int k = 0; k < 36; ++k) {
        const double s = warp_sum(acc[k]);
        if ((threadIdx.x & 31) == 0) atomicAdd(&s_acc[k], s);
    }
    __syncthreads();
    if (threadIdx.x < 36) {
        const int r = threadIdx.x / 6, col = threadIdx.x % 6;
        atomicAdd(Ac + (size_t)(6 * c + r) * nC + 6 * c + col, s_acc[threadIdx.x]);
    }
}

// A_c -= W^T W with W = Es^T P, in two kernels so that a sharded engine can sum the per-rank parts of W
// (ncclAllReduce) in between:
//   k_coarse_w      W_{j,c} = sum_{k in j, cam(k) in c} Es_k^T P_cam(k)   over this rank's observations,
//                   written to Wbuf[(mcl_ptr[j] + position of c in marker j's cluster list) * 36 ..]
//   k_coarse_outer  every pair (a, b) of clusters touched by marker j receives -W_a^T W_b
// The clusters a marker touches are few (mcl_ptr / mcl list them — over ALL ranks' observations; mo_lc[k]
// is the position of observation k's cluster in that list).  One CTA per marker.
constexpr int kCoarseMaxTouched = 24;
constexpr int kCwPad = 37;   // row pitch of the staged G blocks (odd: conflict-free column writes)
__global__ void __launch_bounds__(kCoarseThreads)
k_coarse_w(const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active, const int* __restrict__ mtile,
           const int* __restrict__ mo_cam, const unsigned char* __restrict__ mo_lc, const int* __restrict__ mcl_ptr,
           const int* __restrict__ cl_of_cam, const double* __restrict__ xbar, const double* __restrict__ camP,
           const unsigned char* __restrict__ Es, double* __restrict__ Wbuf) {
    const int j = blockIdx.x;
    const int tb = mcl_ptr[j], nt = mcl_ptr[j + 1] - tb;
    if (nt == 0) return;
    __shared__ double s_g[kCoarseThreads][kCwPad];
    __shared__ double s_w[kCoarseMaxTouched * 36];
    __shared__ unsigned char s_lc[kCoarseThreads];
    for (int i = threadIdx.x; i < nt * 36; i += kCoarseThreads) s_w[i] = 0.0;
    const int kb = mptr[j], n = tag_active[j] ? mptr[j + 1] - kb : 0, t0 = mtile[j];
    for (int base = 0; base < n; base += kCoarseThreads) {
        const int i = base + threadIdx.x;
        __syncthreads();
        if (i < n) {
            const int cam = __ldg(mo_cam + kb + i);
            const int c = __ldg(cl_of_cam + cam);
            s_lc[threadIdx.x] = mo_lc[kb + i];
            const double xb[3] = { xbar[3 * c], xbar[3 * c + 1], xbar[3 * c + 2] };
            double P[36], E[36];
            coarse_basis(camP + (size_t)cam * CAM_PREP, xb, P);
#pragma unroll
            for (int p = 0; p < 18; ++p) {
                const double2 v = __ldcs(e_entry(Es, t0, i, p));
                E[2 * p] = v.x; E[2 * p + 1] = v.y;
            }
            // G = Es^T P  (marker dof x coarse dof)
#pragma unroll
            for (int m = 0; m < 6; ++m)
#pragma unroll
                for (int col = 0; col < 6; ++col) {
                    double g = 0.0;
#pragma unroll
                    for (int a = 0; a < 6; ++a) g += E[a * 6 + m] * P[a * 6 + col];
                    s_g[threadIdx.x][m * 6 + col] = g;
                }
        } else {
            s_lc[threadIdx.x] = 255;
        }
        __syncthreads();
        // thread (cluster slot t, entry) owns s_w[t*36 + entry]: no atomics
        const int cnt = min(kCoarseThreads, n - base);
        for (int id = threadIdx.x; id < nt * 36; id += kCoarseThreads) {
            const int t = id / 36, en = id % 36;
            double acc = 0.0;
            for (int q = 0; q < cnt; ++q)
                if (s_lc[q] == t) acc += s_g[q][en];
            s_w[id] += acc;
        }
    }
    __syncthreads();
    for (int id = threadIdx.x; id < nt * 36; id += kCoarseThreads) Wbuf[(size_t)tb * 36 + id] = s_w[id];
}

// Set-up (one GPU): the distinct clusters every marker touches, in order of first appearance in the marker order,
// and per observation the position of its cluster in that list.  One warp per marker.
__global__ void __launch_bounds__(128)
k_marker_cluster_lists(int nm, const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active,
                       const int* __restrict__ mo_cam, const int* __restrict__ cl_of_cam, int* __restrict__ lists,
                       int* __restrict__ counts, unsigned char* __restrict__ mo_lc, int* __restrict__ overflow) {
    const int j = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= nm) return;
    __shared__ int s_list[4][kCoarseMaxTouched];
    int* list = s_list[threadIdx.x >> 5];
    int cnt = 0;
    if (tag_active[j]) {
        const int kb = mptr[j], ke = mptr[j + 1];
        for (int base = kb; base < ke; base += 32) {
            const int k = base + lane;
            const bool valid = k < ke;
            const int cc = valid ? __ldg(cl_of_cam + __ldg(mo_cam + k)) : -1;
            int pos = -1;
            for (int t = 0; t < cnt; ++t) if (list[t] == cc) pos = t;
            unsigned pending = __ballot_sync(0xffffffffu, valid && pos < 0);
            while (pending) {
                const int leader = __ffs(pending) - 1;
                const int v = __shfl_sync(0xffffffffu, cc, leader);
                if (cnt < kCoarseMaxTouched) { if (lane == 0) list[cnt] = v; }
                else if (lane == 0) *overflow = 1;
                if (valid && pos < 0 && cc == v) pos = min(cnt, kCoarseMaxTouched - 1);
                cnt = min(cnt + 1, kCoarseMaxTouched);
                pending &= ~__ballot_sync(0xffffffffu, cc == v);
                __syncwarp();
            }
            if (valid) mo_lc[k] = (unsigned char)pos;
        }
    }
    __syncwarp();
    if (lane == 0) counts[j] = cnt;
    for (int t = lane; t < cnt; t += 32) lists[(size_t)j * kCoarseMaxTouched + t] = list[t];
}

// Sharded set-up: positions in this rank's local cluster list -> positions in the merged (all ranks) list.
// map[j * kCoarseMaxTouched + local position] = merged position.  One warp per marker.
__global__ void __launch_bounds__(128)
k_remap_cluster_positions(int nm, const int* __restrict__ mptr, const unsigned char* __restrict__ map, unsigned char* __restrict__ mo_lc) {
    const int j = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= nm) return;
    for (int k = mptr[j] + lane; k < mptr[j + 1]; k += 32) mo_lc[k] = map[(size_t)j * kCoarseMaxTouched + mo_lc[k]];
}

// markers j with j % stride == first are handled by this rank (stride = world size)
__global__ void __launch_bounds__(kCoarseThreads)
k_coarse_outer(int nm, int first, int stride, const unsigned char* __restrict__ tag_active, const int* __restrict__ mcl_ptr,
               const int* __restrict__ mcl, const double* __restrict__ Wbuf, int nC, double* __restrict__ Ac) {
    const int j = first + blockIdx.x * stride;
    if (j >= nm || !tag_active[j]) return;
    const int tb = mcl_ptr[j], nt = mcl_ptr[j + 1] - tb;
    __shared__ double s_w[kCoarseMaxTouched * 36];
    for (int i = threadIdx.x; i < nt * 36; i += kCoarseThreads) s_w[i] = Wbuf[(size_t)tb * 36 + i];
    __syncthreads();
    for (int t = threadIdx.x; t < nt * nt * 36; t += kCoarseThreads) {
        const int pair = t / 36, en = t % 36;
        const int ia = pair / nt, ib = pair % nt;
        const int r = en / 6, col = en % 6;
        double s = 0.0;
#pragma unroll
        for (int m = 0; m < 6; ++m) s += s_w[ia * 36 + m * 6 + r] * s_w[ib * 36 + m * 6 + col];
        atomicAdd(Ac + (size_t)(6 * mcl[tb + ia] + r) * nC + 6 * mcl[tb + ib] + col, -s);
    }
}

__global__ void k_identity(int n, double* __restrict__ A) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)n * n) return;
    A[idx] = (idx / n == idx % n) ? 1.0 : 0.0;
}

// identity on the padding diagonal [n, ld) of a square matrix with leading dimension ld
__global__ void k_pad_identity(int n, int ld, double* __restrict__ A) {
    for (int i = n + threadIdx.x; i < ld; i += blockDim.x) A[(size_t)i * ld + i] = 1.0;
}

// m x m identity into a block of a column-major matrix with leading dimension ld
__global__ void k_identity_block(int m, int ld, double* __restrict__ A) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)m * m) return;
    const int r = (int)(idx % m), c = (int)(idx / m);
    A[(size_t)c * ld + r] = (r == c) ? 1.0 : 0.0;
}

// syrk leaves the inverse in one triangle: mirror it (column-major lower == row-major upper).
__global__ void k_symmetrize(int n, double* __restrict__ A) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (int64_t)n * n) return;
    const int r = (int)(idx / n), c = (int)(idx % n);
    if (c > r) A[(size_t)c * n + r] = A[(size_t)r * n + c];   // row-major view: upper triangle is valid
}

// rc = P^T r restricted per cluster (one CTA per cluster, no atomics).
__global__ void __launch_bounds__(kCoarseThreads)
k_coarse_restrict(const int* __restrict__ cl_ptr, const int* __restrict__ cl_cams, const double* __restrict__ xbar,
                  const double* __restrict__ camP, const double* __restrict__ r, double* __restrict__ rc,
                  const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    const int c = blockIdx.x;
    const double xb[3] = { xbar[3 * c], xbar[3 * c + 1], xbar[3 * c + 2] };
    double acc[6] = { 0, 0, 0, 0, 0, 0 };
    for (int q = cl_ptr[c] + threadIdx.x; q < cl_ptr[c + 1]; q += kCoarseThreads) {
        const int i = cl_cams[q];
        double rv[6], o[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) rv[a] = r[(size_t)i * 6 + a];
        coarse_restrict(camP + (size_t)i * CAM_PREP, xb, rv, o);
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[a] += o[a];
    }
    __shared__ double s_p[kCoarseThreads / 32][6];
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        const double s = warp_sum(acc[a]);
        if ((threadIdx.x & 31) == 0) s_p[threadIdx.x >> 5][a] = s;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kCoarseThreads / 32; ++w) s += s_p[w][threadIdx.x];
        rc[(size_t)c * 6 + threadIdx.x] = s;
    }
}

// yc = Ainv rc (dense symmetric, row-major), one warp per row; dot_acc += rc . yc (the coarse part of r.z)
__global__ void __launch_bounds__(256)
k_coarse_apply(int n, int ld, const double* __restrict__ Ainv, const double* __restrict__ rc, double* __restrict__ yc,
               double* __restrict__ dot_acc, double* __restrict__ dot_acc2, const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    double s = 0.0;
    if (row < n) {
        const double2* a2 = reinterpret_cast<const double2*>(Ainv + (size_t)row * ld);
        const double2* x2 = reinterpret_cast<const double2*>(rc);
        for (int k = lane; k < n / 2; k += 32) {
            const double2 a = __ldg(a2 + k), x = __ldg(x2 + k);
            s += a.x * x.x + a.y * x.y;
        }
        s = warp_sum(s);
        if (lane == 0) yc[row] = s;
    }
    __shared__ double s_d[8];
    if (lane == 0) s_d[threadIdx.x >> 5] = (row < n) ? s * rc[row] : 0.0;
    __syncthreads();
    if (threadIdx.x == 0 && dot_acc) {
        double d = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) d += s_d[w];
        atomicAdd(dot_acc, d);
        if (dot_acc2) atomicAdd(dot_acc2, d);
    }
}

// Sharded engines: the r.z partials travel in the tail of the rc all-reduce (pack before, unpack after),
// and the coarse part of r.z is summed in a fixed order so that every rank takes the same stopping decision.
__global__ void k_pack2(const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ tail,
                        const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    if (threadIdx.x == 0) tail[0] = *a;
    if (threadIdx.x == 1) tail[1] = b ? *b : 0.0;
}
__global__ void __launch_bounds__(256)
k_coarse_dot(int n, const double* __restrict__ rc, const double* __restrict__ yc, const double* __restrict__ tail,
             double* __restrict__ acc, double* __restrict__ acc2, const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    double s = 0.0;
    for (int k = threadIdx.x; k < n; k += 256) s += rc[k] * yc[k];
    s = warp_sum(s);
    __shared__ double s_p[8];
    if ((threadIdx.x & 31) == 0) s_p[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double d = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) d += s_p[w];
        // tail[0] (and tail[1]): the block-Jacobi part of r.z, complete after the all-reduce
        *acc = tail[0] + d;
        if (acc2) *acc2 = tail[1] + d;
    }
}

}  // namespace vmm
