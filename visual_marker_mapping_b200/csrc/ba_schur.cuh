// ba_schur.cuh — the Schur-complement builder and the products with the reduced camera system.
//
// The normal equations of the BA problem (src/TagReconstructor.cpp:699-724) have the block form
//     [ B   E ] [dc]   [gc]        B_i = sum Jc^T Jc (cameras, 6x6)   C_j = sum Jm^T Jm (markers, 6x6)
//     [ E^T C ] [dm] = -[gm]       E_k = Jc^T Jm of tag observation k (4 corners = 8 residual rows)
// The marker blocks are eliminated: S = (B+L_c) - E (C+L_m)^-1 E^T acts on camera space only and is
// never formed; one product with S streams the E blocks ONCE from HBM.
//
// Two device orders of the tag observations exist side by side (both built in vmm_ba_set_problem):
//   camera order  — sorted by camera: B_i / g_c / cost by warp-segmented reduction (no hot atomics)
//   marker order  — sorted by marker, CSR offsets mptr[]: one CTA owns one marker, so C_j / g_m are
//                   plain stores and w_j = sum_k E_k^T p needs no atomics at all
// E is stored in marker order as 32-observation tiles of 18 double2 planes ("tile-SoA"): lane l of a
// warp reading plane c touches Et[(tile*18 + c)*32 + l], i.e. every load/store instruction of a warp
// is one contiguous 512 B run.
#pragma once
#include "ba_kernels.cuh"

namespace vmm {

__device__ __forceinline__ size_t e_index(int k, int plane) {
    return ((size_t)(k >> 5) * 18 + plane) * 32 + (k & 31);
}

// ------------------------------------------------------------------------------------------
// Camera half of the linearisation (camera order, one thread per tag observation):
//   B_i += Jc^T Jc (upper triangle, 21), gc_i += Jc^T r, robust cost.
// Algorithmic bytes per tag observation: 64 (uv) + 8 (indices); FP64-bound.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_linearize_cameras(Intrinsics in, int nto, const double2* __restrict__ uv, const int* __restrict__ ocam,
                    const int* __restrict__ otag, const double* __restrict__ camP, const double* __restrict__ mkP,
                    int loss_kind, double loss_a, double* __restrict__ Bt, double* __restrict__ gc,
                    double* __restrict__ cost_partials) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    double Bv[21], gv[6];
#pragma unroll
    for (int i = 0; i < 21; ++i) Bv[i] = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) gv[i] = 0.0;
    int ci = -1;
    if (k < nto) {
        ci = __ldg(ocam + k);
        const int mi = __ldg(otag + k);
        double cam[12], mk[12];
        load12(camP + (size_t)ci * CAM_PREP, cam);
        load12(mkP + (size_t)mi * MARK_PREP, mk);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const double2 o = __ldcs(uv + 4 * (size_t)k + c);
            double r[2], Jc[12];
            cost += corner_camera_side(in, cam, mk, c, o.x, o.y, loss_kind, loss_a, r, Jc);
#pragma unroll
            for (int row = 0; row < 2; ++row) {
                const double* J = Jc + 6 * row;
                int t = 0;
#pragma unroll
                for (int a = 0; a < 6; ++a) {
#pragma unroll
                    for (int b = a; b < 6; ++b) Bv[t++] += J[a] * J[b];
                    gv[a] += J[a] * r[row];
                }
            }
        }
    }
    warp_segmented_add<21>(Bv, ci, Bt, 21);
    warp_segmented_add<6>(gv, ci, gc, 6);
    const double s = block_sum(cost);
    if (threadIdx.x == 0) cost_partials[blockIdx.x] = s;
}

// ------------------------------------------------------------------------------------------
// Marker half of the linearisation (marker order, one CTA per marker, one thread per tag
// observation): E_k = Jc^T Jm written as coalesced tiles, C_j = sum Jm^T Jm (upper triangle, 21) and
// gm_j = sum Jm^T r reduced inside the CTA and stored once — no atomics.
// Algorithmic bytes per tag observation: 64 (uv) + 4 (camera index) + 288 (E) = 356; FP64-bound
// (~2.4 kflop per tag observation).
// ------------------------------------------------------------------------------------------
constexpr int kLinThreads = 128;

__global__ void __launch_bounds__(kLinThreads)
k_linearize_markers(Intrinsics in, const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active,
                    const int* __restrict__ mo_cam, const double2* __restrict__ uvm, size_t uv_stride,
                    const double* __restrict__ camP, const double* __restrict__ mkP, int loss_kind, double loss_a,
                    double2* __restrict__ Et, double* __restrict__ Ct, double* __restrict__ gm) {
    const int j = blockIdx.x;
    if (!tag_active[j]) return;   // constant origin tag / unobserved tag: no columns (Ct, gm stay zero)
    const int kb = mptr[j], ke = mptr[j + 1];
    __shared__ double red[27][kLinThreads + 1];
    double mk[12];
    load12(mkP + (size_t)j * MARK_PREP, mk);
    double acc[27];
#pragma unroll
    for (int i = 0; i < 27; ++i) acc[i] = 0.0;
    for (int k = kb + threadIdx.x; k < ke; k += kLinThreads) {
        const int ci = __ldg(mo_cam + k);
        double cam[12];
        load12(camP + (size_t)ci * CAM_PREP, cam);
        double E[36];
#pragma unroll
        for (int i = 0; i < 36; ++i) E[i] = 0.0;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            const double2 o = __ldcs(uvm + (size_t)c * uv_stride + k);
            double r[2], J[24];
            corner_full(in, cam, mk, c, o.x, o.y, loss_kind, loss_a, r, J);
#pragma unroll
            for (int row = 0; row < 2; ++row) {
                const double* Jc = J + row * 12;
                const double* Jm = J + row * 12 + 6;
#pragma unroll
                for (int a = 0; a < 6; ++a) {
#pragma unroll
                    for (int b = 0; b < 6; ++b) E[a * 6 + b] += Jc[a] * Jm[b];
                }
                int t = 0;
#pragma unroll
                for (int a = 0; a < 6; ++a) {
#pragma unroll
                    for (int b = a; b < 6; ++b) acc[t++] += Jm[a] * Jm[b];
                    acc[21 + a] += Jm[a] * r[row];
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 18; ++i) Et[e_index(k, i)] = make_double2(E[2 * i], E[2 * i + 1]);
    }
    // CTA reduction of the 27 marker-side sums through shared memory
#pragma unroll
    for (int i = 0; i < 27; ++i) red[i][threadIdx.x] = acc[i];
    __syncthreads();
    const int v = threadIdx.x >> 2, part = threadIdx.x & 3;   // 32 value slots x 4 parts of 32 entries
    double s = 0.0;
    if (v < 27) {
#pragma unroll 8
        for (int i = 0; i < 32; ++i) s += red[v][part * 32 + ((i + 4 * part) & 31)];
    }
    s += shfl_xor_d(s, 1);
    s += shfl_xor_d(s, 2);
    if (part == 0 && v < 27) {
        if (v < 21) Ct[(size_t)j * 21 + v] = s;
        else gm[(size_t)j * 6 + (v - 21)] = s;
    }
}

// ------------------------------------------------------------------------------------------
// Products with the E blocks, one CTA per marker (marker order).
//   SP_FULL    w_j = sum_k E_k^T v[cam(k)];  z_j = Cinv_j w_j;  out[cam(k)] += sign * E_k z_j;  dot += w_j.z_j
//              (the whole off-diagonal part of S*p; E is read from HBM once and kept in registers)
//   SP_GATHER  w_j = sum_k E_k^T v[cam(k)]  stored to mvec[j]          (multi-GPU: all-reduced next)
//   SP_SCATTER out[cam(k)] += sign * E_k mvec[j]
// Algorithmic bytes per tag observation: 288 (E) + 4 (camera index); the camera-space vectors
// (48 B per camera) live in L2.  `done` != 0 turns the launch into a no-op (device-side PCG stop).
// ------------------------------------------------------------------------------------------
enum { SP_FULL = 0, SP_GATHER = 1, SP_SCATTER = 2 };
constexpr int kSpThreads = 256;

__device__ __forceinline__ void load_e_tile(const double2* __restrict__ Et, int k, double2 (&E)[18]) {
#pragma unroll
    for (int i = 0; i < 18; ++i) E[i] = __ldcs(Et + e_index(k, i));
}
__device__ __forceinline__ void et_times_x(const double2 (&E)[18], const double (&x)[6], double (&y)[6]) {
#pragma unroll
    for (int a = 0; a < 6; ++a) {
#pragma unroll
        for (int h = 0; h < 3; ++h) {
            y[2 * h] += E[3 * a + h].x * x[a];
            y[2 * h + 1] += E[3 * a + h].y * x[a];
        }
    }
}
__device__ __forceinline__ void scatter_e_z(const double2 (&E)[18], const double (&z)[6], double sign,
                                            double* __restrict__ out) {
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        double s = 0.0;
#pragma unroll
        for (int h = 0; h < 3; ++h) s += E[3 * a + h].x * z[2 * h] + E[3 * a + h].y * z[2 * h + 1];
        atomicAdd(out + a, sign * s);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kSpThreads, 2)
k_schur_product(const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active,
                const int* __restrict__ mo_cam, const double2* __restrict__ Et, const double* __restrict__ Cinv,
                const double* __restrict__ vin_c, double* __restrict__ mvec, double sign, double* __restrict__ out_c,
                double* __restrict__ dot_acc, const double* __restrict__ done, double* zero0, double* zero1,
                double* zero2) {
    if (done && *done != 0.0) return;
    const int j = blockIdx.x;
    const int tid = threadIdx.x;
    if (j == 0 && tid == 0) {   // recycle the other parity's PCG accumulators (nobody touches them in this launch)
        if (zero0) *zero0 = 0.0;
        if (zero1) *zero1 = 0.0;
        if (zero2) *zero2 = 0.0;
    }
    if (!tag_active[j]) {
        if (MODE == SP_GATHER && tid < 6) mvec[(size_t)j * 6 + tid] = 0.0;
        return;
    }
    const int kb = mptr[j], ke = mptr[j + 1];
    __shared__ double s_part[kSpThreads / 32][6];
    __shared__ double s_w[6], s_z[6];
    double2 Er[18];
    const int k0 = kb + tid;
    int cam0 = 0;
    if (MODE != SP_SCATTER) {
        double y[6] = { 0, 0, 0, 0, 0, 0 };
        if (k0 < ke) {
            cam0 = __ldg(mo_cam + k0);
            load_e_tile(Et, k0, Er);
            double x[6];
            const double2* x2 = reinterpret_cast<const double2*>(vin_c + (size_t)cam0 * 6);
#pragma unroll
            for (int i = 0; i < 3; ++i) { const double2 t = __ldg(x2 + i); x[2 * i] = t.x; x[2 * i + 1] = t.y; }
            et_times_x(Er, x, y);
        }
        for (int k = k0 + kSpThreads; k < ke; k += kSpThreads) {   // markers with more than 256 observations
            const int cam = __ldg(mo_cam + k);
            double2 Ek[18];
#pragma unroll
            for (int i = 0; i < 18; ++i) Ek[i] = __ldg(Et + e_index(k, i));
            double x[6];
            const double2* x2 = reinterpret_cast<const double2*>(vin_c + (size_t)cam * 6);
#pragma unroll
            for (int i = 0; i < 3; ++i) { const double2 t = __ldg(x2 + i); x[2 * i] = t.x; x[2 * i + 1] = t.y; }
            et_times_x(Ek, x, y);
        }
#pragma unroll
        for (int b = 0; b < 6; ++b) y[b] = warp_sum(y[b]);
        if ((tid & 31) == 0) {
#pragma unroll
            for (int b = 0; b < 6; ++b) s_part[tid >> 5][b] = y[b];
        }
        __syncthreads();
        if (tid < 6) {
            double w = 0.0;
#pragma unroll
            for (int q = 0; q < kSpThreads / 32; ++q) w += s_part[q][tid];
            if (MODE == SP_GATHER) mvec[(size_t)j * 6 + tid] = w;
            else s_w[tid] = w;
        }
        if (MODE == SP_GATHER) return;
        __syncthreads();
        if (tid < 6) {
            const double* Ci = Cinv + (size_t)j * 36 + tid * 6;
            double z = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) z += Ci[b] * s_w[b];
            s_z[tid] = z;
        }
        __syncthreads();
        if (tid == 0 && dot_acc) {
            double d = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) d += s_w[b] * s_z[b];
            atomicAdd(dot_acc, d);
        }
    } else {
        if (tid < 6) s_z[tid] = mvec[(size_t)j * 6 + tid];
        __syncthreads();
        if (k0 < ke) {
            cam0 = __ldg(mo_cam + k0);
            load_e_tile(Et, k0, Er);
        }
    }
    double z[6];
#pragma unroll
    for (int b = 0; b < 6; ++b) z[b] = s_z[b];
    if (k0 < ke) scatter_e_z(Er, z, sign, out_c + (size_t)cam0 * 6);
    for (int k = k0 + kSpThreads; k < ke; k += kSpThreads) {
        const int cam = __ldg(mo_cam + k);
        double2 Ek[18];
#pragma unroll
        for (int i = 0; i < 18; ++i) Ek[i] = __ldg(Et + e_index(k, i));   // L2 hit: read moments ago
        scatter_e_z(Ek, z, sign, out_c + (size_t)cam * 6);
    }
}

// ------------------------------------------------------------------------------------------
// Block diagonal of the reduced camera system (the SCHUR_JACOBI preconditioner):
//   Sd_i = B_i + diag(lam_c) - sum_k E_k Cinv_{m(k)} E_k^T     (upper triangle; mirrored afterwards)
// ------------------------------------------------------------------------------------------
__global__ void k_schur_diag_init(int n, const double* __restrict__ B, const double* __restrict__ lam,
                                  double* __restrict__ Sd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 36) return;
    const int blk = i / 36, e = i % 36;
    const int a = e / 6, b = e % 6;
    Sd[i] = B[i] + ((a == b) ? lam[(size_t)blk * 6 + a] : 0.0);
}

__global__ void __launch_bounds__(kSpThreads, 2)
k_schur_diag(const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active, const int* __restrict__ mo_cam,
             const double2* __restrict__ Et, const double* __restrict__ Cinv, double* __restrict__ Sd) {
    const int j = blockIdx.x;
    if (!tag_active[j]) return;
    __shared__ double s_c[36];
    if (threadIdx.x < 36) s_c[threadIdx.x] = Cinv[(size_t)j * 36 + threadIdx.x];
    __syncthreads();
    const int kb = mptr[j], ke = mptr[j + 1];
    for (int k = kb + threadIdx.x; k < ke; k += kSpThreads) {
        const int cam = __ldg(mo_cam + k);
        double2 E2[18];
        load_e_tile(Et, k, E2);
        const double* E = reinterpret_cast<const double*>(E2);
        double* o = Sd + (size_t)cam * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double W[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int l = 0; l < 6; ++l) s += E[a * 6 + l] * s_c[l * 6 + c];
                W[c] = s;
            }
#pragma unroll
            for (int b = a; b < 6; ++b) {
                double s = 0.0;
#pragma unroll
                for (int c = 0; c < 6; ++c) s += W[c] * E[b * 6 + c];
                atomicAdd(o + a * 6 + b, -s);
            }
        }
    }
}

__global__ void k_mirror_blocks(int n, double* __restrict__ M) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* f = M + (size_t)i * 36;
    for (int a = 0; a < 6; ++a)
        for (int b = a + 1; b < 6; ++b) f[b * 6 + a] = f[a * 6 + b];
}

}  // namespace vmm
