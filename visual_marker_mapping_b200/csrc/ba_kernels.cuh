// ba_kernels.cuh — sm_100a kernels of the bundle-adjustment engine.
//
// Data layout in HBM (per GPU; a distributed engine holds one camera shard, markers replicated):
//   uv        double2[4*Nto]   observed corners, (u,v) interleaved, tag observations sorted by camera
//   ocam/otag int32[Nto]       local camera index / marker index of each tag observation
//   camP      double[12*Nc]    prepared camera blocks (R row-major | t)
//   mkP       double[12*Nm]    prepared marker blocks (a | b | t | pad)
//   Et, Es    bytes            E_k = Jc^T Jm (camera rows, marker cols) and the scaled blocks Es_k = E_k L_j^-T, ONE copy each, in
//                              marker order as 8-observation tiles of 18 double2 planes (ba_schur.cuh): what cp.async.bulk moves
//   B, C      double[36*Nc], double[36*Nm]   diagonal blocks of J^T J; gc, gm gradients
//   res/jac   double[2][Nobs], double[24][Nobs]  component-major output of the materialising pass
//
// Reference semantics: TagReconstructionCostFunction.h:101-159 and src/TagReconstructor.cpp:699-724
// (what is evaluated), SURVEY.md Appendix A (Ceres semantics).  Nothing here is translated from the
// reference: it has no device code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "ba_math.h"

namespace vmm {

constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------
// small device helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ double shfl_up_d(double v, int d) { return __shfl_up_sync(0xffffffffu, v, d); }
__device__ __forceinline__ double shfl_xor_d(double v, int d) { return __shfl_xor_sync(0xffffffffu, v, d); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += shfl_xor_d(v, d);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v = fmax(v, shfl_xor_d(v, d));
    return v;
}

// Block-wide sum, result valid in thread 0. blockDim.x must be a multiple of 32, <= 1024.
__device__ __forceinline__ double block_sum(double v) {
    __shared__ double sm[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();   // protect sm reuse across calls
    if (lane == 0) sm[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = (lane < (int)(blockDim.x >> 5)) ? sm[lane] : 0.0;
        v = warp_sum(v);
    }
    return v;
}
__device__ __forceinline__ double block_max(double v) {
    __shared__ double smx[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) smx[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = (lane < (int)(blockDim.x >> 5)) ? smx[lane] : 0.0;
        v = warp_max(v);
    }
    return v;
}

// max for non-negative doubles through the integer ordering of their bit patterns
__device__ __forceinline__ void atomic_max_nonneg(double* addr, double v) {
    atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

// Warp-level segmented reduction: lanes hold N values and a key; keys are non-decreasing across
// the warp (observations sorted by camera, or by marker).  After the inclusive segmented scan the
// last lane of every run holds the run total and adds it to dst[key*stride + i]; runs that cross
// a warp boundary meet in memory, so atomics are issued once per (run, warp) instead of once per
// observation.  key < 0 marks an idle lane.
template <int N>
__device__ __forceinline__ void warp_segmented_add(double (&v)[N], int key, double* __restrict__ dst, int stride) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int kup = __shfl_up_sync(0xffffffffu, key, d);
        const bool take = (lane >= d) && (kup == key);
#pragma unroll
        for (int i = 0; i < N; ++i) {
            const double up = shfl_up_d(v[i], d);
            if (take) v[i] += up;
        }
    }
    const int knext = __shfl_down_sync(0xffffffffu, key, 1);
    const bool tail = (lane == 31) || (knext != key);
    if (tail && key >= 0) {
        double* o = dst + (size_t)key * stride;
#pragma unroll
        for (int i = 0; i < N; ++i) atomicAdd(o + i, v[i]);
    }
}

// ------------------------------------------------------------------------------------------
// pose preparation
// ------------------------------------------------------------------------------------------
__global__ void k_prep_cameras(int n, const double* __restrict__ q, const double* __restrict__ t,
                               double* __restrict__ camP) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double out[CAM_PREP];
    prep_camera(q + 4 * (size_t)i, t + 3 * (size_t)i, out);
#pragma unroll
    for (int k = 0; k < CAM_PREP; ++k) camP[(size_t)i * CAM_PREP + k] = out[k];
}
__global__ void k_prep_markers(int n, const double* __restrict__ q, const double* __restrict__ t,
                               const double* __restrict__ wh, double* __restrict__ mkP) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double out[MARK_PREP];
    prep_marker(q + 4 * (size_t)i, t + 3 * (size_t)i, wh[2 * (size_t)i], wh[2 * (size_t)i + 1], out);
#pragma unroll
    for (int k = 0; k < MARK_PREP; ++k) mkP[(size_t)i * MARK_PREP + k] = out[k];
}

__device__ __forceinline__ void load12(const double* __restrict__ p, double (&o)[12]) {
    const double2* p2 = reinterpret_cast<const double2*>(p);
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const double2 a = __ldg(p2 + k);
        o[2 * k] = a.x; o[2 * k + 1] = a.y;
    }
}

// ------------------------------------------------------------------------------------------
// Residual + analytic Jacobian, materialised (the HBM-roofline kernel; vmm_ba_evaluate).
// One thread per corner.  Reads 16 B (u,v) + shared indices, writes 2 + 24 doubles component-major
// so that every store instruction of a warp is one contiguous 256 B run.
// Algorithmic bytes per corner: 16 (uv) + 2 (indices, 8 B per tag observation) + 208 (r, J) = 226.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_residual_jacobian(Intrinsics in, int64_t nobs, const double2* __restrict__ uv, const int* __restrict__ ocam,
                    const int* __restrict__ otag, const double* __restrict__ camP, const double* __restrict__ mkP,
                    int loss_kind, double loss_a, double* __restrict__ res, double* __restrict__ jac,
                    double* __restrict__ cost_partials) {
    const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    if (c < nobs) {
        const int64_t k = c >> 2;
        const int corner = (int)(c & 3);
        const int ci = __ldg(ocam + k), mi = __ldg(otag + k);
        double cam[12], mk[12];
        load12(camP + (size_t)ci * CAM_PREP, cam);
        load12(mkP + (size_t)mi * MARK_PREP, mk);
        const double2 o = __ldcs(uv + c);
        double r[2], J[24];
        cost = corner_full(in, cam, mk, corner, o.x, o.y, loss_kind, loss_a, r, J);
        __stcs(res + c, r[0]);
        __stcs(res + nobs + c, r[1]);
#pragma unroll
        for (int e = 0; e < 24; ++e) __stcs(jac + (size_t)e * nobs + c, J[e]);
    }
    if (cost_partials) {
        const double s = block_sum(cost);
        if (threadIdx.x == 0) cost_partials[blockIdx.x] = s;
    }
}

// Residual-only cost at a (candidate) point: reads 16 B per corner, writes one partial per block.
__global__ void __launch_bounds__(kThreads)
k_cost(Intrinsics in, int64_t nto, const double2* __restrict__ uv, const int* __restrict__ ocam,
       const int* __restrict__ otag, const double* __restrict__ camP, const double* __restrict__ mkP, int loss_kind,
       double loss_a, double* __restrict__ cost_partials) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double cost = 0.0;
    if (k < nto) {
        const int ci = __ldg(ocam + k), mi = __ldg(otag + k);
        double cam[12], mk[12];
        load12(camP + (size_t)ci * CAM_PREP, cam);
        load12(mkP + (size_t)mi * MARK_PREP, mk);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const double2 o = __ldcs(uv + 4 * k + c);
            cost += corner_cost(in, cam, mk, c, o.x, o.y, loss_kind, loss_a);
        }
    }
    const double s = block_sum(cost);
    if (threadIdx.x == 0) cost_partials[blockIdx.x] = s;
}

// Deterministic second stage: one block sums n partials in a fixed order into out[0].
__global__ void k_sum_partials(int n, const double* __restrict__ partials, double* __restrict__ out) {
    double s = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s += partials[i];
    s = block_sum(s);
    if (threadIdx.x == 0) out[0] = s;
}

__device__ __forceinline__ int tri_index(int a, int b) { return a * 6 - a * (a - 1) / 2 + (b - a); }   // a <= b

// Expand packed upper triangles (21) into full symmetric 6x6 blocks (36) and extract the diagonal.
__global__ void k_expand_blocks(int n, const double* __restrict__ tri, double* __restrict__ full,
                                double* __restrict__ diag) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* t = tri + (size_t)i * 21;
    double* f = full + (size_t)i * 36;
    int idx = 0;
    for (int a = 0; a < 6; ++a)
        for (int b = a; b < 6; ++b) {
            const double v = t[idx++];
            f[a * 6 + b] = v;
            f[b * 6 + a] = v;
            if (a == b) diag[(size_t)i * 6 + a] = v;
        }
}

// Jacobi scaling (iteration 0): scale = 1 / (1 + sqrt(H_ii))   
__global__ void k_jacobi_scale(int n, const double* __restrict__ hdiag, double* __restrict__ scale, int enabled) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    scale[i] = enabled ? 1.0 / (1.0 + sqrt(hdiag[i])) : 1.0;
}
// LM diagonal: d = clamp(scale^2 H_ii, lo, hi)  (LevenbergMarquardtStrategy, on the scaled Jacobian)
__global__ void k_lm_diagonal(int n, const double* __restrict__ hdiag, const double* __restrict__ scale,
                              double lo, double hi, double* __restrict__ d) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double s = scale[i];
    d[i] = fmin(fmax(hdiag[i] * s * s, lo), hi);
}
// Damping in unscaled coordinates: lambda = d / (radius * scale^2)
__global__ void k_lambda(int n, const double* __restrict__ d, const double* __restrict__ scale, double radius,
                         double* __restrict__ lam) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double s = scale[i];
    lam[i] = d[i] / (radius * s * s);
}

// out = (A + diag(lam))^-1 per 6x6 block; inactive blocks -> 0.  fail flag set on a non-PD block.
__global__ void k_invert_blocks(int n, const double* __restrict__ A, const double* __restrict__ lam,
                                const unsigned char* __restrict__ active, double* __restrict__ out,
                                int* __restrict__ fail) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double M[36];
    if (active && !active[i]) {
#pragma unroll
        for (int k = 0; k < 36; ++k) out[(size_t)i * 36 + k] = 0.0;
        return;
    }
#pragma unroll
    for (int k = 0; k < 36; ++k) M[k] = A[(size_t)i * 36 + k];
    if (lam) {
#pragma unroll
        for (int a = 0; a < 6; ++a) M[a * 6 + a] += lam[(size_t)i * 6 + a];
    }
    if (!spd_inverse6(M)) {
        atomicExch(fail, 1);
#pragma unroll
        for (int k = 0; k < 36; ++k) M[k] = 0.0;
    }
#pragma unroll
    for (int k = 0; k < 36; ++k) out[(size_t)i * 36 + k] = M[k];
}

// Marker blocks: Linv_j = G_j^-1 with C_j + diag(lam) = G_j G_j^T, so (C_j + diag(lam))^-1 = Linv^T Linv.
// Inactive blocks (constant origin tag, unobserved tags) -> 0.  fail flag set on a non-PD block.
__global__ void k_factor_markers(int n, const double* __restrict__ C, const double* __restrict__ lam,
                                 const unsigned char* __restrict__ active, double* __restrict__ Linv,
                                 int* __restrict__ fail) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double M[36], L[36];
#pragma unroll
    for (int k = 0; k < 36; ++k) L[k] = 0.0;
    if (active[i]) {
#pragma unroll
        for (int k = 0; k < 36; ++k) M[k] = C[(size_t)i * 36 + k];
#pragma unroll
        for (int a = 0; a < 6; ++a) M[a * 6 + a] += lam[(size_t)i * 6 + a];
        if (!spd_inverse6(M, L)) {
            atomicExch(fail, 1);
#pragma unroll
            for (int k = 0; k < 36; ++k) L[k] = 0.0;
        }
    }
#pragma unroll
    for (int k = 0; k < 36; ++k) Linv[(size_t)i * 36 + k] = L[k];
}

// out = alpha * op(Linv) (v + add) per marker; op = transpose when `transposed`.
//   rhs:                z = Linv gm                 (E Cinv gm = Es z)
//   back-substitution:  dm = -Linv^T (z + Es^T dc)  (= -Cinv (gm + E^T dc))
__global__ void __launch_bounds__(kThreads)
k_marker_apply(int nm, const double* __restrict__ Linv, int transposed, const double* __restrict__ v,
               const double* __restrict__ add, double alpha, double* __restrict__ out) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nm) return;
    double x[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) x[a] = alpha * (v[(size_t)j * 6 + a] + (add ? add[(size_t)j * 6 + a] : 0.0));
    const double* L = Linv + (size_t)j * 36;
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        double s = 0.0;
#pragma unroll
        for (int b = 0; b < 6; ++b) s += (transposed ? L[b * 6 + a] : L[a * 6 + b]) * x[b];
        out[(size_t)j * 6 + a] = s;
    }
}

// acc += sum v_i^2 (the w.w term of the sharded PCG, after the all-reduce of w)
__global__ void __launch_bounds__(kThreads)
k_sum_squares(int n, const double* __restrict__ v, double* __restrict__ acc, const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const double x = (i < n) ? v[i] : 0.0;
    const double s = block_sum(x * x);
    if (threadIdx.x == 0) atomicAdd(acc, s);
}

// sharded engines, start of a PCG iteration: clears the boundary vector and the three accumulators of the
// next iteration (one GPU does the latter inside the product kernel)
__global__ void k_zero_boundary(int n, double* __restrict__ wb, double* __restrict__ a, double* __restrict__ b,
                                double* __restrict__ c, const double* __restrict__ done) {
    if (done && *done != 0.0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) wb[i] = 0.0;
    else if (i == n) *a = 0.0;
    else if (i == n + 1) *b = 0.0;
    else if (i == n + 2) *c = 0.0;
}

// PCG scalars, device resident. Slot parity s = iteration & 1.
//   sc[0+s] rz = r.M^-1 r   sc[2+s] wz = w.w (w = Es^T p)   sc[4] rz0      pbp[s] = p.(B+lam)p
//   sc[5] done flag (set on device when rz <= tol^2 rz0; every later PCG launch is then a no-op)
//   sc[6] iterations performed when the flag was set      sc[7] last rz / rz0
// (pbp lives at the tail of the marker-space vector w so that one all-reduce carries both).
enum { SC_RZ = 0, SC_WZ = 2, SC_RZ0 = 4, SC_DONE = 5, SC_ITERS = 6, SC_REL2 = 7, SC_N = 8 };

// r = b (already in r): x = 0, p = 0, z = Minv r, rz[0] = rz0 = r.z.  The first direction (p = z [+ coarse
// part], q = (B + lam) p, pBp[0]) is set by k_pcg_direction(s = 1) right after.
__global__ void __launch_bounds__(kThreads)
k_pcg_init(int nc, const double* __restrict__ Minv, const double* __restrict__ r, double* __restrict__ x,
           double* __restrict__ z, double* __restrict__ p, double* __restrict__ sc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double rz = 0.0;
    if (i < nc) {
        double rv[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) rv[a] = r[(size_t)i * 6 + a];
        const double* Mi = Minv + (size_t)i * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double s = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) s += Mi[a * 6 + b] * rv[b];
            rz += s * rv[a];
            x[(size_t)i * 6 + a] = 0.0;
            p[(size_t)i * 6 + a] = 0.0;
            z[(size_t)i * 6 + a] = s;
        }
    }
    const double s1 = block_sum(rz);
    if (threadIdx.x == 0) { atomicAdd(sc + 0, s1); atomicAdd(sc + 4, s1); }
}

// The two per-iteration vector kernels run SIX threads per camera, thread = 6 * camera + block row: every vector
// access is one coalesced double per thread, a block row of Minv / B is 48 contiguous bytes per thread (three
// 16-byte loads, contiguous across the warp), and the camera's 6-vector is exchanged through shared memory.  One
// thread per camera (100 k threads at C3, a third of the SM thread slots, 288-byte strides) ran at 20 us per launch.
constexpr int kPcgThreads = 192;   // 32 cameras x 6 rows

__device__ __forceinline__ double row_dot6(const double* __restrict__ row, const double* v) {
    const double2* r2 = reinterpret_cast<const double2*>(row);
    const double2 a = r2[0], b = r2[1], c = r2[2];
    double t = a.x * v[0];
    t += a.y * v[1];
    t += b.x * v[2];
    t += b.y * v[3];
    t += c.x * v[4];
    t += c.y * v[5];
    return t;
}

// alpha = rz[s] / (pbp[s] - wz[s]); x += alpha p; r -= alpha q; z = Minv r; rz[o] += r.z
__global__ void __launch_bounds__(kPcgThreads)
k_pcg_update(int nc, int s, const double* __restrict__ Minv, const double* __restrict__ p,
             const double* __restrict__ q, double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
             double* __restrict__ sc, const double* __restrict__ pbp, const double* __restrict__ wz,
             double* __restrict__ rz_out) {
    if (sc[SC_DONE] != 0.0) return;
    __shared__ double s_v[kPcgThreads];
    const int t = blockIdx.x * kPcgThreads + threadIdx.x;
    const bool live = t < 6 * nc;
    const double denom = pbp[s] - wz[s];
    const double alpha = (sc[0 + s] != 0.0) ? sc[0 + s] / denom : 0.0;
    double rv = 0.0;
    if (live) {
        x[t] += alpha * p[t];
        rv = r[t] - alpha * q[t];
        r[t] = rv;
    }
    s_v[threadIdx.x] = rv;
    __syncthreads();
    double rz = 0.0;
    if (live) {
        const double zt = row_dot6(Minv + (size_t)t * 6, s_v + (threadIdx.x / 6) * 6);
        z[t] = zt;
        rz = zt * rv;
    }
    const double s1 = block_sum(rz);
    if (threadIdx.x == 0) atomicAdd(rz_out, s1);
}

// beta = rz[o] / rz[s]; p = z + beta p; q = (B + lam) p; pBp[o] += p.q
// With the two-level preconditioner z holds only the block-Jacobi part; the coarse part P yc is added
// here on the fly (cl_of_cam != nullptr; cameras outside every cluster have cl_of_cam < 0).
struct CoarseView {
    const int* cl_of_cam;
    const double* xbar;
    const double* camP;
    const double* yc;
};
__device__ __forceinline__ void coarse_add(const CoarseView& cv, int i, double (&v)[6]);   // ba_coarse.cuh
__device__ __forceinline__ double coarse_row(const CoarseView& cv, int i, int a);           // ba_coarse.cuh: row a of P_i yc

__global__ void __launch_bounds__(kPcgThreads)
k_pcg_direction(int nc, int s, const double* __restrict__ B, const double* __restrict__ lam,
                const double* __restrict__ z, double* __restrict__ p, double* __restrict__ q,
                double* __restrict__ sc, double* __restrict__ pbp, int iters_done, int min_iters, double tol2,
                CoarseView cv) {
    if (sc[SC_DONE] != 0.0) return;
    __shared__ double s_v[kPcgThreads];
    const int t = blockIdx.x * kPcgThreads + threadIdx.x;
    const bool live = t < 6 * nc;
    const int o = s ^ 1;
    // device-side stopping test: every block evaluates the same predicate on the same (complete) sums
    const double rz_new = sc[0 + o], rz0 = sc[SC_RZ0];
    const bool stop = (iters_done >= min_iters) && (rz_new <= tol2 * rz0);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        sc[SC_REL2] = (rz0 > 0.0) ? rz_new / rz0 : 0.0;
        if (stop) sc[SC_ITERS] = (double)iters_done;
    }
    if (stop) {   // blocks that start after the flag is up leave at the top: same effect
        if (blockIdx.x == 0 && threadIdx.x == 0) sc[SC_DONE] = 1.0;
        return;
    }
    const double beta = (sc[0 + s] != 0.0) ? rz_new / sc[0 + s] : 0.0;
    double pv = 0.0;
    if (live) {
        pv = z[t];
        if (cv.cl_of_cam) pv += coarse_row(cv, t / 6, threadIdx.x % 6);   // kPcgThreads is a multiple of 6
        pv += beta * p[t];
        p[t] = pv;
    }
    s_v[threadIdx.x] = pv;
    __syncthreads();
    double pq = 0.0;
    if (live) {
        const double qt = lam[t] * pv + row_dot6(B + (size_t)t * 6, s_v + (threadIdx.x / 6) * 6);
        q[t] = qt;
        pq = qt * pv;
    }
    const double s1 = block_sum(pq);
    if (threadIdx.x == 0) atomicAdd(pbp + o, s1);
}

// One thread per camera: the variant the six-thread kernels replaced, kept for A/B runs (VMM_BA_PCG_NARROW=1).
// alpha = rz[s] / (pbp[s] - wz[s]); x += alpha p; r -= alpha q; z = Minv r; rz[o] += r.z
__global__ void __launch_bounds__(kThreads)
k_pcg_update_1(int nc, int s, const double* __restrict__ Minv, const double* __restrict__ p,
             const double* __restrict__ q, double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
             double* __restrict__ sc, const double* __restrict__ pbp, const double* __restrict__ wz,
             double* __restrict__ rz_out) {
    if (sc[SC_DONE] != 0.0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int o = s ^ 1;
    const double denom = pbp[s] - wz[s];
    const double alpha = (sc[0 + s] != 0.0) ? sc[0 + s] / denom : 0.0;
    double rz = 0.0;
    if (i < nc) {
        double rv[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const size_t e = (size_t)i * 6 + a;
            x[e] += alpha * p[e];
            rv[a] = r[e] - alpha * q[e];
            r[e] = rv[a];
        }
        const double* Mi = Minv + (size_t)i * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double t = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) t += Mi[a * 6 + b] * rv[b];
            z[(size_t)i * 6 + a] = t;
            rz += t * rv[a];
        }
    }
    const double s1 = block_sum(rz);
    if (threadIdx.x == 0) atomicAdd(rz_out, s1);
}

// beta = rz[o] / rz[s]; p = z + beta p; q = (B + lam) p; pBp[o] += p.q
// With the two-level preconditioner z holds only the block-Jacobi part; the coarse part P yc is added
// here on the fly (cl_of_cam != nullptr; cameras outside every cluster have cl_of_cam < 0).
__global__ void __launch_bounds__(kThreads)
k_pcg_direction_1(int nc, int s, const double* __restrict__ B, const double* __restrict__ lam,
                const double* __restrict__ z, double* __restrict__ p, double* __restrict__ q,
                double* __restrict__ sc, double* __restrict__ pbp, int iters_done, int min_iters, double tol2,
                CoarseView cv) {
    if (sc[SC_DONE] != 0.0) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int o = s ^ 1;
    // device-side stopping test: every block evaluates the same predicate on the same (complete) sums
    const double rz_new = sc[0 + o], rz0 = sc[SC_RZ0];
    const bool stop = (iters_done >= min_iters) && (rz_new <= tol2 * rz0);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        sc[SC_REL2] = (rz0 > 0.0) ? rz_new / rz0 : 0.0;
        if (stop) sc[SC_ITERS] = (double)iters_done;
    }
    if (stop) {   // blocks that start after the flag is up leave at the top: same effect
        if (blockIdx.x == 0 && threadIdx.x == 0) sc[SC_DONE] = 1.0;
        return;
    }
    const double beta = (sc[0 + s] != 0.0) ? rz_new / sc[0 + s] : 0.0;
    double pq = 0.0;
    if (i < nc) {
        double pv[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) pv[a] = z[(size_t)i * 6 + a];
        if (cv.cl_of_cam) coarse_add(cv, i, pv);
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            const size_t e = (size_t)i * 6 + a;
            pv[a] += beta * p[e];
            p[e] = pv[a];
        }
        const double* Bi = B + (size_t)i * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double t = lam[(size_t)i * 6 + a] * pv[a];
#pragma unroll
            for (int b = 0; b < 6; ++b) t += Bi[a * 6 + b] * pv[b];
            q[(size_t)i * 6 + a] = t;
            pq += t * pv[a];
        }
    }
    const double s1 = block_sum(pq);
    if (threadIdx.x == 0) atomicAdd(pbp + o, s1);
}

// y = (B + diag(lam)) x per 6x6 block (test entry vmm_ba_schur_multiply)
__global__ void k_block_multiply(int n, const double* __restrict__ B, const double* __restrict__ lam,
                                 const double* __restrict__ x, double* __restrict__ y) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double xv[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) xv[a] = x[(size_t)i * 6 + a];
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        double t = lam[(size_t)i * 6 + a] * xv[a];
#pragma unroll
        for (int b = 0; b < 6; ++b) t += B[(size_t)i * 36 + a * 6 + b] * xv[b];
        y[(size_t)i * 6 + a] = t;
    }
}

// r = -gc (right-hand side start); the E Cinv gm term is added by k_schur_product<SP_SCATTER>
__global__ void k_negate(int n, const double* __restrict__ a, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = -a[i];
}

// ------------------------------------------------------------------------------------------
// LM step bookkeeping.  stats (camera part or marker part):
//   [0] step_norm^2 (ambient)  [1] delta.g  [2] delta^T Lambda delta  [3] delta_c . r_pcg  [4] x_norm^2 of candidate
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
k_candidate(int n, const unsigned char* __restrict__ active, const double* __restrict__ q, const double* __restrict__ t,
            const double* __restrict__ delta, const double* __restrict__ g, const double* __restrict__ lam,
            const double* __restrict__ rpcg, double* __restrict__ qc, double* __restrict__ tc,
            double* __restrict__ stats) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double sn = 0.0, dg = 0.0, dld = 0.0, dr = 0.0, xn = 0.0;
    if (i < n) {
        double qi[4], ti[3], qo[4], to[3];
#pragma unroll
        for (int k = 0; k < 4; ++k) qi[k] = q[(size_t)i * 4 + k];
#pragma unroll
        for (int k = 0; k < 3; ++k) ti[k] = t[(size_t)i * 3 + k];
        if (active[i]) {
            double d[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) d[k] = delta[(size_t)i * 6 + k];
#pragma unroll
            for (int k = 0; k < 3; ++k) to[k] = ti[k] + d[k];
            quat_plus(qi, d + 3, qo);
#pragma unroll
            for (int k = 0; k < 3; ++k) { const double e = to[k] - ti[k]; sn += e * e; xn += to[k] * to[k]; }
#pragma unroll
            for (int k = 0; k < 4; ++k) { const double e = qo[k] - qi[k]; sn += e * e; xn += qo[k] * qo[k]; }
#pragma unroll
            for (int k = 0; k < 6; ++k) {
                const size_t e = (size_t)i * 6 + k;
                dg += d[k] * g[e];
                dld += d[k] * d[k] * lam[e];
                if (rpcg) dr += d[k] * rpcg[e];
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) qo[k] = qi[k];
#pragma unroll
            for (int k = 0; k < 3; ++k) to[k] = ti[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) qc[(size_t)i * 4 + k] = qo[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) tc[(size_t)i * 3 + k] = to[k];
    }
    const double a0 = block_sum(sn), a1 = block_sum(dg), a2 = block_sum(dld), a3 = block_sum(dr), a4 = block_sum(xn);
    if (threadIdx.x == 0) {
        atomicAdd(stats + 0, a0); atomicAdd(stats + 1, a1); atomicAdd(stats + 2, a2);
        atomicAdd(stats + 3, a3); atomicAdd(stats + 4, a4);
    }
}

// gradient_max_norm = |x - Plus(x, -g)|_inf and x_norm^2 over the active blocks  
//   out[0] = max (atomic max), out[1] += x_norm^2
__global__ void __launch_bounds__(kThreads)
k_gradient_norm(int n, const unsigned char* __restrict__ active, const double* __restrict__ q,
                const double* __restrict__ t, const double* __restrict__ g, double* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double m = 0.0, xn = 0.0;
    if (i < n && active[i]) {
        double qi[4], qo[4], d[3];
#pragma unroll
        for (int k = 0; k < 4; ++k) { qi[k] = q[(size_t)i * 4 + k]; xn += qi[k] * qi[k]; }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double tv = t[(size_t)i * 3 + k];
            xn += tv * tv;
            m = fmax(m, fabs(g[(size_t)i * 6 + k]));
            d[k] = -g[(size_t)i * 6 + 3 + k];
        }
        quat_plus(qi, d, qo);
#pragma unroll
        for (int k = 0; k < 4; ++k) m = fmax(m, fabs(qi[k] - qo[k]));
    }
    const double bm = block_max(m);
    const double bx = block_sum(xn);
    if (threadIdx.x == 0) { atomic_max_nonneg(out + 0, bm); atomicAdd(out + 1, bx); }
}


// ------------------------------------------------------------------------------------------
// Reprojection statistics (src/TagReconstructor.cpp:340-428): per tag observation, the sum of the
// four Euclidean corner errors under CameraModel::projectPoint (src/CameraModel.cpp:6-26).
// Reproduces the reference's quirks: Eigen's toRotationMatrix does not normalise q, and the
// distorted x is reused inside the y formula.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ void eigen_rot(const double* q, double R[9]) {
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
                 tyz = tz * y, tzz = tz * z;
    R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}
__global__ void __launch_bounds__(kThreads)
k_reprojection_sums(Intrinsics in, int nto, const double2* __restrict__ uv, const int* __restrict__ ocam,
                    const int* __restrict__ otag, const double* __restrict__ cam_q, const double* __restrict__ cam_t,
                    const double* __restrict__ tag_q, const double* __restrict__ tag_t,
                    const double* __restrict__ tag_wh, double* __restrict__ sums) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nto) return;
    const int ci = ocam[k], mi = otag[k];
    double Rm[9], Rc[9];
    eigen_rot(tag_q + 4 * (size_t)mi, Rm);
    eigen_rot(cam_q + 4 * (size_t)ci, Rc);
    const double w = tag_wh[2 * (size_t)mi], h = tag_wh[2 * (size_t)mi + 1];
    double sum = 0.0;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const double px = corner_sx(c) * w / 2.0, py = corner_sy(c) * h / 2.0;
        double pw[3], pc[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) pw[i] = Rm[3 * i] * px + Rm[3 * i + 1] * py + tag_t[3 * (size_t)mi + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) pc[i] = Rc[3 * i] * pw[0] + Rc[3 * i + 1] * pw[1] + Rc[3 * i + 2] * pw[2] + cam_t[3 * (size_t)ci + i];
        double x = pc[0] / pc[2], y = pc[1] / pc[2];
        const double r2 = x * x + y * y;
        x = x * (1 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3))) + 2 * in.p1 * x * y + in.p2 * (r2 + 2 * x * x);
        y = y * (1 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3))) + 2 * in.p2 * x * y + in.p1 * (r2 + 2 * y * y);
        const double2 o = uv[4 * (size_t)k + c];
        const double du = in.fx * x + in.cx - o.x, dv = in.fy * y + in.cy - o.y;
        sum += sqrt(du * du + dv * dv);
    }
    sums[k] = sum;
}

// ------------------------------------------------------------------------------------------
// Measurement kernels
// ------------------------------------------------------------------------------------------
// DFMA throughput: 8 independent chains per thread.
__global__ void __launch_bounds__(256)
k_dfma_peak(int iters, double seed, double* __restrict__ out) {
    double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    const double s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 123.456) out[0] = s;   // never true: keeps the chains alive
}
// streaming copy, 16 B per thread per step
__global__ void __launch_bounds__(256)
k_copy16(size_t n16, const double2* __restrict__ src, double2* __restrict__ dst) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n16; i += stride) __stcs(dst + i, __ldcs(src + i));
}

}  // namespace vmm
