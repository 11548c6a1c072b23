// ba_pcg.cuh — the whole preconditioned conjugate-gradient solve of one LM step as ONE persistent kernel.
//
// Round 1 ran every PCG iteration as five launches (streamed Schur product, update, coarse restriction, coarse
// mat-vec, direction) with a host synchronisation every eight iterations: at C3 on one B200 the four small kernels
// cost 58 us of a 308 us iteration, on eight GPUs they were the iteration.  Here one cooperative launch (one CTA per
// SM, the CTAs of the streamed product) runs all iterations; the phases of an iteration are separated by grid
// barriers (one atomic + one polled word in L2, ~2 us) and the stopping test runs on the device:
//
//   A  q -= Es Es^T p                 the streamed product (sp_pass<SP_FULL>, TMA ring), w.w into the scalar table
//   B  x += a p, r -= a q, z = M^-1 r, r.z, per-unit partial sums of the coarse restriction P^T r
//   C  yc = A_c^-1 P^T r              dense symmetric mat-vec, one warp per row, rows spread over the whole grid
//   D  stop test; p = z + P yc + b p; q = (B + lambda) p; p.q
//
// Phases B and D work on *units*: at most kUnitCams cameras of one cluster of the coarse space (six threads per
// camera, one pass of the CTA).  Every CTA owns a fixed, contiguous range of units, so x, r, z, p are only ever
// touched by the CTA that owns them; q (reduced into by phase A of all CTAs), the restriction partials and yc cross
// CTAs and are read with ld.cg after the barrier.  The PCG scalars live in a table with one row per iteration
// (zeroed by the host before the launch), so no accumulator is ever recycled inside the kernel.
#pragma once
#include "ba_coarse.cuh"
#include "ba_kernels.cuh"
#include "ba_schur.cuh"

namespace vmm {

constexpr int kUnitCams = kSpThreads / 6;   // 80 cameras: one pass of the CTA's 480 threads

struct __align__(16) PcgUnit {
    int cluster;   // cluster of the coarse space (-1: none)
    int first;     // position of the unit's first camera in the unit camera list
    int count;     // cameras (<= kUnitCams)
    int pad;
};

enum { PS_RZ = 0, PS_WW = 1, PS_PBP = 2, PS_STRIDE = 4 };   // scalar table: row k = (r.z, w.w, p.(B+lambda)p) of iteration k

struct PcgArgs {
    // streamed product (SP_FULL segment list)
    const SpSeg* segs; const int* cta_seg; int nseg; const int2* slot_desc; int nslots_total;
    const unsigned char* Es; const int* mo_cam;
    // units
    const PcgUnit* units; const int* cta_unit; const int* unit_cams;
    // camera-space data
    const double* Minv; const double* B; const double* lam;
    double *x, *r, *z, *p, *q;
    // coarse level (ncl == 0: block-Jacobi only)
    int ncl, nC, ldc;         // ldc: leading dimension of Ainv (nC rounded up to the inversion kernel's block order)
    const int* coarse_info;   // != 0: the inversion of A_c met a non-positive pivot, run block-Jacobi only
    const int* cl_unit_ptr;   // units of cluster c: [cl_unit_ptr[c], cl_unit_ptr[c + 1])
    const double* xbar; const double* camP; const double* Ainv;
    double* rc_part;          // 6 per unit
    double* yc;
    // control
    double* scal;             // (max_it + 2) rows of PS_STRIDE
    double* result;           // sc[]: SC_DONE, SC_ITERS, SC_REL2, SC_RZ0 are written at the end
    unsigned* barrier;        // [0] arrivals, [1] generation, [2] abort flag (grid_barrier)
    int max_it, min_it;
    double tol2;
};

// bar[0] arrivals, bar[1] generation, bar[2] abort flag.  A CTA that polls for more than ~2 s (a lost CTA: the launch was
// not co-resident, or another CTA faulted) raises the abort flag; every barrier then returns false and the kernel
// winds down instead of hanging the device.
__device__ __forceinline__ bool grid_barrier(unsigned* bar, unsigned nblocks) {
    __shared__ int s_abort;
    __syncthreads();
    if (threadIdx.x == 0) {
        // release / acquire at gpu scope on the two words instead of three sequentially consistent fences: the CTA's
        // writes (ordered before this thread by the bar.sync above) are released by the arriving atomic, the last
        // arrival releases the new generation, the waiters acquire it.
        unsigned g, old, cur, ab;
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(bar + 1) : "memory");
        asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(ab) : "l"(bar + 2) : "memory");
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(old) : "l"(bar) : "memory");
        if (old == nblocks - 1) {
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(bar), "r"(0u) : "memory");
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(bar + 1), "r"(g + 1) : "memory");
        } else {
            const long long t0 = clock64();
            unsigned spins = 0;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(cur) : "l"(bar + 1) : "memory");
                if (cur != g) break;
                if ((++spins & 1023u) == 0) {
                    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(ab) : "l"(bar + 2) : "memory");
                    if (ab) break;
                    if (clock64() - t0 > 4000000000LL) {
                        asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(bar + 2), "r"(1u) : "memory");
                        ab = 1;
                        break;
                    }
                }
            } while (true);
        }
        s_abort = (int)ab;
    }
    __syncthreads();
    return s_abort == 0;
}

// block sum over all kSpThreads threads (15 warps); result valid in every thread
__device__ __forceinline__ double cta_sum(double v, double* s_red) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSpThreads / 32; ++w) t += s_red[w];
    return t;
}

// Phase B of one unit.  first: x = 0, r is the right-hand side (no p, q yet).
__device__ __forceinline__ double pcg_update_unit(const PcgArgs& A, const PcgUnit u, bool first, double alpha, double* s_v, double* s_c,
                                                  double* s_tot, int unit_index, bool coarse) {
    const int t = threadIdx.x, cs = t / 6, a = t - cs * 6;
    const bool valid = cs < u.count;
    int cam = 0;
    double rv = 0.0;
    size_t idx = 0;
    if (valid) {
        cam = __ldg(A.unit_cams + u.first + cs);
        idx = (size_t)cam * 6 + a;
        if (first) {
            A.x[idx] = 0.0;
            rv = A.r[idx];
        } else {
            A.x[idx] += alpha * A.p[idx];
            rv = A.r[idx] - alpha * __ldcg(A.q + idx);
            A.r[idx] = rv;
        }
    }
    __syncthreads();   // s_v / s_c of the previous unit are no longer read
    s_v[t] = rv;
    __syncthreads();
    double rz = 0.0, contrib = 0.0;
    const bool restrict_unit = coarse && u.cluster >= 0;   // uniform for the CTA
    if (valid) {
        const double* rr = s_v + cs * 6;
        const double zt = row_dot6(A.Minv + idx * 6, rr);
        A.z[idx] = zt;
        rz = zt * rv;
        if (restrict_unit) {
            // column k of R_i against the translation (a < 3) or rotation (a >= 3) half of r: V = sum R^T r_t, W = sum R^T r_theta
            const int k = a < 3 ? a : a - 3;
            const double* R = A.camP + (size_t)cam * CAM_PREP;
            const double* r3 = rr + (a < 3 ? 0 : 3);
            contrib = __ldg(R + k) * r3[0] + __ldg(R + 3 + k) * r3[1] + __ldg(R + 6 + k) * r3[2];
        }
    }
    if (restrict_unit) {
        s_c[t] = contrib;
        __syncthreads();
        // 96 threads: component a2 = t / 16 summed over 16 parts of 5 cameras, then over the half-warp
        if (t < 96) {
            const int a2 = t >> 4, part = t & 15;
            double s = 0.0;
#pragma unroll
            for (int i = 0; i < kUnitCams / 16; ++i) s += s_c[(part * (kUnitCams / 16) + i) * 6 + a2];
            s += __shfl_xor_sync(0xffffffffu, s, 8);
            s += __shfl_xor_sync(0xffffffffu, s, 4);
            s += __shfl_xor_sync(0xffffffffu, s, 2);
            s += __shfl_xor_sync(0xffffffffu, s, 1);
            if (part == 0) s_tot[a2] = s;
        }
        __syncthreads();
        if (t < 6) {
            // P^T r = [ -V ; -(V x xbar) - W / 2 ]
            const double* xb = A.xbar + 3 * (size_t)u.cluster;
            const double V0 = s_tot[0], V1 = s_tot[1], V2 = s_tot[2];
            double o;
            if (t < 3) o = -s_tot[t];
            else if (t == 3) o = -(V1 * xb[2] - V2 * xb[1]) - 0.5 * s_tot[3];
            else if (t == 4) o = -(V2 * xb[0] - V0 * xb[2]) - 0.5 * s_tot[4];
            else o = -(V0 * xb[1] - V1 * xb[0]) - 0.5 * s_tot[5];
            A.rc_part[(size_t)unit_index * 6 + t] = o;
        }
    }
    return rz;
}

// Phase D of one unit: p = z + P yc + beta p, q = (B + lambda) p; returns the thread's share of p.q
__device__ __forceinline__ double pcg_direction_unit(const PcgArgs& A, const PcgUnit u, bool first, double beta, double* s_v, bool coarse) {
    const int t = threadIdx.x, cs = t / 6, a = t - cs * 6;
    const bool valid = cs < u.count;
    double pv = 0.0;
    size_t idx = 0;
    if (valid) {
        const int cam = __ldg(A.unit_cams + u.first + cs);
        idx = (size_t)cam * 6 + a;
        pv = A.z[idx];
        if (u.cluster >= 0 && coarse) {
            const double* y = A.yc + (size_t)u.cluster * 6;
            const double* xb = A.xbar + 3 * (size_t)u.cluster;
            const int k = a < 3 ? a : a - 3;
            const double* R = A.camP + (size_t)cam * CAM_PREP + 3 * k;
            const double y3 = __ldcg(y + 3), y4 = __ldcg(y + 4), y5 = __ldcg(y + 5);
            if (a >= 3) {
                pv += -0.5 * (__ldg(R) * y3 + __ldg(R + 1) * y4 + __ldg(R + 2) * y5);
            } else {
                const double ux = __ldcg(y) + (xb[1] * y5 - xb[2] * y4);
                const double uy = __ldcg(y + 1) + (xb[2] * y3 - xb[0] * y5);
                const double uz = __ldcg(y + 2) + (xb[0] * y4 - xb[1] * y3);
                pv += -(__ldg(R) * ux + __ldg(R + 1) * uy + __ldg(R + 2) * uz);
            }
        }
        if (!first) pv += beta * A.p[idx];
        A.p[idx] = pv;
    }
    __syncthreads();
    s_v[t] = pv;
    __syncthreads();
    double pq = 0.0;
    if (valid) {
        const double qt = __ldg(A.lam + idx) * pv + row_dot6(A.B + idx * 6, s_v + cs * 6);
        A.q[idx] = qt;
        pq = qt * pv;
    }
    return pq;
}

// Phase C: rc = sum of the unit partials (every CTA forms the whole vector in shared memory), then the rows of
// A_c^-1 rc this CTA's warps own.  Returns (in every thread) the CTA's share of rc . yc.
__device__ __forceinline__ double pcg_coarse_apply(const PcgArgs& A, double* s_rc, double* s_red) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    __syncthreads();
    for (int e = t; e < A.nC; e += kSpThreads) {
        const int c = e / 6, comp = e - c * 6;
        double s = 0.0;
        for (int u = __ldg(A.cl_unit_ptr + c); u < __ldg(A.cl_unit_ptr + c + 1); ++u) s += __ldcg(A.rc_part + (size_t)u * 6 + comp);
        s_rc[e] = s;
    }
    __syncthreads();
    double dot = 0.0;
    const int nwarps = kSpThreads / 32;
    for (int row = blockIdx.x * nwarps + warp; row < A.nC; row += gridDim.x * nwarps) {
        const double2* a2 = reinterpret_cast<const double2*>(A.Ainv + (size_t)row * A.ldc);
        const double2* x2 = reinterpret_cast<const double2*>(s_rc);
        double s = 0.0;
        for (int k = lane; k < A.nC / 2; k += 32) {
            const double2 av = __ldg(a2 + k), xv = x2[k];
            s += av.x * xv.x + av.y * xv.y;
        }
        s = warp_sum(s);
        if (lane == 0) {
            A.yc[row] = s;
            dot += s * s_rc[row];
        }
    }
    const double total = cta_sum(dot, s_red);   // (its barriers also end the reads of s_rc)
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // s_rc aliases the ring: generic writes before the next bulk copies
    return total;
}

__global__ void __launch_bounds__(kSpThreads, 1)
k_pcg_persistent(const PcgArgs A) {
    const bool coarse = A.ncl > 0 && __ldcg(A.coarse_info) == 0;   // uniform for the grid
    extern __shared__ __align__(128) unsigned char sp_ring[];
    __shared__ SpShared sh;
    __shared__ double s_v[kSpThreads], s_c[kSpThreads], s_tot[8], s_red[kSpThreads / 32];
    double* s_rc = reinterpret_cast<double*>(sp_ring);   // phase C scratch: the ring is idle between products
    const unsigned G = gridDim.x;
    const int ub = A.cta_unit[blockIdx.x], ue = A.cta_unit[blockIdx.x + 1];
    const PcgUnit* units = A.units;
    volatile double* scal = A.scal;
    bool aborted = false;
    int it = 0, iterations = 0;
    bool ok = true, ring_used = false;
    double rel2 = 1.0, rz0 = 0.0;

    // ---- start: z = M^-1 r, r.z, restriction; coarse solve; first direction ----
    {
        double rz = 0.0;
        for (int u = ub; u < ue; ++u) rz += pcg_update_unit(A, units[u], true, 0.0, s_v, s_c, s_tot, u, coarse);
        rz = cta_sum(rz, s_red);
        if (threadIdx.x == 0) red_add(A.scal + PS_RZ, rz);
    }
    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    if (coarse) {
        const double d = pcg_coarse_apply(A, s_rc, s_red);
        if (threadIdx.x == 0) red_add(A.scal + PS_RZ, d);
        if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    }
    {
        double pq = 0.0;
        for (int u = ub; u < ue; ++u) pq += pcg_direction_unit(A, units[u], true, 0.0, s_v, coarse);
        pq = cta_sum(pq, s_red);
        if (threadIdx.x == 0) red_add(A.scal + PS_PBP, pq);
    }
    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    rz0 = scal[PS_RZ];
    ok = isfinite(rz0);
    if (ok && rz0 > 0.0) {
        for (it = 0; it < A.max_it; ++it) {
            double* row = A.scal + (size_t)it * PS_STRIDE;
            double* nxt = row + PS_STRIDE;
            // ---- A: q -= Es Es^T p ----
            sp_pass<SP_FULL, true>(sh, sp_ring, ring_used, A.segs, A.cta_seg, A.nseg, A.slot_desc, A.nslots_total, A.Es, A.mo_cam, A.p, nullptr,
                                   -1.0, A.q, row + PS_WW);
            ring_used = true;
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            // ---- B ----
            const double rz_s = scal[(size_t)it * PS_STRIDE + PS_RZ];
            const double denom = scal[(size_t)it * PS_STRIDE + PS_PBP] - scal[(size_t)it * PS_STRIDE + PS_WW];
            const double alpha = (rz_s != 0.0) ? rz_s / denom : 0.0;
            {
                double rz = 0.0;
                for (int u = ub; u < ue; ++u) rz += pcg_update_unit(A, units[u], false, alpha, s_v, s_c, s_tot, u, coarse);
                rz = cta_sum(rz, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_RZ, rz);
            }
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            // ---- C ----
            if (coarse) {
                const double d = pcg_coarse_apply(A, s_rc, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_RZ, d);
                if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            }
            // ---- D ----
            const double rz_new = scal[(size_t)(it + 1) * PS_STRIDE + PS_RZ];
            iterations = it + 1;
            rel2 = rz_new / rz0;
            if (!isfinite(rz_new)) { ok = false; break; }
            if (iterations >= A.min_it && rz_new <= A.tol2 * rz0) break;
            const double beta = (rz_s != 0.0) ? rz_new / rz_s : 0.0;
            {
                double pq = 0.0;
                for (int u = ub; u < ue; ++u) pq += pcg_direction_unit(A, units[u], false, beta, s_v, coarse);
                pq = cta_sum(pq, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_PBP, pq);
            }
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
        }
    } else if (ok) {
        rel2 = 0.0;   // zero right-hand side: x = 0 is the solution
    }
finish:
    if (aborted) ok = false;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        A.result[SC_RZ0] = rz0;
        A.result[SC_ITERS] = (double)iterations;
        A.result[SC_REL2] = ok ? rel2 : nan("");
        A.result[SC_DONE] = 1.0;
    }
}

}  // namespace vmm
