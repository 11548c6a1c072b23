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
// Phases B and D work on the cameras in cluster order: every CTA owns a fixed, contiguous range of that list (six lanes
// per camera, five cameras per warp step, no block barrier inside the loops), so x, r, z, p are only ever touched by
// the CTA that owns them; q (reduced into by phase A of all CTAs), the restriction partials and yc cross
// CTAs and are read with ld.cg after the barrier.  The PCG scalars live in a table with one row per iteration
// (zeroed by the host before the launch), so no accumulator is ever recycled inside the kernel.
#pragma once
#include "ba_coarse.cuh"
#include "ba_kernels.cuh"
#include "ba_schur.cuh"

namespace vmm {

constexpr int kPcgMaxSlots = 16;   // clusters one CTA's camera range may touch (host: else the clusters are too small for this path)

enum { PS_RZ = 0, PS_WW = 1, PS_PBP = 2, PS_STRIDE = 4 };   // scalar table: row k = (r.z, w.w, p.(B+lambda)p) of iteration k

struct PcgArgs {
    // streamed product (SP_FULL segment list)
    const SpSeg* segs; const int* cta_seg; int nseg; const int2* slot_desc; int nslots_total;
    const unsigned char* Es; const int* mo_cam;
    // cameras in cluster order (cluster of every position; -1 without a coarse space), per-CTA ranges of that list,
    // per CTA its first cluster and the index of its first restriction partial
    const int* cams; const int* cam_cl; const int* cta_cam; const int* cta_cl_first; const int* cta_part; int nparts;
    // camera-space data
    const double* Minv; const double* B; const double* lam;
    double *x, *r, *z, *p, *q;
    // coarse level (ncl == 0: block-Jacobi only)
    int ncl, nC, ldc;         // ldc: leading dimension of Ainv (nC rounded up to the inversion kernel's block order)
    const int* coarse_info;   // != 0: the inversion of A_c met a non-positive pivot, run block-Jacobi only
    const int* cl_part_ptr;   // restriction partials of cluster c: [cl_part_ptr[c], cl_part_ptr[c + 1])
    const double* xbar; const double* camP; const double* Ainv;
    double* rc_part;          // 6 per (CTA, cluster it touches)
    double* yc;
    // control
    double* scal;             // (max_it + 2) rows of PS_STRIDE
    double* result;           // sc[]: SC_DONE, SC_ITERS, SC_REL2, SC_RZ0 are written at the end
    unsigned* barrier;        // [0] arrivals, [1] generation, [2] abort flag (grid_barrier)
    int max_it, min_it;
    double tol2;
    int resume;               // 1: continue the solve the previous launch stopped (same system, tighter tolerance): x, r, z, p and the
                              //    scalars it left in `result` are taken up where they were
    unsigned long long* phase_ns;   // optional (debug / bench): CTA 0 accumulates the time of the phases A, B, C, D, barriers excluded [0..3], barriers [4];
                                    // [8 + b]: time CTA b spent in its streamed product (load balance of phase A)
};

// The arguments live in constant memory (copied there, stream-ordered, before every launch): every phase — each its own
// function with its own register allocation — reads the fields it needs from the constant bank with immediate offsets,
// so no pointer is passed, kept in a register across a call or spilled.  (With the arguments as a kernel parameter the
// phases had to be inlined into one body to see the parameter bank, and that body spilled inside the camera loops; with
// 184 KB of the SM's memory in the ring there is no L1 left to catch local memory.)  One solve at a time per device and
// process, which the handle's threading rule (include/vmm_ba.h) already demands.
__constant__ PcgArgs c_pcg;

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
                asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(cur) : "l"(bar + 1) : "memory");   // the acquire comes once, after the loop
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
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
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

// ---- phases B and D: warp-autonomous.  A warp works on five cameras at a time (six lanes per camera, lanes 30 / 31
// idle); a camera's 6-vector is exchanged with shuffles, so there is no block barrier inside the loop and the fifteen
// warps stream through the CTA's camera range independently, each with the next chunk's loads in flight. ----
// the camera (and its cluster) at position pos of the cluster-ordered list; -1 beyond the CTA's range / for the idle lanes
__device__ __forceinline__ int2 pcg_chunk_index(int pos, int pos_end, bool lane_ok) {
    if (lane_ok && pos < pos_end) return make_int2(__ldg(c_pcg.cams + pos), __ldg(c_pcg.cam_cl + pos));
    return make_int2(-1, -1);
}

// x += alpha p, r -= alpha q, z = M^-1 r over the CTA's cameras, the restriction P^T r per cluster the CTA touches
// (written to rc_part); returns the thread's share of r.z.  s_acc: kPcgMaxSlots x 6 doubles.
__device__ __noinline__ double pcg_phase_update(bool first, double alpha, bool coarse, double* s_acc) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nwarps = kSpThreads / 32;
    const int g = lane / 6, a = lane - 6 * g, k = a < 3 ? a : a - 3;
    const bool lane_ok = lane < 30;
    const int c0 = __ldg(c_pcg.cta_cam + blockIdx.x), c1 = __ldg(c_pcg.cta_cam + blockIdx.x + 1);
    const int cl_first = __ldg(c_pcg.cta_cl_first + blockIdx.x);
    for (int i = t; i < kPcgMaxSlots * 6; i += kSpThreads) s_acc[i] = 0.0;
    __syncthreads();
    double rz = 0.0, acc = 0.0;
    int acc_cl = -1;   // lanes 0..5: running restriction sum of component `lane` for cluster acc_cl
    // only the camera index of the NEXT chunk is requested ahead (one dependent L2 access less per chunk)
    int base = c0 + warp * 5;
    int2 cc = pcg_chunk_index(base + g, c1, lane_ok);
    for (; base < c1; base += nwarps * 5) {
        const int cam = cc.x, cl = cc.y;
        cc = pcg_chunk_index(base + nwarps * 5 + g, c1, lane_ok);
        const bool valid = cam >= 0;
        const size_t idx = (size_t)max(cam, 0) * 6 + a;
        double rv = 0.0;
        double2 m0 = make_double2(0.0, 0.0), m1 = m0, m2 = m0;
        double R0 = 0.0, R1 = 0.0, R2 = 0.0;
        if (valid) {
            const double2* m = reinterpret_cast<const double2*>(c_pcg.Minv + idx * 6);
            m0 = __ldg(m); m1 = __ldg(m + 1); m2 = __ldg(m + 2);
            if (coarse && cl >= 0) {
                const double* R = c_pcg.camP + (size_t)cam * CAM_PREP;
                R0 = R[k]; R1 = R[3 + k]; R2 = R[6 + k];
            }
            // all operands first, then the stores: the vectors may alias as far as the compiler knows, so a load written
            // behind a store waits for the store's operands (one L2 round trip each)
            rv = c_pcg.r[idx];
            double xv = 0.0, pvv = 0.0, qv = 0.0;
            if (!first) { xv = c_pcg.x[idx]; pvv = c_pcg.p[idx]; qv = c_pcg.q[idx]; }
            rv -= alpha * qv;
            c_pcg.x[idx] = xv + alpha * pvv;
            if (!first) c_pcg.r[idx] = rv;
        }
        const int src = min(6 * g, 26);
        const double r0 = __shfl_sync(0xffffffffu, rv, src), r1 = __shfl_sync(0xffffffffu, rv, src + 1), r2 = __shfl_sync(0xffffffffu, rv, src + 2);
        const double r3 = __shfl_sync(0xffffffffu, rv, src + 3), r4 = __shfl_sync(0xffffffffu, rv, src + 4), r5 = __shfl_sync(0xffffffffu, rv, src + 5);
        double contrib = 0.0;
        if (valid) {
            const double zt = m0.x * r0 + m0.y * r1 + m1.x * r2 + m1.y * r3 + m2.x * r4 + m2.y * r5;
            c_pcg.z[idx] = zt;
            rz += zt * rv;
            // column k of R_i against the translation (a < 3) or rotation (a >= 3) half of r: V = sum R^T r_t, W = sum R^T r_theta
            contrib = a < 3 ? R0 * r0 + R1 * r1 + R2 * r2 : R0 * r3 + R1 * r4 + R2 * r5;
        }
        if (coarse) {
            // lanes 0..5 collect component `lane` of the five cameras, cluster by cluster
#pragma unroll
            for (int gg = 0; gg < 5; ++gg) {
                const double v = __shfl_sync(0xffffffffu, contrib, min(6 * gg + lane, 31));
                const int clg = __shfl_sync(0xffffffffu, cl, 6 * gg);
                if (lane < 6 && clg >= 0) {
                    if (clg != acc_cl) {
                        if (acc_cl >= 0) atomicAdd(&s_acc[(acc_cl - cl_first) * 6 + lane], acc);
                        acc_cl = clg;
                        acc = 0.0;
                    }
                    acc += v;
                }
            }
        }
    }
    if (coarse && lane < 6 && acc_cl >= 0) atomicAdd(&s_acc[(acc_cl - cl_first) * 6 + lane], acc);
    __syncthreads();
    if (coarse) {
        // P^T r = [ -V ; -(V x xbar) - W / 2 ] per touched cluster
        const int p0 = __ldg(c_pcg.cta_part + blockIdx.x), nslots = __ldg(c_pcg.cta_part + blockIdx.x + 1) - p0;
        if (t < nslots * 6) {
            const int slot = t / 6, comp = t - 6 * slot;
            const double* sv = s_acc + slot * 6;
            const double* xb = c_pcg.xbar + 3 * (size_t)(cl_first + slot);
            double o;
            if (comp < 3) o = -sv[comp];
            else if (comp == 3) o = -(sv[1] * xb[2] - sv[2] * xb[1]) - 0.5 * sv[3];
            else if (comp == 4) o = -(sv[2] * xb[0] - sv[0] * xb[2]) - 0.5 * sv[4];
            else o = -(sv[0] * xb[1] - sv[1] * xb[0]) - 0.5 * sv[5];
            c_pcg.rc_part[(size_t)(p0 + slot) * 6 + comp] = o;
        }
    }
    return rz;
}

// p = z + P yc + beta p, q = (B + lambda) p over the CTA's cameras; returns the thread's share of p.q
__device__ __noinline__ double pcg_phase_direction(bool first, double beta, bool coarse) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5, nwarps = kSpThreads / 32;
    const int g = lane / 6, a = lane - 6 * g, k = a < 3 ? a : a - 3;
    const bool lane_ok = lane < 30;
    const int c0 = __ldg(c_pcg.cta_cam + blockIdx.x), c1 = __ldg(c_pcg.cta_cam + blockIdx.x + 1);
    double pq = 0.0;
    // the coarse correction of a cluster as this lane needs it: u = y_tau + xbar x y_theta (rows 0-2) or y_theta (rows 3-5)
    int cl_cached = -1;
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    int base = c0 + warp * 5;
    int2 cc = pcg_chunk_index(base + g, c1, lane_ok);
    for (; base < c1; base += nwarps * 5) {
        const int cam = cc.x, cl = cc.y;
        cc = pcg_chunk_index(base + nwarps * 5 + g, c1, lane_ok);
        const bool valid = cam >= 0;
        const size_t idx = (size_t)max(cam, 0) * 6 + a;
        double pv = 0.0, lam = 0.0;
        double2 b0 = make_double2(0.0, 0.0), b1 = b0, b2 = b0;
        if (valid) {
            const double2* b = reinterpret_cast<const double2*>(c_pcg.B + idx * 6);
            b0 = __ldg(b); b1 = __ldg(b + 1); b2 = __ldg(b + 2);
            lam = __ldg(c_pcg.lam + idx);
            pv = c_pcg.z[idx];
            const double p_old = first ? 0.0 : c_pcg.p[idx];   // requested with the other operands (see pcg_phase_update)
            if (coarse && cl >= 0) {
                const double* R = c_pcg.camP + (size_t)cam * CAM_PREP + 3 * k;
                const double R0 = R[0], R1 = R[1], R2 = R[2];
                if (cl != cl_cached) {
                    cl_cached = cl;
                    const double* y = c_pcg.yc + (size_t)cl * 6;
                    const double y3 = y[3], y4 = y[4], y5 = y[5];
                    if (a >= 3) {
                        u0 = y3; u1 = y4; u2 = y5;
                    } else {
                        const double* xb = c_pcg.xbar + 3 * (size_t)cl;
                        const double x0 = __ldg(xb), x1 = __ldg(xb + 1), x2 = __ldg(xb + 2);
                        u0 = y[0] + (x1 * y5 - x2 * y4);
                        u1 = y[1] + (x2 * y3 - x0 * y5);
                        u2 = y[2] + (x0 * y4 - x1 * y3);
                    }
                }
                pv += (a >= 3 ? -0.5 : -1.0) * (R0 * u0 + R1 * u1 + R2 * u2);
            }
            pv += beta * p_old;
            c_pcg.p[idx] = pv;
        }
        const int src = min(6 * g, 26);
        const double p0 = __shfl_sync(0xffffffffu, pv, src), p1 = __shfl_sync(0xffffffffu, pv, src + 1), p2 = __shfl_sync(0xffffffffu, pv, src + 2);
        const double p3 = __shfl_sync(0xffffffffu, pv, src + 3), p4 = __shfl_sync(0xffffffffu, pv, src + 4), p5 = __shfl_sync(0xffffffffu, pv, src + 5);
        if (valid) {
            const double qt = lam * pv + b0.x * p0 + b0.y * p1 + b1.x * p2 + b1.y * p3 + b2.x * p4 + b2.y * p5;
            c_pcg.q[idx] = qt;
            pq += qt * pv;
        }
    }
    return pq;
}

// The interior pass of the plain kernel: all arguments from constant memory.
__device__ __noinline__ void pcg_phase_product(SpShared& sh, unsigned char* sp_ring, bool ring_used, double* dot_acc) {
    const PcgArgs& A = c_pcg;
    sp_pass<SP_FULL, true>(sh, sp_ring, ring_used, A.segs, A.cta_seg, A.nseg, A.slot_desc, A.nslots_total, A.Es, A.mo_cam, A.p, nullptr, -1.0, A.q, dot_acc);
}

#ifdef VMM_DBG_PASS_INLINE
#define VMM_PASS_CALL_ATTR __forceinline__
#else
#define VMM_PASS_CALL_ATTR __noinline__
#endif
template <int MODE>
__device__ VMM_PASS_CALL_ATTR void sp_pass_call(SpShared& sh, unsigned char* sp_ring, bool ring_used, const SpSeg* segs, const int* cta_seg, int nseg,
                                          const int2* slot_desc, int nslots_total, const unsigned char* Es, const int* mo_cam, const double* vin_c,
                                          double* mvec, double sign, double* out_c, double* dot_acc) {
    sp_pass<MODE, true>(sh, sp_ring, ring_used, segs, cta_seg, nseg, slot_desc, nslots_total, Es, mo_cam, vin_c, mvec, sign, out_c, dot_acc);
}

// rc = sum of the restriction partials: all partials into shared memory with independent loads, then summed per cluster.
// s_rc: nC (+1) doubles; s_part: 6 per partial (both in the idle ring).
__device__ __forceinline__ void pcg_assemble_rc(int nC, double* s_rc, double* s_part) {
    const PcgArgs& A = c_pcg;
    const int t = threadIdx.x;
    for (int e = t; e < 6 * A.nparts; e += kSpThreads) s_part[e] = A.rc_part[e];
    __syncthreads();
    for (int e = t; e < nC; e += kSpThreads) {
        const int c = e / 6, comp = e - c * 6;
        double s = 0.0;
        for (int u = __ldg(A.cl_part_ptr + c); u < __ldg(A.cl_part_ptr + c + 1); ++u) s += s_part[u * 6 + comp];
        s_rc[e] = s;
    }
}

// Phase C: every CTA forms the whole restricted residual in shared memory, then the rows of A_c^-1 rc this CTA's warps
// own.  Returns (in every thread) the CTA's share of rc . yc.
__device__ __noinline__ double pcg_coarse_apply(double* s_rc, double* s_red) {
    const PcgArgs& A = c_pcg;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    __syncthreads();
    pcg_assemble_rc(A.nC, s_rc, s_rc + ((A.nC + 8) & ~1));
    __syncthreads();
    double dot = 0.0;
    const int nwarps = kSpThreads / 32;
    for (int row = blockIdx.x * nwarps + warp; row < A.nC; row += gridDim.x * nwarps) {
        const double2* a2 = reinterpret_cast<const double2*>(A.Ainv + (size_t)row * A.ldc);
        const double2* x2 = reinterpret_cast<const double2*>(s_rc);
        double s = 0.0;
        for (int k = lane; k < A.nC / 2; k += 32) {
            const double2 av = a2[k], xv = x2[k];
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
k_pcg_persistent() {
    const PcgArgs& A = c_pcg;
    const bool coarse = A.ncl > 0 && __ldcg(A.coarse_info) == 0;   // uniform for the grid
    extern __shared__ __align__(128) unsigned char sp_ring[];
    __shared__ SpShared sh;
    __shared__ double s_acc[kPcgMaxSlots * 6], s_red[kSpThreads / 32];
    double* s_rc = reinterpret_cast<double*>(sp_ring);   // phase C scratch: the ring is idle between products
    const unsigned G = gridDim.x;
    volatile double* scal = A.scal;
    unsigned long long t_prev = 0;
    // phase clock (CTA 0, thread 0; only when the host asked for it): time since the previous tick into slot `k`
    auto tick = [&](int k) {
        if (A.phase_ns && blockIdx.x == 0 && threadIdx.x == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (k >= 0) A.phase_ns[k] += now - t_prev;
            t_prev = now;
        }
    };
    // ONE loop, every phase inlined exactly once (the start, the resumed start and the iterations are the same code
    // under flags: three inlined copies of each phase made the register allocator spill inside the camera loops, and
    // with 184 KB of the SM's memory in the ring there is no L1 left for local memory).
    //   pass -1 (fresh):    B(first) -> C -> D(first)           rows: r.z_0, p.q_0
    //   pass -1 (resumed):  D(beta from the stopped launch)
    //   pass it >= 0:       A -> B -> C -> stop test -> D
    bool first = !A.resume, resumed = A.resume != 0, aborted = false, ok = true, ring_used = false;
    int it = -1, iterations = 0, it_base = 0;
    double rel2 = 1.0, rz0 = 0.0, rz_s = 0.0, rz_new = 0.0, rz_last = 0.0, rz_before = 0.0;
    if (resumed) {
        rz0 = __ldcg(A.result + SC_RZ0);
        rz_new = __ldcg(A.result + SC_RZ);
        rz_s = __ldcg(A.result + SC_RZ + 1);
        it_base = (int)__ldcg(A.result + SC_ITERS);
        if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }   // every CTA has read the scalars of the stopped launch
        if (blockIdx.x == 0 && threadIdx.x == 0) A.scal[PS_RZ] = rz_new;
    }
    tick(-1);
    for (;;) {
        double* row = A.scal + (size_t)max(it, 0) * PS_STRIDE;
        double* nxt = A.scal + (size_t)(it + 1) * PS_STRIDE;
        if (it >= 0) {
            // ---- A: q -= Es Es^T p ----
            unsigned long long ta = 0;
            if (A.phase_ns && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ta));
            pcg_phase_product(sh, sp_ring, ring_used, row + PS_WW);
            ring_used = true;
            if (A.phase_ns) {
                __syncthreads();
                if (threadIdx.x == 0) {
                    unsigned long long tb;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tb));
                    A.phase_ns[8 + blockIdx.x] += tb - ta;
                }
            }
            tick(0);
            if (!grid_barrier(A.barrier, G)) { aborted = true; break; }
            tick(4);
        }
        if (!resumed) {
            // ---- B ----
            double alpha = 0.0;
            if (it >= 0) {
                rz_s = scal[(size_t)it * PS_STRIDE + PS_RZ];
                const double denom = scal[(size_t)it * PS_STRIDE + PS_PBP] - scal[(size_t)it * PS_STRIDE + PS_WW];
                alpha = (rz_s != 0.0) ? rz_s / denom : 0.0;
            }
            {
                double rz = pcg_phase_update(first, alpha, coarse, s_acc);
                rz = cta_sum(rz, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_RZ, rz);
            }
            tick(1);
            if (!grid_barrier(A.barrier, G)) { aborted = true; break; }
            tick(4);
            // ---- C ----
            if (coarse) {
                const double d = pcg_coarse_apply(s_rc, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_RZ, d);
                tick(2);
                if (!grid_barrier(A.barrier, G)) { aborted = true; break; }
                tick(4);
            }
            rz_new = scal[(size_t)(it + 1) * PS_STRIDE + PS_RZ];
        }
        // ---- stop test ----
        if (first) {
            rz0 = rz_new;
            ok = isfinite(rz0);
            if (!ok) break;
            if (!(rz0 > 0.0)) { rel2 = 0.0; break; }   // zero right-hand side: x = 0 is the solution
        } else if (!resumed) {
            rz_last = rz_new; rz_before = rz_s;
            iterations = it + 1;
            rel2 = rz_new / rz0;
            if (!isfinite(rz_new)) { ok = false; break; }
            if (iterations >= A.min_it && rz_new <= A.tol2 * rz0) break;
            if (iterations >= A.max_it) break;
        }
        // ---- D ----
        {
            const double beta = (first || rz_s == 0.0) ? 0.0 : rz_new / rz_s;
            double pq = pcg_phase_direction(first, beta, coarse);
            pq = cta_sum(pq, s_red);
            if (threadIdx.x == 0) red_add(nxt + PS_PBP, pq);
        }
        tick(3);
        if (!grid_barrier(A.barrier, G)) { aborted = true; break; }
        tick(4);
        first = false;
        resumed = false;
        ++it;
    }
finish:
    if (aborted) ok = false;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        A.result[SC_RZ0] = rz0;
        A.result[SC_RZ] = rz_last;          // for a resumed launch: r.z at the stop and the one before it
        A.result[SC_RZ + 1] = rz_before;
        A.result[SC_ITERS] = (double)(it_base + iterations);
        A.result[SC_REL2] = ok ? rel2 : nan("");
        A.result[SC_DONE] = 1.0;
    }
}

}  // namespace vmm
