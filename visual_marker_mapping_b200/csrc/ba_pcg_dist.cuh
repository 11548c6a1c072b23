// ba_pcg_dist.cuh — the persistent PCG kernel of a camera-sharded engine (one rank per GPU) and of problems whose
// markers do not all fit one SP_FULL segment: ba_pcg.cuh's kernel plus the boundary path of the streamed product and
// the two per-iteration sums over ranks, exchanged INSIDE the kernel through NVLink peer memory.
//
// One PCG iteration on every rank (GB = grid barrier of this rank's CTAs):
//   A1  q -= Es Es^T p over the markers whose blocks all live here (SP_FULL, one pass)           \  no barrier
//   A2  wbl += Es^T p over this rank's blocks of the boundary markers (SP_GATHER, compact vector) /  in between
//   GB  X1: wb = sum over ranks of wbl, with this rank's p.(B+lambda)p and w.w partials in the tail — every thread
//       pushes its elements into its slot of EVERY rank's exchange buffer (each double as two 8-byte words that carry
//       the exchange's sequence number: data and flag arrive together, no fence, no flag round trip) and then sums the
//       slots of its own buffer in rank order, polling until the words carry the current number.  Bitwise identical
//       on all ranks.  Squares of wb per CTA for the boundary part of w.w.
//   GB  A3: q -= Es wb over the boundary blocks (SP_SCATTER)
//   GB  B: update, block-Jacobi part of z, r.z partial, restriction partials per unit
//   GB  X2 + C: CTA k (k < ranks) pushes this rank's restricted residual (r.z partial in its tail) to rank k; every CTA
//       polls its own buffer for all ranks' vectors, sums them in rank order into shared memory and applies its rows
//       of A_c^-1 (the inverse is replicated)
//   GB  D: r.z = tail + rc.yc in a FIXED order (every CTA of every rank computes the same bits, so all take the same
//       stopping decision), direction, q = (B + lambda) p, p.q partial
//   GB
// Without ranks to talk to (one GPU, split markers) X1 / X2 are local sums and the kernel is ba_pcg.cuh's with the
// A2 / A3 passes added.
#pragma once
#include "ba_pcg.cuh"
#include "ba_peer.cuh"

namespace vmm {

constexpr size_t kPcgxCap1 = 256 * 1024;   // words (one double each) per slot of the boundary exchange
constexpr size_t kPcgxCap2 = 2048;         // ... of the coarse exchange (order of A_c + 1 <= 1537)

struct PcgDistArgs {
    // boundary path of the product: segment list of the split / shared markers (marker field = position in wbl)
    const SpSeg* bsegs; const int* bcta_seg; int bnseg; const int2* bslot_desc; int bnslots_total;
    int n_wb;                 // entries of the boundary vector (6 per boundary marker); 0: no boundary path
    double* wbl;              // this rank's partial (zeroed by the kernel for the next iteration)
    double* wb;               // the sum over ranks (== wbl on one GPU)
    double* cta_part;         // gridDim doubles: squares of wb per CTA
    double* rc_local;         // nC + 1: this rank's restricted residual, r.z partial in the tail
    // ranks
    int rank, world;
    uint4* xbuf[kPeerMaxWorld];   // every rank's exchange buffer as mapped here: [X1: 2 parities][world][cap1] [X2: 2 parities][world][cap2]
    size_t cap1, cap2;
    unsigned* seq;            // device-resident sequence base (advanced by the kernel, the same on every rank)
    int debug_skip;           // -DVMM_DBG_TRACE builds only (VMM_BA_DIST_SKIP), 0 otherwise: bit 0 / 1 / 2 skip the interior / gather / scatter pass (wrong numbers)
};

__constant__ PcgDistArgs c_pcgd;

__device__ __forceinline__ uint4* x1_slot(const PcgDistArgs& D, int dst_rank, int parity, int src_rank) {
    return D.xbuf[dst_rank] + ((size_t)parity * D.world + src_rank) * D.cap1;
}
__device__ __forceinline__ uint4* x2_slot(const PcgDistArgs& D, int dst_rank, int parity, int src_rank) {
    return D.xbuf[dst_rank] + (size_t)2 * D.world * D.cap1 + ((size_t)parity * D.world + src_rank) * D.cap2;
}

// ld_word_pair of ba_peer.cuh with a way out: a word that does not arrive within ~2 s (a peer died or left the kernel)
// raises this rank's abort flag (the third word of the grid barrier), after which every barrier winds the kernel down.
__device__ __forceinline__ double ld_word_pair_guarded(const uint4* p, uint32_t seq, unsigned* abort_flag) {
    uint32_t lo, f1, hi, f2;
    unsigned spins = 0;
    long long t0 = 0;
    do {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(lo), "=r"(f1), "=r"(hi), "=r"(f2) : "l"(p) : "memory");
        if (f1 == seq && f2 == seq) break;
        if ((++spins & 4095u) == 0) {
            if (t0 == 0) t0 = clock64();
            if (clock64() - t0 > 4000000000LL || *reinterpret_cast<volatile unsigned*>(abort_flag) != 0) {
                *reinterpret_cast<volatile unsigned*>(abort_flag) = 1;
                return 0.0;
            }
        }
    } while (true);
    return __longlong_as_double((long long)(((unsigned long long)hi << 32) | lo));
}

// rc . yc in a fixed order, the same bits in every CTA of every rank: thread t sums its strided entries, the 15 warp
// totals are added in warp order.
__device__ __forceinline__ double pcg_fixed_dot(const double* s_rc, const double* yc, int n, double* s_red) {
    double s = 0.0;
    for (int k = threadIdx.x; k < n; k += kSpThreads) s += s_rc[k] * __ldcg(yc + k);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSpThreads / 32; ++w) t += s_red[w];
    return t;
}

// sum of v[0 .. n) in a fixed order, the same bits in every CTA of every rank (see pcg_fixed_dot)
__device__ __forceinline__ double pcg_fixed_sum(const double* v, int n, double* s_red) {
    double s = 0.0;
    for (int k = threadIdx.x; k < n; k += kSpThreads) s += __ldcg(v + k);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kSpThreads / 32; ++w) t += s_red[w];
    return t;
}

// X1: wb = sum over ranks of wbl (+ two scalars in the tail).  Returns the thread's share of sum wb_i^2.
template <bool SHARDED>
__device__ __forceinline__ double pcg_exchange_boundary(const PcgDistArgs& D, unsigned seq, double tail0, double tail1) {
    const int n = D.n_wb;
    const int stride = gridDim.x * kSpThreads;
    const int t0 = blockIdx.x * kSpThreads + threadIdx.x;
    double sq = 0.0;
    if (!SHARDED) {
        for (int i = t0; i < n; i += stride) { const double v = __ldcg(D.wbl + i); sq += v * v; }
        return sq;
    }
    const int parity = (int)(seq & 1u);
    for (int i = t0; i < n + 2; i += stride) {
        const double v = i < n ? __ldcg(D.wbl + i) : (i == n ? tail0 : tail1);
        for (int r = 0; r < D.world; ++r) st_word_pair(x1_slot(D, r, parity, D.rank) + i, v, seq);
    }
    for (int i = t0; i < n + 2; i += stride) {
        double s = 0.0;
        for (int r = 0; r < D.world; ++r) s += ld_word_pair_guarded(x1_slot(D, D.rank, parity, r) + i, seq, c_pcg.barrier + 2);
        D.wb[i] = s;
        if (i < n) sq += s * s;
    }
    return sq;
}

// X2 + C: the restricted residual summed over ranks into s_rc[0 .. nC] (tail: the block-Jacobi part of r.z), then this
// CTA's rows of yc = A_c^-1 rc.
template <bool SHARDED>
__device__ __forceinline__ void pcg_coarse_exchange_apply(const PcgDistArgs& D, const PcgArgs& A, unsigned seq, double rz_local, double* s_rc, int nC) {
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    __syncthreads();
    if (!SHARDED || (int)blockIdx.x < D.world) {
        // this rank's restricted residual from the unit partials (tail: its r.z partial)
        pcg_assemble_rc(nC, s_rc, s_rc + ((nC + 8) & ~1));
        if (t == 0) s_rc[nC] = rz_local;
        __syncthreads();
        if (SHARDED) {
            for (int e = t; e <= nC; e += kSpThreads) st_word_pair(x2_slot(D, (int)blockIdx.x, (int)(seq & 1u), D.rank) + e, s_rc[e], seq);
            __syncthreads();   // s_rc is overwritten with the sums over ranks next
        }
    }
    if (SHARDED) {
        const int parity = (int)(seq & 1u);
        for (int e = t; e <= nC; e += kSpThreads) {
            double s = 0.0;
            for (int r = 0; r < D.world; ++r) s += ld_word_pair_guarded(x2_slot(D, D.rank, parity, r) + e, seq, c_pcg.barrier + 2);
            s_rc[e] = s;
        }
    }
    __syncthreads();
    const int nwarps = kSpThreads / 32;
    for (int row = blockIdx.x * nwarps + warp; row < nC; row += gridDim.x * nwarps) {
        const double2* a2 = reinterpret_cast<const double2*>(A.Ainv + (size_t)row * A.ldc);
        const double2* x2 = reinterpret_cast<const double2*>(s_rc);
        double s = 0.0;
        for (int k = lane; k < nC / 2; k += 32) {
            const double2 av = a2[k], xv = x2[k];
            s += av.x * xv.x + av.y * xv.y;
        }
        s = warp_sum(s);
        if (lane == 0) A.yc[row] = s;
    }
}

#ifdef VMM_DBG_TRACE
#define KT(stage) do { if (g_sp_trace && threadIdx.x == 0) *reinterpret_cast<volatile unsigned*>(g_sp_trace + ((size_t)blockIdx.x * 16 + 15) * 4) = ((unsigned)it << 8) | (unsigned)(stage); } while (0)
#else
#define KT(stage) do { } while (0)
#endif
template <bool SHARDED>
__global__ void __launch_bounds__(kSpThreads, 1)
k_pcg_persistent_dist() {
    const PcgDistArgs& D = c_pcgd;   // constant memory, like c_pcg (ba_pcg.cuh)
    const PcgArgs& A = c_pcg;
    const bool coarse = A.ncl > 0 && __ldcg(A.coarse_info) == 0;   // uniform for the grid (and, the matrix being replicated, for all ranks)
    extern __shared__ __align__(128) unsigned char sp_ring[];
    __shared__ SpShared sh;
    __shared__ double s_acc[kPcgMaxSlots * 6], s_red[kSpThreads / 32];
    double* s_rc = reinterpret_cast<double*>(sp_ring);   // restricted residual (+ tail): the ring is idle between products
    const unsigned G = gridDim.x;
    const int nC = coarse ? A.nC : 0;
    const bool boundary = D.n_wb > 0;
    volatile double* scal = A.scal;
    const unsigned seq0 = SHARDED ? __ldcg(D.seq) : 0u;   // exchanges of this launch: X2 uses seq0 + 1, seq0 + 2, ...; X1 uses seq0 + 1 + it
    unsigned x2_count = 0;
    bool aborted = false, ok = true, ring_used = false;
    unsigned long long t_prev = 0;
    // phase clock as in k_pcg_persistent (CTA 0): [0] A1 + A2, [1] B, [2] X2 + C, [3] D, [4] barriers, [5] X1, [6] A3
    auto tick = [&](int k) {
        if (A.phase_ns && blockIdx.x == 0 && threadIdx.x == 0) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (k >= 0) A.phase_ns[k] += now - t_prev;
            t_prev = now;
        }
    };
    int it = 0, iterations = 0;
    double rel2 = 1.0, rz0 = 0.0;

    // ---- start ----
    {
        double rz = pcg_phase_update(true, 0.0, coarse, s_acc);
        rz = cta_sum(rz, s_red);
        if (threadIdx.x == 0) red_add(A.scal + PS_RZ, rz);
        if (boundary) for (int i = blockIdx.x * kSpThreads + threadIdx.x; i < D.n_wb; i += G * kSpThreads) D.wbl[i] = 0.0;
    }
    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    pcg_coarse_exchange_apply<SHARDED>(D, A, seq0 + (++x2_count), scal[PS_RZ], s_rc, nC);
    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    rz0 = s_rc[nC] + pcg_fixed_dot(s_rc, A.yc, nC, s_red);
    {
        double pq = pcg_phase_direction(true, 0.0, coarse);
        pq = cta_sum(pq, s_red);
        if (threadIdx.x == 0) red_add(A.scal + PS_PBP, pq);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // s_rc aliases the ring
    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
    ok = isfinite(rz0);
    tick(-1);
    if (ok && rz0 > 0.0) {
        double rz_s = rz0;
        for (it = 0; it < A.max_it; ++it) {
            double* row = A.scal + (size_t)it * PS_STRIDE;
            double* nxt = row + PS_STRIDE;
            // ---- A1 / A2 ----
            KT(1);
            if (!(D.debug_skip & 1)) sp_pass_call<SP_FULL>(sh, sp_ring, ring_used, A.segs, A.cta_seg, A.nseg, A.slot_desc, A.nslots_total, A.Es, A.mo_cam, A.p, nullptr,
                                   -1.0, A.q, row + PS_WW);
            ring_used = ring_used || A.cta_seg[blockIdx.x] < A.cta_seg[blockIdx.x + 1];
            KT(2);
            if (boundary && !(D.debug_skip & 2)) {
                sp_pass_call<SP_GATHER>(sh, sp_ring, ring_used, D.bsegs, D.bcta_seg, D.bnseg, D.bslot_desc, D.bnslots_total, A.Es, A.mo_cam, A.p,
                                         D.wbl, 1.0, nullptr, nullptr);
                ring_used = ring_used || D.bcta_seg[blockIdx.x] < D.bcta_seg[blockIdx.x + 1];
            }
            KT(3);
            tick(0);
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            tick(4);
            KT(4);
            double pbp_tot = scal[(size_t)it * PS_STRIDE + PS_PBP], ww_tot = scal[(size_t)it * PS_STRIDE + PS_WW];
            if (boundary || SHARDED) {
                // ---- X1 ----
                double sq = pcg_exchange_boundary<SHARDED>(D, seq0 + 1u + (unsigned)it, pbp_tot, ww_tot);
                sq = cta_sum(sq, s_red);
                if (threadIdx.x == 0) D.cta_part[blockIdx.x] = sq;
                tick(5);
                if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
                tick(4);
                if (SHARDED) { pbp_tot = __ldcg(D.wb + D.n_wb); ww_tot = __ldcg(D.wb + D.n_wb + 1); }
                ww_tot += pcg_fixed_sum(D.cta_part, (int)G, s_red);   // fixed order: the same alpha in every CTA of every rank
                // ---- A3 ----
                KT(5);
                if (boundary) {
                    if (!(D.debug_skip & 4)) sp_pass_call<SP_SCATTER>(sh, sp_ring, ring_used, D.bsegs, D.bcta_seg, D.bnseg, D.bslot_desc, D.bnslots_total, A.Es, A.mo_cam,
                                              nullptr, D.wb, -1.0, A.q, nullptr);
                    KT(6);
                    tick(6);
                    if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
                    tick(4);
                }
            }
            // ---- B ----
            KT(7);
            const double alpha = (rz_s != 0.0) ? rz_s / (pbp_tot - ww_tot) : 0.0;
            {
                double rz = pcg_phase_update(false, alpha, coarse, s_acc);
                rz = cta_sum(rz, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_RZ, rz);
                if (boundary) for (int i = blockIdx.x * kSpThreads + threadIdx.x; i < D.n_wb; i += G * kSpThreads) D.wbl[i] = 0.0;
            }
            tick(1);
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            tick(4);
            // ---- X2 + C ----
            KT(8);
            pcg_coarse_exchange_apply<SHARDED>(D, A, seq0 + (++x2_count), scal[(size_t)(it + 1) * PS_STRIDE + PS_RZ], s_rc, nC);
            tick(2);
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            tick(4);
            // ---- D ----
            KT(9);
            const double rz_new = s_rc[nC] + pcg_fixed_dot(s_rc, A.yc, nC, s_red);
            iterations = it + 1;
            rel2 = rz_new / rz0;
            if (!isfinite(rz_new)) { ok = false; break; }
            if (iterations >= A.min_it && rz_new <= A.tol2 * rz0) break;
            const double beta = (rz_s != 0.0) ? rz_new / rz_s : 0.0;
            rz_s = rz_new;
            {
                double pq = pcg_phase_direction(false, beta, coarse);
                pq = cta_sum(pq, s_red);
                if (threadIdx.x == 0) red_add(nxt + PS_PBP, pq);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tick(3);
            if (!grid_barrier(A.barrier, G)) { aborted = true; goto finish; }
            tick(4);
        }
    } else if (ok) {
        rel2 = 0.0;
    }
finish:
    if (aborted) ok = false;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        A.result[SC_RZ0] = rz0;
        A.result[SC_ITERS] = (double)iterations;
        A.result[SC_REL2] = ok ? rel2 : nan("");
        A.result[SC_DONE] = 1.0;
        if (SHARDED) *D.seq = seq0 + (unsigned)A.max_it + 4u;   // past every number this launch can have used, whatever the iteration count
    }
}

}  // namespace vmm
