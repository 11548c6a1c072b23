// ba_schur.cuh — the Schur-complement builder and the products with the reduced camera system.
//
// The normal equations of the BA problem (src/TagReconstructor.cpp:699-724) have the block form
//     [ B   E ] [dc]   [gc]        B_i = sum Jc^T Jc (cameras, 6x6)   C_j = sum Jm^T Jm (markers, 6x6)
//     [ E^T C ] [dm] = -[gm]       E_k = Jc^T Jm of tag observation k (4 corners = 8 residual rows)
// The marker blocks are eliminated: S = (B+L_c) - E (C+L_m)^-1 E^T acts on camera space only and is
// never formed.  With (C_j+L_m)^-1 = U_j U_j^T (U_j = G_j^-T from the Cholesky factor G_j) and the
// scaled blocks  Es_k = E_k U_j  the off-diagonal part is a plain Gram product,
//     E (C+L_m)^-1 E^T p = Es (Es^T p),
// so one product with S streams the Es blocks ONCE from HBM, needs no 6x6 solve between its two
// halves and costs one group barrier per marker.  Es is rebuilt (k_scale_e) whenever the damping
// changes, i.e. once per LM iteration against hundreds of products.
//
// Two device orders of the tag observations exist side by side (both built in vmm_ba_set_problem):
//   camera order  — sorted by camera: B_i / g_c / cost by warp-segmented reduction (no hot atomics)
//   marker order  — sorted by marker, CSR offsets mptr[]: one CTA (or consumer group) owns one marker,
//                   so C_j / g_m are plain stores and w_j = sum_k Es_k^T p needs no atomics at all
// E and Es are stored in marker order as byte streams of 8-observation tiles: 18 double2 planes of 8
// entries (2304 B).  A marker's tiles are contiguous and its observation count is padded to a
// multiple of 8 (padding: zeros), so that the observations of any run of markers are ONE contiguous
// byte range that cp.async.bulk can move, and 8 consecutive lanes touch 128 contiguous bytes in every
// plane (shared-memory conflict free, whole sectors in HBM).
#pragma once
#include "ba_kernels.cuh"

namespace vmm {

constexpr int kTileObs = 8;
constexpr int kTileBytes = 18 * kTileObs * 16;           // 2304
constexpr int kSlotTiles = 4;                            // a ring slot: 32 observations = one warp task
constexpr int kSlotBytes = kSlotTiles * kTileBytes;      // 9216

// E plane `plane` of observation i (0-based inside its marker) of a marker whose first tile is tile0
__device__ __forceinline__ double2* e_entry(unsigned char* Et, int tile0, int i, int plane) {
    return reinterpret_cast<double2*>(Et + (size_t)(tile0 + (i >> 3)) * kTileBytes + plane * (kTileObs * 16) + (i & 7) * 16);
}
__device__ __forceinline__ const double2* e_entry(const unsigned char* Et, int tile0, int i, int plane) {
    return reinterpret_cast<const double2*>(Et + (size_t)(tile0 + (i >> 3)) * kTileBytes + plane * (kTileObs * 16) + (i & 7) * 16);
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
                    const int* __restrict__ mtile, unsigned char* __restrict__ Et, double* __restrict__ Ct,
                    double* __restrict__ gm) {
    const int j = blockIdx.x;
    if (!tag_active[j]) return;   // constant origin tag / unobserved tag: no columns (Ct, gm stay zero)
    const int kb = mptr[j], ke = mptr[j + 1], t0 = mtile[j];
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
        for (int i = 0; i < 18; ++i) *e_entry(Et, t0, k - kb, i) = make_double2(E[2 * i], E[2 * i + 1]);
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
// Es_k = E_k U_j with U_j = Linv_j^T (Linv_j: inverse Cholesky factor of C_j + L_m, lower triangular,
// row-major 6x6 with zeros above the diagonal; k_factor_markers).  One CTA per marker, one thread per
// tag observation; reads and writes whole tiles.  Algorithmic bytes per tag observation: 2 x 288.
// ------------------------------------------------------------------------------------------
constexpr int kScaleThreads = 128;
// rhs != nullptr: the right-hand side of the reduced system rides along, rhs[cam(k)] += Es_k zm_j (zm = Linv g_m),
// which saves the separate scatter pass over the freshly written blocks.
__global__ void __launch_bounds__(kScaleThreads)
k_scale_e(const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active, const int* __restrict__ mtile,
          const double* __restrict__ Linv, const unsigned char* __restrict__ Et, unsigned char* __restrict__ Es,
          const int* __restrict__ mo_cam, const double* __restrict__ zm, double* __restrict__ rhs) {
    const int j = blockIdx.x;
    if (!tag_active[j]) return;
    __shared__ double s_l[36];
    __shared__ double s_z[6];
    if (threadIdx.x < 36) s_l[threadIdx.x] = Linv[(size_t)j * 36 + threadIdx.x];
    if (rhs && threadIdx.x >= 64 && threadIdx.x < 70) s_z[threadIdx.x - 64] = zm[(size_t)j * 6 + threadIdx.x - 64];
    __syncthreads();
    const int kb = mptr[j], n = mptr[j + 1] - kb, t0 = mtile[j];
    for (int i = threadIdx.x; i < n; i += kScaleThreads) {
        double E[36];
#pragma unroll
        for (int p = 0; p < 18; ++p) {
            const double2 v = __ldcs(e_entry(Et, t0, i, p));
            E[2 * p] = v.x; E[2 * p + 1] = v.y;
        }
        double* out = rhs ? rhs + (size_t)__ldg(mo_cam + kb + i) * 6 : nullptr;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double o[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
                double s = 0.0;
#pragma unroll
                for (int b = 0; b <= c; ++b) s += E[a * 6 + b] * s_l[c * 6 + b];   // (E U)[a][c] = sum_b E[a][b] Linv[c][b]
                o[c] = s;
            }
#pragma unroll
            for (int h = 0; h < 3; ++h) *e_entry(Es, t0, i, 3 * a + h) = make_double2(o[2 * h], o[2 * h + 1]);
            if (out) {
                double r = 0.0;
#pragma unroll
                for (int c = 0; c < 6; ++c) r += o[c] * s_z[c];
                atomicAdd(out + a, r);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// Products with the scaled E blocks: a persistent, TMA-fed kernel (one CTA per SM).
//
//   SP_FULL    w_j = sum_k Es_k^T v[cam(k)];  out[cam(k)] += sign * Es_k w_j;  dot += w_j.w_j
//              (the whole off-diagonal part of S*p with ONE pass over Es)
//   SP_GATHER  mvec[j] += sum_k Es_k^T v[cam(k)]                        (multi-GPU: all-reduced next)
//   SP_SCATTER out[cam(k)] += sign * Es_k mvec[j]
//
// The unit of work is a *segment*: up to kSegMaxObs tag observations of one marker (a whole marker in
// SP_FULL).  Every CTA owns a contiguous range of segments.  A producer warp moves their tiles through
// a ring of shared-memory slots with cp.async.bulk (one slot = 32 observations = the task of one
// consumer warp; lane r of the producer owns ring position r) under full/empty mbarriers; two
// consumer groups work on alternate segments, each thread keeping the Es block of its observation
// in registers from the gather pass to the scatter pass, so a slot is released as soon as it has been
// consumed and the ring mostly holds data that is still in flight.  A segment longer than the group
// has lanes keeps its tail in the ring slot for both passes.  One named group barrier per segment.
// The camera indices and the camera-space vector entries of the NEXT segment are requested while
// the current one is processed, so no global-memory latency sits between the slot and the FMAs.
// Algorithmic bytes per tag observation: 288 (Es) + 4 (camera index).
// ------------------------------------------------------------------------------------------
enum { SP_FULL = 0, SP_GATHER = 1, SP_SCATTER = 2 };
#ifndef VMM_SP_GROUPS
#define VMM_SP_GROUPS 2
#endif
constexpr int kSpGroups = VMM_SP_GROUPS;
#ifndef VMM_SP_GROUP_WARPS
#define VMM_SP_GROUP_WARPS 7   // 2 x 7 consumer warps + producer = 15 warps: 128 registers per thread (17 warps: 96, spills)
#endif
constexpr int kSpGroupWarps = VMM_SP_GROUP_WARPS;
constexpr int kSpConsumerThreads = kSpGroups * kSpGroupWarps * 32;
constexpr int kSpThreads = kSpConsumerThreads + 32;                  // + producer warp
#ifndef VMM_SP_RING
#define VMM_SP_RING 16
#endif
constexpr int kSpRing = VMM_SP_RING;                                 // slots of kSlotBytes: 184,320 B at 20
constexpr int kSegRegObs = kSpGroupWarps * 32;                       // observations of a segment held in registers
constexpr int kSegMaxObs = 2 * kSegRegObs;                           // the second half stays in its ring slot
constexpr size_t kSpSmemBytes = (size_t)kSpRing * kSlotBytes;

struct __align__(16) SpSeg {
    int marker;   // marker index j
    int slot0;    // sequence number of the segment's first slot in its list
    int nobs;     // observations in this segment (<= kSegMaxObs; <= kSegRegObs outside SP_FULL)
    int k0;       // marker-order index of the segment's first observation (mo_cam[k0 + i])
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" ::"r"(smem_u32(b)), "r"(parity)
        : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* b, uint32_t parity) {   // non-blocking
    uint32_t ok;
    asm volatile(
        "{\n.reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// The same with an L2 eviction policy (createpolicy): the persistent PCG kernel streams the scaled blocks (1.2 GB per pass at
// C3, ten times the L2) evict-first, so that the camera-space vectors and blocks of the phases between two passes survive in L2.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar, uint64_t policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
                 : "memory");
}
// Slot release.  The consumer's reads of the slot (generic proxy, LDS) must be ordered before the producer's next
// cp.async.bulk into it (async proxy): a cross-proxy fence on shared memory, then the arrive.  The arrive itself is
// RELAXED — a releasing arrive makes ptxas track every earlier global reduction of the thread (ATOMG with a
// scoreboard instead of fire-and-forget RED), which serialises the scatter pass; the proxy fence orders exactly what
// the ring needs (FENCE.VIEW.ASYNC.S in SASS, shared memory only) and leaves the reductions in flight.  It is
// issued by every lane of the warp that read the slot, followed by __syncwarp() at the call sites, so that lane 0's
// arrive comes after all 32 lanes' fences.
#ifndef VMM_SP_RELEASE_MODE
#define VMM_SP_RELEASE_MODE 1   // 1: proxy fence + relaxed arrive (product); 0: relaxed arrive alone, 2: releasing arrive (A/B builds, tools/)
#endif
__device__ __forceinline__ void slot_reads_done() {
#if VMM_SP_RELEASE_MODE == 1
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
#endif
}
__device__ __forceinline__ void mbar_arrive_relaxed(uint64_t* b) {
#if VMM_SP_RELEASE_MODE == 2
    asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
#else
    asm volatile("mbarrier.arrive.relaxed.cta.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
#endif
}
__device__ __forceinline__ void red_add(double* addr, double v) {
    asm volatile("red.relaxed.gpu.global.add.f64 [%0], %1;" ::"l"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ void group_barrier(int g) {
    asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "r"(kSpGroupWarps * 32) : "memory");
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
#ifdef VMM_SP_NO_RED   // timing experiment only (tools/): results are wrong
        if (s == 1.2345e300) red_add(out + a, sign * s);
#else
        red_add(out + a, sign * s);
#endif
    }
}
// COHERENT: the vector changes during the kernel's lifetime (persistent PCG).  Plain (L1-cached, coherent) loads: the
// acquire of the grid barrier that precedes the pass invalidates this SM's L1, so the entries are fetched from L2 once per
// pass and then shared by the ~40 observations of a camera; the non-coherent path (ld.global.nc) must not be used.
template <bool COHERENT>
__device__ __forceinline__ void load_x6(const double* __restrict__ v, int cam, double (&x)[6]) {
#ifdef VMM_SP_NO_X    // timing experiment only (tools/): results are wrong
    const double2* x2 = reinterpret_cast<const double2*>(v + (size_t)(cam & 31) * 6);
#else
    const double2* x2 = reinterpret_cast<const double2*>(v + (size_t)cam * 6);
#endif
#pragma unroll
    for (int i = 0; i < 3; ++i) { const double2 t = COHERENT ? x2[i] : __ldg(x2 + i); x[2 * i] = t.x; x[2 * i + 1] = t.y; }
}
// lane's block inside a ring slot (tile = lane / 8, row = lane % 8)
__device__ __forceinline__ void slot_load_e(const unsigned char* slot, int lane, double2 (&E)[18]) {
    const unsigned char* b = slot + (lane >> 3) * kTileBytes + (lane & 7) * 16;
#ifdef VMM_SP_EMU32   // timing experiment only (tools/): half the planes are read and kept (the register and shared-memory footprint of an fp32 stream); results are wrong
#pragma unroll
    for (int i = 0; i < 9; ++i) { E[i] = *reinterpret_cast<const double2*>(b + i * (kTileObs * 16)); E[i + 9] = E[i]; }
    return;
#endif
#pragma unroll
    for (int i = 0; i < 18; ++i) E[i] = *reinterpret_cast<const double2*>(b + i * (kTileObs * 16));
}

// Sum of 6 values over the 32 lanes with 9 exchanges instead of 30: every stage halves the number of
// values a lane carries (it keeps the half selected by one bit of its lane index and hands the other
// half to its partner).  Returns the total of component warp_reduce6_index(lane) — every component
// < 6 ends up, complete, in four lanes.
__device__ __forceinline__ int warp_reduce6_index(int lane) { return ((lane >> 4) & 1) * 4 + ((lane >> 3) & 1) * 2 + ((lane >> 2) & 1); }
__device__ __forceinline__ double warp_reduce6(const double (&y)[6], int lane) {
    const bool h16 = lane & 16, h8 = lane & 8, h4 = lane & 4;
    double a[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const double lo = y[i], hi = (i + 4 < 6) ? y[i + 4] : 0.0;
        a[i] = (h16 ? hi : lo) + shfl_xor_d(h16 ? lo : hi, 16);
    }
    double b[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) b[i] = (h8 ? a[i + 2] : a[i]) + shfl_xor_d(h8 ? a[i] : a[i + 2], 8);
    double c = (h4 ? b[1] : b[0]) + shfl_xor_d(h4 ? b[0] : b[1], 4);
    c += shfl_xor_d(c, 2);
    c += shfl_xor_d(c, 1);
    return c;
}

// Consumer side of the ring: slot sequence number q -> ring position, after its data has landed.
// Two groups (and, without group barriers, the warps of a group) consume out of order, so a warp can
// be a whole lap ahead of the slot it maps to; an mbarrier parity only distinguishes adjacent phases,
// hence the wait on the full barrier starts only once the producer has issued lap q / kSpRing of that
// ring position (then the barrier is in the phase of q or the next one).  No fence: shared memory has
// one copy and the try_wait is control-dependent on the load of s_lap.
__device__ __forceinline__ int ring_acquire(int q, uint64_t* s_full, const volatile int* s_lap, const int* s_lap_base) {
    const int ring = q % kSpRing, lap = s_lap_base[ring] + q / kSpRing;   // laps count from the first pass of the kernel (SpShared::lap_base)
    if ((threadIdx.x & 31) == 0) {
        while (s_lap[ring] <= lap) { }
    }
    __syncwarp();
    mbar_wait(&s_full[ring], (uint32_t)lap & 1);
    return ring;
}

#ifdef VMM_DBG_TRACE   // debug builds (tools/): progress words in mapped host memory, readable after a device fault or during a hang
__device__ unsigned* g_sp_trace = nullptr;   // [CTA][16 warps][4] consumer words, then [CTA][32] producer words at offset 148*64... see SP_TRACE
#define SP_TRACE(k, v) do { if (g_sp_trace && (threadIdx.x & 31) == 0) { \
    *reinterpret_cast<volatile unsigned*>(g_sp_trace + ((size_t)blockIdx.x * 16 + (threadIdx.x >> 5)) * 4 + (k)) = (unsigned)(v); } } while (0)
#define SP_TRACE_P(v) do { if (g_sp_trace) { \
    *reinterpret_cast<volatile unsigned*>(g_sp_trace + (size_t)160 * 64 + (size_t)blockIdx.x * 32 + (threadIdx.x & 31)) = (unsigned)(v); } } while (0)
#else
#define SP_TRACE(k, v) do { } while (0)
#define SP_TRACE_P(v) do { } while (0)
#endif

// Shared state of one streamed pass (static shared memory of the calling kernel; the ring itself is dynamic).
struct SpShared {
    uint64_t full[kSpRing], empty[kSpRing];
    double part[2][kSpGroups][kSpGroupWarps][6];
    volatile int lap[kSpRing];   // laps the producer has issued per ring position since the barriers were initialised (see ring_acquire)
    int lap_base[kSpRing];       // ... at the start of the current pass
    int pass_no;                 // debug builds: passes this CTA has started
};

// One streamed pass over a segment list, executed by ALL kSpThreads threads of the CTA (consumer groups + producer
// warp); returns with every thread (no early exits: the persistent PCG kernel goes on to a grid barrier).  `ring_used`:
// an earlier pass of the same kernel has initialised the ring barriers of this CTA; their phases are carried on.
// COHERENT: see load_x6.
template <int MODE, bool COHERENT>
__device__ __forceinline__ void sp_pass(SpShared& sh, unsigned char* sp_ring, bool ring_used, const SpSeg* __restrict__ segs,
                                        const int* __restrict__ cta_seg, int nseg, const int2* __restrict__ slot_desc, int nslots_total,
                                        const unsigned char* __restrict__ Es, const int* __restrict__ mo_cam,
                                        const double* __restrict__ vin_c, double* __restrict__ mvec, double sign,
                                        double* __restrict__ out_c, double* __restrict__ dot_acc) {
    uint64_t* s_full = sh.full;
    uint64_t* s_empty = sh.empty;
    volatile int* s_lap = sh.lap;
    const int tid = threadIdx.x;
    const int sb = cta_seg[blockIdx.x], se = cta_seg[blockIdx.x + 1];
    if (sb >= se) return;   // uniform for the CTA
    __syncthreads();        // an earlier pass has left the ring
    // The ring barriers are initialised ONCE per kernel and never again: a pass ends with every slot it issued consumed and
    // released, so full[r] and empty[r] have both completed lap[r] phases and the next pass simply goes on counting
    // from there (lap_base).  (Invalidating and re-initialising them per pass — the first version — let a pass start
    // with an empty barrier in the wrong phase once in a while: the producer then overwrote a slot nobody had read.)
    if (tid < kSpRing) {
        if (!ring_used) { mbar_init(&s_full[tid], 1); mbar_init(&s_empty[tid], 1); s_lap[tid] = 0; }
        sh.lap_base[tid] = ring_used ? s_lap[tid] : 0;
        if (!ring_used) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#ifdef VMM_DBG_TRACE
    if (tid == 0) sh.pass_no = ring_used ? sh.pass_no + 1 : 1;
#endif
    __syncthreads();
    const int4* seg4 = reinterpret_cast<const int4*>(segs);
    const int slot_base = __ldg(seg4 + sb).y;

    if (tid >= kSpConsumerThreads) {
        // ---------------- producer warp: lane r owns ring position r and feeds it slot r, r + kSpRing, ... ----------------
        // The lanes poll their empty barriers together (test_wait, non-blocking): a lane parked in a
        // blocking try_wait would hold up the copies of the 19 other ring positions.
        const int lane = tid & 31;
        const int nslots = ((se < nseg) ? __ldg(seg4 + se).y : nslots_total) - slot_base;
        const int2* desc = slot_desc + slot_base;
        unsigned char* dst = sp_ring + (size_t)(lane < kSpRing ? lane : 0) * kSlotBytes;
        int q = (lane < kSpRing) ? lane : nslots;
        int lap = (lane < kSpRing) ? sh.lap_base[lane] : 0;   // laps of this ring position since the barriers were initialised
        int2 sd = (q < nslots) ? __ldg(desc + q) : make_int2(0, 0);
        const uint64_t l2_evict_first = COHERENT ? l2_policy_evict_first() : 0;
#ifdef VMM_DBG_TRACE
        const unsigned tp = ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18);
        SP_TRACE_P(tp | 0x20000u | (unsigned)(nslots & 0xffff));
#endif
        while (__any_sync(0xffffffffu, q < nslots)) {
            const bool ready = (q < nslots) && (lap == 0 || mbar_test(&s_empty[lane], (uint32_t)(lap - 1) & 1));
            if (ready) {
#ifdef VMM_SP_HALF_BYTES   // timing experiment only (tools/): half of every slot is fetched, results are wrong
                const uint32_t bytes = (uint32_t)sd.y * kTileBytes / 2;
#else
                const uint32_t bytes = (uint32_t)sd.y * kTileBytes;
#endif
                mbar_expect_tx(&s_full[lane], bytes);
#ifdef VMM_SP_NO_L2_HINT   // A/B builds (tools/)
                bulk_g2s(dst, Es + (size_t)sd.x * kTileBytes, bytes, &s_full[lane]);
#else
                if (COHERENT) bulk_g2s_hint(dst, Es + (size_t)sd.x * kTileBytes, bytes, &s_full[lane], l2_evict_first);   // persistent PCG kernels
                else bulk_g2s(dst, Es + (size_t)sd.x * kTileBytes, bytes, &s_full[lane]);
#endif
                s_lap[lane] = ++lap;   // after the expect_tx in this thread's shared-memory order
#ifdef VMM_DBG_TRACE
                SP_TRACE_P(tp | (unsigned)(lap & 0xffff));
#endif
                q += kSpRing;
                if (q < nslots) sd = __ldg(desc + q);   // (two laps ahead measured the same: the producer does not wait for descriptors)
            }
        }
#ifdef VMM_DBG_TRACE
        SP_TRACE_P(tp | 0x10000u | (unsigned)(lap & 0xffff));
#endif
        return;
    }

    // ---------------- consumers ----------------
    const int g = tid / (kSpGroupWarps * 32);        // group
    const int gt = tid - g * (kSpGroupWarps * 32);   // thread within group
    const int wg = gt >> 5, lane = tid & 31;
    const int o0 = wg * 32 + lane;                   // observation of the segment held in registers
    const int o1 = o0 + kSegRegObs;                  // observation that stays in its slot (long segments, SP_FULL only)
    const int ridx = warp_reduce6_index(lane);
    double dacc = 0.0;
#ifdef VMM_SP_TIMING   // debug build: phase time stamps of warp 0 of each group into mvec (tools/, not the product)
#define SP_TS(k, dep) do { if (MODE == SP_FULL && mvec && wg == 0 && lane == 0 && it >= 8 && it < 16) \
        mvec[(((size_t)blockIdx.x * 2 + g) * 8 + (it - 8)) * 8 + (k)] = (double)(clock64() + ((dep) == 1.2345e300 ? 1 : 0)); } while (0)
#else
#define SP_TS(k, dep) do { } while (0)
#endif
    // software pipeline: descriptor two segments ahead, camera indices one segment ahead
    const int s0 = sb + g;
    int4 d_cur = (s0 < se) ? __ldg(seg4 + s0) : make_int4(0, 0, 0, 0);
    int4 d_nxt = (s0 + kSpGroups < se) ? __ldg(seg4 + s0 + kSpGroups) : make_int4(0, 0, 0, 0);
    int cam0 = (o0 < d_cur.z) ? __ldg(mo_cam + d_cur.w + o0) : 0;
    int cam1 = (o1 < d_cur.z) ? __ldg(mo_cam + d_cur.w + o1) : 0;
    int it = 0;
    for (int s = s0; s < se; s += kSpGroups, ++it) {
        const SpSeg d{ d_cur.x, d_cur.y, d_cur.z, d_cur.w };
        const int j = d.marker;
        const bool has0 = wg * 32 < d.nobs, has1 = wg * 32 + kSegRegObs < d.nobs;   // warp-uniform
        const bool v0 = o0 < d.nobs, v1 = o1 < d.nobs;
        SP_TS(0, 0.0);
#ifdef VMM_DBG_TRACE
        SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 1u);
        SP_TRACE(1, s - sb);
        SP_TRACE(2, (d.slot0 - slot_base + wg) | (has0 ? 0x80000000u : 0u) | (has1 ? 0x40000000u : 0u));
        SP_TRACE(3, d.nobs);
#endif
        // ---- requests that must not wait for the slot: this segment's vector entries, the next one's indices ----
        double x0[6], z[6];
        const int c0 = cam0, c1 = cam1;
        if (MODE != SP_SCATTER) {
            if (v0) load_x6<COHERENT>(vin_c, c0, x0);
        } else {
#pragma unroll
#ifdef VMM_DBG_SCATTER_LDG
            for (int b = 0; b < 6; ++b) z[b] = __ldg(mvec + (size_t)j * 6 + b);
#else
            for (int b = 0; b < 6; ++b) z[b] = COHERENT ? mvec[(size_t)j * 6 + b] : __ldg(mvec + (size_t)j * 6 + b);
#endif
        }
        d_cur = d_nxt;
        if (s + 2 * kSpGroups < se) d_nxt = __ldg(seg4 + s + 2 * kSpGroups);
        cam0 = (o0 < d_cur.z && s + kSpGroups < se) ? __ldg(mo_cam + d_cur.w + o0) : 0;
        cam1 = (o1 < d_cur.z && s + kSpGroups < se) ? __ldg(mo_cam + d_cur.w + o1) : 0;

        // ---- gather pass ----
        double2 Er[18];
        const unsigned char* slot1 = nullptr;
        int ring0 = 0, ring1 = 0;
        double y[6] = { 0, 0, 0, 0, 0, 0 };
        SP_TS(1, 0.0);
        if (has1) {   // the tail first: its vector entries die before the register-resident block is loaded
            ring1 = ring_acquire(d.slot0 - slot_base + wg + kSpGroupWarps, s_full, s_lap, sh.lap_base);
            slot1 = sp_ring + (size_t)ring1 * kSlotBytes;
            if (MODE != SP_SCATTER && v1) {
                double x1[6];
                load_x6<COHERENT>(vin_c, c1, x1);   // requested here: prefetching it as well costs more in spills than it hides
                double2 Ek[18];
                slot_load_e(slot1, lane, Ek);
                et_times_x(Ek, x1, y);
            }
        }
        if (has0) {
            SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 2u);
            ring0 = ring_acquire(d.slot0 - slot_base + wg, s_full, s_lap, sh.lap_base);
            SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 3u);
            SP_TS(2, 0.0);
            if (v0) {
                slot_load_e(sp_ring + (size_t)ring0 * kSlotBytes, lane, Er);
                if (MODE != SP_SCATTER) et_times_x(Er, x0, y);
            }
        }
        if (MODE != SP_SCATTER) {
            SP_TS(3, y[0] + y[1] + y[2] + y[3] + y[4] + y[5]);
            const double part = warp_reduce6(y, lane);
            SP_TS(4, part);
            if ((lane & 3) == 0 && ridx < 6) sh.part[it & 1][g][wg][ridx] = part;
            if (has0) {   // every lane's reads of the slot are fenced against the async proxy before lane 0 hands it back
                slot_reads_done();
                __syncwarp();
                if (lane == 0) mbar_arrive_relaxed(&s_empty[ring0]);
            }
            group_barrier(g);
            SP_TS(5, 0.0);
            double w = 0.0;
            if (lane < 6) {
#pragma unroll
                for (int q = 0; q < kSpGroupWarps; ++q) w += sh.part[it & 1][g][q][lane];
            }
            if (MODE == SP_GATHER) {
                if (wg == 0 && lane < 6) red_add(mvec + (size_t)j * 6 + lane, w);   // segments of a split marker meet here
                continue;
            }
            if (wg == 0) dacc += w * w;   // lanes 0..5 of the group's first warp
#pragma unroll
            for (int b = 0; b < 6; ++b) z[b] = __shfl_sync(0xffffffffu, w, b);
            SP_TS(6, z[5]);
        }
        // ---- scatter pass ----
        SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 4u);
        if (v0) scatter_e_z(Er, z, sign, out_c + (size_t)c0 * 6);
        SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 5u);
        if (MODE == SP_SCATTER && has0) {
            slot_reads_done();
            __syncwarp();
            if (lane == 0) mbar_arrive_relaxed(&s_empty[ring0]);
        }
        if (has1) {
            if (v1) {
                double2 Ek[18];
                slot_load_e(slot1, lane, Ek);
                scatter_e_z(Ek, z, sign, out_c + (size_t)c1 * 6);
            }
            slot_reads_done();
            __syncwarp();
            if (lane == 0) mbar_arrive_relaxed(&s_empty[ring1]);
        }
        SP_TS(7, 0.0);
        SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 6u);
#ifdef VMM_DBG_SCATTER_BAR
        if (MODE == SP_SCATTER) group_barrier(g);
#endif
    }
    SP_TRACE(0, ((unsigned)sh.pass_no << 20) | ((unsigned)MODE << 18) | 7u);
    if (MODE == SP_FULL && dot_acc && wg == 0) {
        const double dsum = warp_sum(lane < 6 ? dacc : 0.0);
        if (lane == 0) red_add(dot_acc, dsum);
    }
}

template <int MODE>
__global__ void __launch_bounds__(kSpThreads, 1)
k_schur_product(const SpSeg* __restrict__ segs, const int* __restrict__ cta_seg, int nseg, const int2* __restrict__ slot_desc,
                int nslots_total, const unsigned char* __restrict__ Es, const int* __restrict__ mo_cam,
                const double* __restrict__ vin_c, double* __restrict__ mvec, double sign, double* __restrict__ out_c,
                double* __restrict__ dot_acc, const double* __restrict__ done, double* zero0, double* zero1, double* zero2) {
    if (done && *done != 0.0) return;
    extern __shared__ __align__(128) unsigned char sp_ring[];
    __shared__ SpShared sh;
    if (blockIdx.x == 0 && threadIdx.x == 0) {   // recycle the other parity's PCG accumulators (nobody touches them in this launch)
        if (zero0) *zero0 = 0.0;
        if (zero1) *zero1 = 0.0;
        if (zero2) *zero2 = 0.0;
    }
    sp_pass<MODE, false>(sh, sp_ring, false, segs, cta_seg, nseg, slot_desc, nslots_total, Es, mo_cam, vin_c, mvec, sign, out_c, dot_acc);
}

// ------------------------------------------------------------------------------------------
// Block diagonal of the reduced camera system (the SCHUR_JACOBI preconditioner):
//   Sd_i = B_i + diag(lam_c) - sum_k Es_k Es_k^T     (upper triangle; mirrored afterwards)
// ------------------------------------------------------------------------------------------
__global__ void k_schur_diag_init(int n, const double* __restrict__ B, const double* __restrict__ lam,
                                  double* __restrict__ Sd) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 36) return;
    const int blk = i / 36, e = i % 36;
    const int a = e / 6, b = e % 6;
    Sd[i] = B[i] + ((a == b) ? lam[(size_t)blk * 6 + a] : 0.0);
}

constexpr int kDiagThreads = 256;
__global__ void __launch_bounds__(kDiagThreads, 2)
k_schur_diag(const int* __restrict__ mptr, const unsigned char* __restrict__ tag_active, const int* __restrict__ mo_cam,
             const int* __restrict__ mtile, const unsigned char* __restrict__ Es, double* __restrict__ Sd) {
    const int j = blockIdx.x;
    if (!tag_active[j]) return;
    const int kb = mptr[j], ke = mptr[j + 1], t0 = mtile[j];
    for (int k = kb + threadIdx.x; k < ke; k += kDiagThreads) {
        const int cam = __ldg(mo_cam + k);
        double2 E2[18];
#pragma unroll
        for (int i = 0; i < 18; ++i) E2[i] = __ldcs(e_entry(Es, t0, k - kb, i));
        const double* E = reinterpret_cast<const double*>(E2);
        double* o = Sd + (size_t)cam * 36;
#pragma unroll
        for (int a = 0; a < 6; ++a) {
#pragma unroll
            for (int b = a; b < 6; ++b) {
                double s = 0.0;
#pragma unroll
                for (int c = 0; c < 6; ++c) s += E[a * 6 + c] * E[b * 6 + c];   // (Es Es^T)[a][b]
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
