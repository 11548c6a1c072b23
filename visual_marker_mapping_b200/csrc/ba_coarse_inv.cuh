// ba_coarse_inv.cuh — inverse of the dense coarse matrix A_c = P^T S P (order 6 x clusters <= 1536), hand-written.
//
// Round 1 formed it with cusolverDnDpotrf + a recursive cublasDtrsm / cublasDgemm inversion + cublasDsyrk (sm_80 library
// kernels, ~30 launches, 1.8 ms and a blocking synchronisation per refresh).  Here: ONE cooperative kernel that
// inverts the symmetric positive definite matrix in place by the blocked SWEEP operator.  Sweeping the pivot block K
// of a symmetric matrix M,
//     D = M_KK^-1,   M_IJ -= M_IK D M_KJ  (I, J != K),   M_IK <- M_IK D,   M_KK <- -D,
// keeps M symmetric, needs no pivoting for an SPD matrix (every pivot block is a Schur complement of A_c), and after
// all blocks have been swept M = -A_c^-1.  One kind of step, the same amount of work in every step (the whole lower
// triangle is updated: perfectly even tiles for the grid), n^3 multiply-adds in total like potrf + trtri + lauum.
// Per step: every CTA inverts the 64 x 64 pivot block redundantly in shared memory (no barrier for it), the panel
// P_I = M_IK D is formed by the CTAs that own a block row, grid barrier, then the tiles M_IJ -= P_I Q_J^T (Q_J = M_JK
// before the step) are spread over the grid, grid barrier.  fp64 throughout, 4 x 4 register tiles out of shared
// memory.  Only the lower triangle is referenced; the last pass negates and mirrors it.
//
// The matrix is stored row-major with leading dimension ld = order rounded up to a multiple of 64; the padding is the
// identity (set by the caller), its inverse is the identity and never touches the rest.
#pragma once
#include "ba_pcg.cuh"

namespace vmm {

constexpr int kInvNB = 64;
constexpr int kInvThreads = 256;
constexpr int kInvPitch = kInvNB + 1;   // shared-memory row pitch of the pivot block (odd: conflict-free columns)
constexpr size_t kInvSmemBytes = sizeof(double) * ((size_t)kInvNB * kInvPitch + 2 * (size_t)kInvNB * kInvNB + 2 * kInvNB);

// element (i, j), i >= j, of the lower triangle
__device__ __forceinline__ double* inv_at(double* M, int ld, int i, int j) { return M + (size_t)i * ld + j; }

// Q_I = M_IK as a 64 x 64 block in shared memory, TRANSPOSED (k-major): dst[c][r] = M_IK[r][c].
// For I > K the block is stored at (I, K); for I < K it is the transpose of the stored block (K, I).
__device__ __forceinline__ void inv_load_panel_block_T(const double* M, int ld, int I, int K, double* dst) {
    for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
        const int r = e / kInvNB, c = e % kInvNB;   // coalesced over the stored block's rows
        if (I > K) dst[c * kInvNB + r] = __ldcg(M + (size_t)(I * kInvNB + r) * ld + K * kInvNB + c);   // M_IK[r][c]
        else dst[r * kInvNB + c] = __ldcg(M + (size_t)(K * kInvNB + r) * ld + I * kInvNB + c);          // stored (K,I)[r][c] = M_IK[c][r]
    }
}

// Round-2 first version (every entry through shared memory, four cases per entry): kept as the checked reference.
// In-place symmetric sweep of all 64 pivots of the block S (pitch kInvPitch): S <- -S^-1.  Returns false on a
// non-positive pivot.
__device__ __forceinline__ bool inv_sweep_block_v1(double* S, double* col) {
    __shared__ int s_bad;
    if (threadIdx.x == 0) s_bad = 0;
    const int ti = threadIdx.x / 16, tj = threadIdx.x % 16;   // 16 x 16 threads, 4 x 4 entries each (interleaved)
    for (int p = 0; p < kInvNB; ++p) {
        __syncthreads();
        if (threadIdx.x < kInvNB) col[threadIdx.x] = S[threadIdx.x * kInvPitch + p];
        __syncthreads();
        const double piv = col[p];
        if (!(piv > 0.0)) { if (threadIdx.x == 0) s_bad = 1; break; }   // uniform: every thread reads the same value
        const double d = 1.0 / piv;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int i = ti + 16 * a;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int j = tj + 16 * b;
                double v;
                if (i == p && j == p) v = -d;
                else if (i == p) v = d * col[j];
                else if (j == p) v = col[i] * d;
                else v = S[i * kInvPitch + j] - col[i] * d * col[j];
                S[i * kInvPitch + j] = v;
            }
        }
    }
    __syncthreads();
    return s_bad == 0;
}


// In-place symmetric sweep of all 64 pivots of the block S (pitch kInvPitch): S <- -S^-1.  Returns false on a
// non-positive pivot.  Thread (ti, tj) keeps its 16 entries S[ti + 16 a][tj + 16 b] in registers for all 64 pivots; per
// pivot only the pivot column travels through shared memory (double-buffered: one block barrier per pivot), and the
// update is S_ij -= c_i d c_j with row / column p selected to d c_j / c_i d, then S_pp = -d.
__device__ __forceinline__ bool inv_sweep_block_v2(double* S, double* col /* 2 x 64 */) {
    const int ti = threadIdx.x / 16, tj = threadIdx.x % 16;
    double s[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) s[a][b] = S[(ti + 16 * a) * kInvPitch + tj + 16 * b];
    bool good = true;
    for (int p = 0; p < kInvNB; ++p) {
        double* c = col + (p & 1) * kInvNB;
        const int pb = p >> 4, pj = p & 15;
        if (tj == pj) {
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (b == pb) {
#pragma unroll
                    for (int a = 0; a < 4; ++a) c[ti + 16 * a] = s[a][b];
                }
        }
        __syncthreads();
        const double piv = c[p];
        if (!(piv > 0.0)) { good = false; break; }   // uniform: every thread reads the same value
        const double d = 1.0 / piv;
        // S_ij -= c_i d c_j;  row p: d c_j;  column p: c_i d  (selected, not folded into the product: the folded form
        // u_p = piv - 1 cancels catastrophically for the large pivots of a coarse matrix — measured: pivots went negative)
        double ci[4], cjd[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) ci[a] = c[ti + 16 * a];
#pragma unroll
        for (int b = 0; b < 4; ++b) cjd[b] = d * c[tj + 16 * b];
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const bool rowp = ti + 16 * a == p;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const bool colp = tj + 16 * b == p;
                double v = s[a][b] - ci[a] * cjd[b];
                v = rowp ? cjd[b] : v;
                v = colp ? ci[a] * d : v;
                s[a][b] = v;
            }
        }
        if (ti == pj && tj == pj) {
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b)
                    if (a == pb && b == pb) s[a][b] = -d;
        }
    }
    __syncthreads();
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) S[(ti + 16 * a) * kInvPitch + tj + 16 * b] = s[a][b];
    __syncthreads();
    return good;
}

#ifndef VMM_INV_SWEEP_V2
#define VMM_INV_SWEEP_V2 1   // register-resident pivot sweep (0: the round-2 first version, every entry through shared memory)
#endif
#ifndef VMM_INV_LOOKAHEAD
#define VMM_INV_LOOKAHEAD 1  // CTA 0 inverts the next pivot block during the tile phase (0: every CTA at the start of the step)
#endif
__device__ __forceinline__ bool inv_sweep_block(double* S, double* col) {
#if VMM_INV_SWEEP_V2
    return inv_sweep_block_v2(S, col);
#else
    return inv_sweep_block_v1(S, col);
#endif
}

// C (64 x 64, registers: thread (ti, tj) owns rows ti*4.., cols tj*4..) = sum_k At[k][i] * Bt[k][j]
__device__ __forceinline__ void inv_tile_product(const double* At, const double* Bt, double (&acc)[4][4]) {
    const int ti = threadIdx.x / 16, tj = threadIdx.x % 16;
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
#pragma unroll 8
    for (int k = 0; k < kInvNB; ++k) {
        const double2 a01 = *reinterpret_cast<const double2*>(At + k * kInvNB + ti * 4);
        const double2 a23 = *reinterpret_cast<const double2*>(At + k * kInvNB + ti * 4 + 2);
        const double2 b01 = *reinterpret_cast<const double2*>(Bt + k * kInvNB + tj * 4);
        const double2 b23 = *reinterpret_cast<const double2*>(Bt + k * kInvNB + tj * 4 + 2);
        const double av[4] = { a01.x, a01.y, a23.x, a23.y }, bv[4] = { b01.x, b01.y, b23.x, b23.y };
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
    }
}

// M (n x n, row-major, leading dimension ld, n a multiple of 64, lower triangle valid) <- M^-1 (full, symmetric).
// Pt, Qt: scratch panels of (n / 64) blocks of 64 x 64 (k-major: Xt[I][k][r] = X_I[r][k]).  info: set to 1 when a pivot
// is not positive (the matrix is then garbage: the caller drops the coarse level for this iteration).
__global__ void __launch_bounds__(kInvThreads, 1)
k_coarse_sweep_invert(double* __restrict__ M, int n, int ld, double* __restrict__ Pt, double* __restrict__ Qt, int* __restrict__ info,
                      unsigned* __restrict__ barrier) {
    double* Dbuf = Qt + (size_t)n * kInvNB;   // 64 x 64 behind the Q panels: the next pivot block, inverted ahead by CTA 0
    extern __shared__ __align__(16) unsigned char inv_smem[];
    double* S = reinterpret_cast<double*>(inv_smem);       // pivot block / its negative inverse, pitch kInvPitch
    double* At = S + kInvNB * kInvPitch;                    // 64 x 64, k-major
    double* Bt = At + kInvNB * kInvNB;
    double* col = Bt + kInvNB * kInvNB;
    const int nb = n / kInvNB;
    const unsigned G = gridDim.x;
    const int ti = threadIdx.x / 16, tj = threadIdx.x % 16;
    bool good = true;
    const bool look = VMM_INV_LOOKAHEAD && G > 1;   // look-ahead: CTA 0 updates and inverts the NEXT pivot block while the others update the tiles
    for (int K = 0; K < nb && good; ++K) {
        if (K == 0 || !look) {
            // ---- pivot block: every CTA inverts it for itself ----
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
                const int r = e / kInvNB, c = e % kInvNB;
                const int i = K * kInvNB + max(r, c), j = K * kInvNB + min(r, c);
                S[r * kInvPitch + c] = __ldcg(M + (size_t)i * ld + j);
            }
            __syncthreads();
            good = inv_sweep_block(S, col);   // S = -D
            if (!good) { if (blockIdx.x == 0 && threadIdx.x == 0) *info = 1; break; }   // the same verdict in every CTA
        } else {
            // ---- pivot block: inverted by CTA 0 during the previous step ----
            good = __ldcg(info) == 0;
            if (!good) break;
            __syncthreads();
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) S[(e / kInvNB) * kInvPitch + (e % kInvNB)] = __ldcg(Dbuf + e);
            __syncthreads();
        }
        // ---- panel: Q_I (old column block) and P_I = Q_I D for the block rows this CTA owns ----
        for (int I = blockIdx.x; I < nb; I += G) {
            if (I == K) continue;
            __syncthreads();
            inv_load_panel_block_T(M, ld, I, K, At);        // At[k][r] = Q_I[r][k]
            // Bt[k][c] = D[k][c] = -S[k][c]
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) Bt[e] = -S[(e / kInvNB) * kInvPitch + (e % kInvNB)];
            __syncthreads();
            double acc[4][4];
            inv_tile_product(At, Bt, acc);                  // P_I[r][c] = sum_k Q_I[r][k] D[k][c]
            double* q = Qt + (size_t)I * kInvNB * kInvNB;
            double* p = Pt + (size_t)I * kInvNB * kInvNB;
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) q[e] = At[e];
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) p[(size_t)(tj * 4 + b) * kInvNB + ti * 4 + a] = acc[a][b];   // k-major: Pt[c][r]
        }
        if (!grid_barrier(barrier, G)) return;
        // ---- tiles: M_IJ -= P_I Q_J^T over the lower triangle without row / column K; panel and pivot written back ----
        const int m = nb - 1;                               // block indices without K
        const int ntiles = m * (m + 1) / 2;
        // the tile of the next pivot block (indices without K: K + 1 -> K) goes to CTA 0 alone, which then inverts it
        const int t_ahead = (look && K + 1 < nb) ? K * (K + 1) / 2 + K : -1;
        const int n_items = ntiles + nb - (t_ahead >= 0 ? 1 : 0);
        const int first = t_ahead >= 0 ? (int)blockIdx.x - 1 : (int)blockIdx.x, step = t_ahead >= 0 ? (int)G - 1 : (int)G;
        for (int it = (blockIdx.x == 0 && t_ahead >= 0) ? -1 : first; it < n_items; it += step) {
            const int t = it < 0 ? t_ahead : ((t_ahead >= 0 && it >= t_ahead) ? it + 1 : it);
            if (t >= ntiles) {
                // write-back of the panel block I (= t - ntiles) and, for I == K, of the pivot block
                const int I = t - ntiles;
                if (I == K) {
                    for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
                        const int r = e / kInvNB, c = e % kInvNB;
                        if (c <= r) M[(size_t)(K * kInvNB + r) * ld + K * kInvNB + c] = S[r * kInvPitch + c];   // -D
                    }
                } else {
                    const double* p = Pt + (size_t)I * kInvNB * kInvNB;   // Pt[c][r] = P_I[r][c]
                    for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
                        const int r = e / kInvNB, c = e % kInvNB;
                        if (I > K) M[(size_t)(I * kInvNB + r) * ld + K * kInvNB + c] = __ldcg(p + (size_t)c * kInvNB + r);        // M_IK = P_I
                        else M[(size_t)(K * kInvNB + r) * ld + I * kInvNB + c] = __ldcg(p + (size_t)r * kInvNB + c);               // M_KI = P_I^T
                    }
                }
                continue;
            }
            // t -> (ri, rj), rj <= ri, among the m remaining block indices
            int ri = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
            while (ri * (ri + 1) / 2 > t) --ri;
            while ((ri + 1) * (ri + 2) / 2 <= t) ++ri;
            const int rj = t - ri * (ri + 1) / 2;
            const int I = ri + (ri >= K), J = rj + (rj >= K);
            __syncthreads();
            const double* p = Pt + (size_t)I * kInvNB * kInvNB;
            const double* q = Qt + (size_t)J * kInvNB * kInvNB;
            for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) { At[e] = __ldcg(p + e); Bt[e] = __ldcg(q + e); }
            __syncthreads();
            double acc[4][4];
            inv_tile_product(At, Bt, acc);                  // (P_I Q_J^T)[r][c] = sum_k P_I[r][k] Q_J[c][k]
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int r = ti * 4 + a, c = tj * 4 + b;
                    if (I > J || c <= r) {
                        double* dst = M + (size_t)(I * kInvNB + r) * ld + J * kInvNB + c;
                        *dst = __ldcg(dst) - acc[a][b];
                    }
                }
            if (it < 0) {
                // CTA 0: the freshly updated block (K+1, K+1) is the next pivot block — invert it now, off the critical path
                __threadfence();
                __syncthreads();
                for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) {
                    const int r = e / kInvNB, c = e % kInvNB;
                    const int i = (K + 1) * kInvNB + max(r, c), j = (K + 1) * kInvNB + min(r, c);
                    S[r * kInvPitch + c] = __ldcg(M + (size_t)i * ld + j);
                }
                __syncthreads();
                const bool ok_next = inv_sweep_block(S, col);
                if (!ok_next && threadIdx.x == 0) *info = 1;
                for (int e = threadIdx.x; e < kInvNB * kInvNB; e += kInvThreads) Dbuf[e] = S[(e / kInvNB) * kInvPitch + (e % kInvNB)];
                break;   // CTA 0 takes no other item of this step
            }
        }
        if (!grid_barrier(barrier, G)) return;
    }
    if (!good) return;
    // ---- M = -A^-1 in the lower triangle: negate and mirror ----
    for (size_t e = (size_t)blockIdx.x * kInvThreads + threadIdx.x; e < (size_t)n * n; e += (size_t)G * kInvThreads) {
        const int r = (int)(e / n), c = (int)(e % n);
        if (c <= r) {
            const double v = -__ldcg(M + (size_t)r * ld + c);
            M[(size_t)r * ld + c] = v;
            M[(size_t)c * ld + r] = v;
        }
    }
}

}  // namespace vmm
