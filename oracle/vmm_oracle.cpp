// =====================================================================================
// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of the bundle-adjustment hot path of schmidtniko/visual_marker_mapping:
//   TagReconstructor::doBundleAdjustment              src/TagReconstructor.cpp:646-738
//   TagReconstructionCostFunction::operator()         include/visual_marker_mapping/TagReconstructionCostFunction.h:101-159
//   ReconstructedTag::computeLocalMarkerCorners3D     include/visual_marker_mapping/TagReconstructor.h:44-52
//   CameraModel::projectPoint (reprojection stats)    src/CameraModel.cpp:6-26, src/TagReconstructor.cpp:340-428
//
// The arithmetic of that path lives in Ceres Solver (un-vendored, unpinned: any 1.11..2.1,
// CMakeLists.txt:35), which is absent from /root/reference and from this image.  This file
// restates Ceres' published algorithms: forward-mode dual numbers (jet.h), QuaternionRotatePoint
// (rotation.h), QuaternionParameterization (local_parameterization.cc), HuberLoss/CauchyLoss
// (loss_function.cc), Corrector (corrector.cc), the trust-region Levenberg-Marquardt loop
// (trust_region_minimizer.cc, levenberg_marquardt_strategy.cc, >=1.13 ordering of the tests)
// with an exact linear solve (what SPARSE_NORMAL_CHOLESKY / DENSE_QR deliver).
//
// PARITY UNPINNED: the reference holds no tests, golden vectors or fixtures for this path and
// cannot be built here (needs Ceres, Eigen, OpenCV C++, Boost).  The restatement is pinned
// instead against independent statements of the same maths: cv2.projectPoints (values and
// Jacobians), 50-digit mpmath finite differences, and zero-noise synthetic scenes
// (tests/golden/, tests/test_oracle_*.py).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this library.  Nothing under visual_marker_mapping_b200/ links or calls it.
// =====================================================================================
#include <algorithm>
#include <cfloat>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---------------------------------------------------------------------------------------------
// Dual numbers: value + N partial derivatives (the mechanism behind ceres::Jet<double, N>).
// ---------------------------------------------------------------------------------------------
template <int N>
struct Dual {
    double a;
    double v[N];
    Dual() : a(0.0) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
    explicit Dual(double s) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
    Dual(double s, int k) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Dual<N> operator+(const Dual<N>& f, const Dual<N>& g) {
    Dual<N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Dual<N> operator-(const Dual<N>& f, const Dual<N>& g) {
    Dual<N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Dual<N> operator-(const Dual<N>& f) {
    Dual<N> h; h.a = -f.a; for (int i = 0; i < N; ++i) h.v[i] = -f.v[i]; return h; }
template <int N> inline Dual<N> operator*(const Dual<N>& f, const Dual<N>& g) {
    Dual<N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Dual<N> operator/(const Dual<N>& f, const Dual<N>& g) {
    // (f/g)' = (f' - (f/g) g') / g
    Dual<N> h; const double gi = 1.0 / g.a; h.a = f.a * gi;
    for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - h.a * g.v[i]) * gi; return h; }
template <int N> inline Dual<N>& operator+=(Dual<N>& f, const Dual<N>& g) { f = f + g; return f; }
template <int N> inline Dual<N> sqrt(const Dual<N>& f) {
    Dual<N> h; h.a = std::sqrt(f.a); const double d = 1.0 / (2.0 * h.a);
    for (int i = 0; i < N; ++i) h.v[i] = f.v[i] * d; return h; }
inline double sqrt(double x) { return std::sqrt(x); }

// ---------------------------------------------------------------------------------------------
// ceres::QuaternionRotatePoint: normalise q = (w,x,y,z), then rotate (rotation.h).
// The residual is therefore invariant to |q|.
// ---------------------------------------------------------------------------------------------
template <typename T>
inline void unit_quaternion_rotate_point(const T q[4], const T pt[3], T out[3]) {
    const T t2 = q[0] * q[1];
    const T t3 = q[0] * q[2];
    const T t4 = q[0] * q[3];
    const T t5 = -(q[1] * q[1]);
    const T t6 = q[1] * q[2];
    const T t7 = q[1] * q[3];
    const T t8 = -(q[2] * q[2]);
    const T t9 = q[2] * q[3];
    const T t1 = -(q[3] * q[3]);
    const T two(2.0);
    out[0] = two * ((t8 + t1) * pt[0] + (t6 - t4) * pt[1] + (t3 + t7) * pt[2]) + pt[0];
    out[1] = two * ((t4 + t6) * pt[0] + (t5 + t1) * pt[1] + (t9 - t2) * pt[2]) + pt[1];
    out[2] = two * ((t7 - t3) * pt[0] + (t2 + t9) * pt[1] + (t5 + t8) * pt[2]) + pt[2];
}
template <typename T>
inline void quaternion_rotate_point(const T q[4], const T pt[3], T out[3]) {
    const T one(1.0);
    const T scale = one / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const T unit[4] = { scale * q[0], scale * q[1], scale * q[2], scale * q[3] };
    unit_quaternion_rotate_point(unit, pt, out);
}

// intr = fx, fy, cx, cy, k1, k2, p1, p2, k3   (CameraModel.cpp:28-36; dist order CameraUtilities.cpp:45-58)
struct Intrinsics { double fx, fy, cx, cy, k1, k2, p1, p2, k3; };

// ---------------------------------------------------------------------------------------------
// The reprojection functor, TagReconstructionCostFunction.h:101-159, statement by statement.
// Parameter order (camera_xyz, camera_q, tagPose_xyz, tagPose_q) as in TagReconstructor.cpp:722.
// ---------------------------------------------------------------------------------------------
template <typename T>
inline void corner_functor(const Intrinsics& in, const double obs[2], const double corner[3],
                           const T* cam_t, const T* cam_q, const T* tag_t, const T* tag_q, T* res) {
    T p[3] = { T(corner[0]), T(corner[1]), T(corner[2]) };
    T pw[3];
    quaternion_rotate_point(tag_q, p, pw);            // :109
    pw[0] += tag_t[0]; pw[1] += tag_t[1]; pw[2] += tag_t[2];   // :112-114
    T pc[3];
    quaternion_rotate_point(cam_q, pw, pc);           // :118
    pc[0] += cam_t[0]; pc[1] += cam_t[1]; pc[2] += cam_t[2];   // :120-122
    pc[0] = pc[0] / pc[2];                            // :125
    pc[1] = pc[1] / pc[2];                            // :126
    const T r2 = pc[0] * pc[0] + pc[1] * pc[1];       // :129
    const T k1(in.k1), k2(in.k2), p1(in.p1), p2(in.p2), k3(in.k3);
    const T one(1.0), two(2.0);
    const T xp = pc[0], yp = pc[1];
    const T xd = xp * (one + r2 * (k1 + r2 * (k2 + r2 * k3))) + two * p1 * xp * yp + p2 * (r2 + two * xp * xp);  // :141-142
    const T yd = yp * (one + r2 * (k1 + r2 * (k2 + r2 * k3))) + two * p2 * xp * yp + p1 * (r2 + two * yp * yp);  // :143-144
    const T px = T(in.fx) * xd + T(in.cx);            // :151
    const T py = T(in.fy) * yd + T(in.cy);            // :152
    res[0] = px - T(obs[0]);                          // :155
    res[1] = py - T(obs[1]);                          // :156
}

// ceres::QuaternionParameterization::ComputeJacobian (4x3, row-major) — SURVEY A.2
inline void quat_local_jacobian(const double x[4], double J[12]) {
    J[0] = -x[1]; J[1]  = -x[2]; J[2]  = -x[3];
    J[3] =  x[0]; J[4]  =  x[3]; J[5]  = -x[2];
    J[6] = -x[3]; J[7]  =  x[0]; J[8]  =  x[1];
    J[9] =  x[2]; J[10] = -x[1]; J[11] =  x[0];
}
// ceres::QuaternionParameterization::Plus — delta on the left, half-angle convention.
inline void quat_plus(const double x[4], const double d[3], double out[4]) {
    const double n = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (n > 0.0) {
        const double s = std::sin(n) / n;
        const double q[4] = { std::cos(n), s * d[0], s * d[1], s * d[2] };
        out[0] = q[0] * x[0] - q[1] * x[1] - q[2] * x[2] - q[3] * x[3];
        out[1] = q[0] * x[1] + q[1] * x[0] + q[2] * x[3] - q[3] * x[2];
        out[2] = q[0] * x[2] - q[1] * x[3] + q[2] * x[0] + q[3] * x[1];
        out[3] = q[0] * x[3] + q[1] * x[2] - q[2] * x[1] + q[3] * x[0];
    } else {
        for (int i = 0; i < 4; ++i) out[i] = x[i];
    }
}

// Loss functions (loss_function.cc). rho[0]=rho(s), rho[1]=rho'(s), rho[2]=rho''(s). SURVEY A.3
enum { LOSS_NONE = 0, LOSS_HUBER = 1, LOSS_CAUCHY = 2 };
inline void loss_eval(int kind, double a, double s, double rho[3]) {
    if (kind == LOSS_HUBER) {
        const double b = a * a;
        if (s > b) {
            const double r = std::sqrt(s);
            rho[0] = 2.0 * a * r - b;
            rho[1] = std::max(std::numeric_limits<double>::min(), a / r);
            rho[2] = -rho[1] / (2.0 * s);
        } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
    } else if (kind == LOSS_CAUCHY) {
        const double b = a * a, c = 1.0 / b;
        const double sum = 1.0 + s * c;
        const double inv = 1.0 / sum;
        rho[0] = b * std::log(sum);
        rho[1] = std::max(std::numeric_limits<double>::min(), inv);
        rho[2] = -c * (inv * inv);
    } else { rho[0] = s; rho[1] = 1.0; rho[2] = 0.0; }
}

struct Problem {
    Intrinsics in;
    int64_t Nc, Nm, Nto;
    const double *cam_q, *cam_t, *tag_q, *tag_t, *tag_wh;
    int64_t fixed_tag;
    const int32_t *to_cam, *to_tag;
    const double* to_uv;
};

// local marker corners, TagReconstructor.h:44-52: LL, LR, UR, UL
inline void local_corner(double w, double h, int i, double c[3]) {
    static const double sx[4] = { -1, 1, 1, -1 }, sy[4] = { -1, -1, 1, 1 };
    c[0] = sx[i] * w / 2.0; c[1] = sy[i] * h / 2.0; c[2] = 0.0;
}

// One residual block (= one corner): residual(2), global Jacobian 2x14 [t_c q_c t_m q_m] by
// dual numbers, local Jacobian 2x12 [t_c d_c t_m d_m], then the loss Corrector (corrector.cc;
// rho''<=0 for Huber/Cauchy so only the sqrt(rho') scaling branch is reachable — SURVEY A.4).
// Returns the block cost 0.5*rho(s).
inline double eval_corner(const Problem& P, const double* cq, const double* ct, const double* mq, const double* mt,
                          const double wh[2], int ci, const double obs[2], int loss_kind, double loss_a,
                          double r[2], double Jl[24], double Jg[28], bool want_jac) {
    double corner[3];
    local_corner(wh[0], wh[1], ci, corner);
    double cost;
    if (want_jac) {
        typedef Dual<14> D;
        D camt[3], camq[4], tagt[3], tagq[4], res[2];
        for (int i = 0; i < 3; ++i) camt[i] = D(ct[i], i);
        for (int i = 0; i < 4; ++i) camq[i] = D(cq[i], 3 + i);
        for (int i = 0; i < 3; ++i) tagt[i] = D(mt[i], 7 + i);
        for (int i = 0; i < 4; ++i) tagq[i] = D(mq[i], 10 + i);
        corner_functor<D>(P.in, obs, corner, camt, camq, tagt, tagq, res);
        double Jc[12], Jm[12];
        quat_local_jacobian(cq, Jc);
        quat_local_jacobian(mq, Jm);
        for (int row = 0; row < 2; ++row) {
            r[row] = res[row].a;
            const double* g = res[row].v;
            if (Jg) for (int k = 0; k < 14; ++k) Jg[row * 14 + k] = g[k];
            double* L = Jl + row * 12;
            for (int k = 0; k < 3; ++k) L[k] = g[k];
            for (int k = 0; k < 3; ++k) L[3 + k] = g[3] * Jc[k] + g[4] * Jc[3 + k] + g[5] * Jc[6 + k] + g[6] * Jc[9 + k];
            for (int k = 0; k < 3; ++k) L[6 + k] = g[7 + k];
            for (int k = 0; k < 3; ++k) L[9 + k] = g[10] * Jm[k] + g[11] * Jm[3 + k] + g[12] * Jm[6 + k] + g[13] * Jm[9 + k];
        }
    } else {
        corner_functor<double>(P.in, obs, corner, ct, cq, mt, mq, r);
    }
    const double s = r[0] * r[0] + r[1] * r[1];
    if (loss_kind == LOSS_NONE) {
        cost = 0.5 * s;
    } else {
        double rho[3];
        loss_eval(loss_kind, loss_a, s, rho);
        cost = 0.5 * rho[0];
        const double w = std::sqrt(rho[1]);
        // Corrector: s == 0 or rho'' <= 0  ->  residual *= sqrt(rho'), jacobian *= sqrt(rho')
        r[0] *= w; r[1] *= w;
        if (want_jac) for (int k = 0; k < 24; ++k) Jl[k] *= w;
    }
    return cost;
}

// ---------------------------------------------------------------------------------------------
// Dense Cholesky (lower, in place), blocked + OpenMP for the reduced marker system.
// ---------------------------------------------------------------------------------------------
bool dense_cholesky(double* A, int64_t n) {
    const int64_t NB = 64;
    for (int64_t k0 = 0; k0 < n; k0 += NB) {
        const int64_t kb = std::min(NB, n - k0);
        // factor diagonal block
        for (int64_t j = k0; j < k0 + kb; ++j) {
            double d = A[j * n + j];
            for (int64_t p = k0; p < j; ++p) d -= A[j * n + p] * A[j * n + p];
            if (!(d > 0.0)) return false;
            d = std::sqrt(d);
            A[j * n + j] = d;
            for (int64_t i = j + 1; i < k0 + kb; ++i) {
                double s = A[i * n + j];
                for (int64_t p = k0; p < j; ++p) s -= A[i * n + p] * A[j * n + p];
                A[i * n + j] = s / d;
            }
        }
        // panel solve: rows below
#pragma omp parallel for schedule(static)
        for (int64_t i = k0 + kb; i < n; ++i) {
            for (int64_t j = k0; j < k0 + kb; ++j) {
                double s = A[i * n + j];
                for (int64_t p = k0; p < j; ++p) s -= A[i * n + p] * A[j * n + p];
                A[i * n + j] = s / A[j * n + j];
            }
        }
        // trailing update (lower triangle)
#pragma omp parallel for schedule(dynamic, 8)
        for (int64_t i = k0 + kb; i < n; ++i) {
            const double* Li = A + i * n + k0;
            for (int64_t j = k0 + kb; j <= i; ++j) {
                const double* Lj = A + j * n + k0;
                double s = 0.0;
                for (int64_t p = 0; p < kb; ++p) s += Li[p] * Lj[p];
                A[i * n + j] -= s;
            }
        }
    }
    return true;
}
void dense_cholesky_solve(const double* L, int64_t n, double* b) {
    for (int64_t i = 0; i < n; ++i) {
        double s = b[i];
        for (int64_t p = 0; p < i; ++p) s -= L[i * n + p] * b[p];
        b[i] = s / L[i * n + i];
    }
    for (int64_t i = n - 1; i >= 0; --i) {
        double s = b[i];
        for (int64_t p = i + 1; p < n; ++p) s -= L[p * n + i] * b[p];
        b[i] = s / L[i * n + i];
    }
}
// small SPD inverse (6x6) via Cholesky; returns false if not PD
bool spd_inverse6(const double* A, double* Ainv) {
    double L[36];
    for (int i = 0; i < 36; ++i) L[i] = A[i];
    for (int j = 0; j < 6; ++j) {
        double d = L[j * 6 + j];
        for (int p = 0; p < j; ++p) d -= L[j * 6 + p] * L[j * 6 + p];
        if (!(d > 0.0)) return false;
        d = std::sqrt(d);
        L[j * 6 + j] = d;
        for (int i = j + 1; i < 6; ++i) {
            double s = L[i * 6 + j];
            for (int p = 0; p < j; ++p) s -= L[i * 6 + p] * L[j * 6 + p];
            L[i * 6 + j] = s / d;
        }
    }
    for (int c = 0; c < 6; ++c) {
        double e[6] = { 0, 0, 0, 0, 0, 0 };
        e[c] = 1.0;
        dense_cholesky_solve(L, 6, e);
        for (int r = 0; r < 6; ++r) Ainv[r * 6 + c] = e[r];
    }
    return true;
}

// ---------------------------------------------------------------------------------------------
// Sparse exact solve for site-scale problems: block-envelope ("skyline") Cholesky.
//
// What SPARSE_NORMAL_CHOLESKY gives the reference (src/TagReconstructor.cpp:725-738, Ceres default
// linear solver) is an EXACT solve of the damped normal equations under a fill-reducing ordering.
// Here: cameras are eliminated block-wise (exact), the free markers are renumbered by reverse
// Cuthill-McKee on the co-observation graph (two markers are adjacent when one camera sees both),
// and the marker-reduced matrix is stored and factored inside its row envelope — Cholesky fill
// never leaves the envelope, so this is an exact factorisation with no dropped entries.  Blocks of
// NB x NB doubles (16 markers), right-looking, OpenMP over the trailing update.  On a site whose
// cameras see a ~18 m neighbourhood the envelope of C3 (120k columns) holds ~4 GB instead of the
// 115 GB of the dense matrix; a scene where every camera sees every marker degenerates to dense.
// ---------------------------------------------------------------------------------------------
constexpr int kNB = 96;             // block order: 16 marker slots of 6
constexpr int kSlotsPerBlock = kNB / 6;
typedef double v4d __attribute__((vector_size(32), aligned(8)));

// C (NB x NB, row-major) -= A (NB x NB, row-major) * Bt (NB x NB, row-major; Bt[k][j] = B[j][k])
inline void block_gemm_sub(const double* __restrict__ A, const double* __restrict__ Bt, double* __restrict__ Cm) {
    for (int i0 = 0; i0 < kNB; i0 += 4) {
        for (int j0 = 0; j0 < kNB; j0 += 12) {
            v4d c[4][3];
            for (int r = 0; r < 4; ++r)
                for (int q = 0; q < 3; ++q) c[r][q] = *(const v4d*)(Cm + (size_t)(i0 + r) * kNB + j0 + 4 * q);
            for (int k = 0; k < kNB; ++k) {
                const v4d b0 = *(const v4d*)(Bt + (size_t)k * kNB + j0), b1 = *(const v4d*)(Bt + (size_t)k * kNB + j0 + 4),
                          b2 = *(const v4d*)(Bt + (size_t)k * kNB + j0 + 8);
                for (int r = 0; r < 4; ++r) {
                    const double av = A[(size_t)(i0 + r) * kNB + k];
                    const v4d a = { av, av, av, av };
                    c[r][0] -= a * b0; c[r][1] -= a * b1; c[r][2] -= a * b2;
                }
            }
            for (int r = 0; r < 4; ++r)
                for (int q = 0; q < 3; ++q) *(v4d*)(Cm + (size_t)(i0 + r) * kNB + j0 + 4 * q) = c[r][q];
        }
    }
}

struct EnvelopeCholesky {
    int64_t n = 0, nb = 0;          // scalar order (padded to a multiple of kNB), block rows
    std::vector<int64_t> lo;        // first block column of block row I (non-decreasing)
    std::vector<size_t> off;        // block offset of row I: block (I, J) lives at (off[I] + J - lo[I]) * kNB * kNB
    std::vector<double> a;

    // slot_lo[s]: smallest marker slot that shares a camera with slot s (s itself when none)
    void set_structure(int64_t nslots, const std::vector<int64_t>& slot_lo) {
        nb = std::max<int64_t>(1, (nslots + kSlotsPerBlock - 1) / kSlotsPerBlock);
        n = nb * kNB;
        lo.assign(nb, 0);
        for (int64_t I = 0; I < nb; ++I) {
            int64_t m = I;
            for (int64_t s = I * kSlotsPerBlock; s < std::min(nslots, (I + 1) * kSlotsPerBlock); ++s) m = std::min(m, slot_lo[s] / kSlotsPerBlock);
            lo[I] = m;
        }
        for (int64_t I = nb - 2; I >= 0; --I) lo[I] = std::min(lo[I], lo[I + 1]);   // monotone: the rows coupled to a column are contiguous
        off.assign(nb + 1, 0);
        for (int64_t I = 0; I < nb; ++I) off[I + 1] = off[I] + (size_t)(I - lo[I] + 1);
    }
    size_t blocks() const { return off[nb]; }
    void clear(int64_t nslots) {
        a.assign(blocks() * (size_t)kNB * kNB, 0.0);
        // padding rows beyond the last marker slot: identity, so that the padded matrix stays positive definite
        for (int64_t i = 6 * nslots; i < n; ++i) *entry(i, i) = 1.0;
    }
    double* block(int64_t I, int64_t J) { return a.data() + (off[I] + (size_t)(J - lo[I])) * (size_t)kNB * kNB; }
    const double* block(int64_t I, int64_t J) const { return a.data() + (off[I] + (size_t)(J - lo[I])) * (size_t)kNB * kNB; }
    double* entry(int64_t i, int64_t j) { return block(i / kNB, j / kNB) + (size_t)(i % kNB) * kNB + (j % kNB); }   // i >= j, inside the envelope
    bool inside(int64_t si, int64_t sj) const { return sj / kSlotsPerBlock >= lo[si / kSlotsPerBlock]; }

    // in-place A = L L^T (lower); false when a pivot is not positive
    bool factor() {
        std::vector<double> panelT;   // transposed copies of the blocks L[I][K] of the current block column
        for (int64_t K = 0; K < nb; ++K) {
            double* D = block(K, K);
            for (int j = 0; j < kNB; ++j) {
                double d = D[(size_t)j * kNB + j];
                for (int p = 0; p < j; ++p) d -= D[(size_t)j * kNB + p] * D[(size_t)j * kNB + p];
                if (!(d > 0.0)) return false;
                d = std::sqrt(d);
                D[(size_t)j * kNB + j] = d;
                for (int i = j + 1; i < kNB; ++i) {
                    double s = D[(size_t)i * kNB + j];
                    for (int p = 0; p < j; ++p) s -= D[(size_t)i * kNB + p] * D[(size_t)j * kNB + p];
                    D[(size_t)i * kNB + j] = s / d;
                }
            }
            int64_t hi = K;   // last block row whose envelope reaches column K
            while (hi + 1 < nb && lo[hi + 1] <= K) ++hi;
            const int64_t rows = hi - K;
            if (rows == 0) continue;
            panelT.resize((size_t)rows * kNB * kNB);
            // L[I][K] = A[I][K] L[K][K]^-T, row by row
#pragma omp parallel for schedule(static)
            for (int64_t q = 0; q < rows * kNB; ++q) {
                const int64_t I = K + 1 + q / kNB;
                const int r = (int)(q % kNB);
                double* x = block(I, K) + (size_t)r * kNB;
                double* xt = panelT.data() + (size_t)(I - K - 1) * kNB * kNB;
                for (int j = 0; j < kNB; ++j) {
                    double s = x[j];
                    const double* Lj = D + (size_t)j * kNB;
                    for (int p = 0; p < j; ++p) s -= x[p] * Lj[p];
                    x[j] = s / Lj[j];
                    xt[(size_t)j * kNB + r] = x[j];
                }
            }
            // trailing update: A[I][J] -= L[I][K] L[J][K]^T for K < J <= I <= hi (all inside the envelope: lo[I] <= K < J)
            const int64_t pairs = rows * (rows + 1) / 2;
#pragma omp parallel for schedule(dynamic, 4)
            for (int64_t t = 0; t < pairs; ++t) {
                // t -> (ri, rj) with rj <= ri, largest rows first
                int64_t ri = (int64_t)((std::sqrt(8.0 * (double)t + 1.0) - 1.0) / 2.0);
                while (ri * (ri + 1) / 2 > t) --ri;
                while ((ri + 1) * (ri + 2) / 2 <= t) ++ri;
                const int64_t rj = t - ri * (ri + 1) / 2;
                const int64_t I = K + 1 + ri, J = K + 1 + rj;
                block_gemm_sub(block(I, K), panelT.data() + (size_t)rj * kNB * kNB, block(I, J));
            }
        }
        return true;
    }
    // b <- A^-1 b (b has n entries, padded)
    void solve(double* b) const {
        for (int64_t I = 0; I < nb; ++I) {          // L y = b
            double* bi = b + I * kNB;
            for (int64_t J = lo[I]; J < I; ++J) {
                const double* L = block(I, J);
                const double* bj = b + J * kNB;
                for (int r = 0; r < kNB; ++r) {
                    double s = 0.0;
                    for (int c = 0; c < kNB; ++c) s += L[(size_t)r * kNB + c] * bj[c];
                    bi[r] -= s;
                }
            }
            const double* D = block(I, I);
            for (int r = 0; r < kNB; ++r) {
                double s = bi[r];
                for (int c = 0; c < r; ++c) s -= D[(size_t)r * kNB + c] * bi[c];
                bi[r] = s / D[(size_t)r * kNB + r];
            }
        }
        for (int64_t I = nb - 1; I >= 0; --I) {     // L^T x = y
            double* bi = b + I * kNB;
            const double* D = block(I, I);
            for (int r = kNB - 1; r >= 0; --r) {
                double s = bi[r];
                for (int c = r + 1; c < kNB; ++c) s -= D[(size_t)c * kNB + r] * bi[c];
                bi[r] = s / D[(size_t)r * kNB + r];
            }
            for (int64_t J = lo[I]; J < I; ++J) {
                const double* L = block(I, J);
                double* bj = b + J * kNB;
                for (int r = 0; r < kNB; ++r) {
                    const double xr = bi[r];
                    for (int c = 0; c < kNB; ++c) bj[c] -= L[(size_t)r * kNB + c] * xr;
                }
            }
        }
    }
};

// Reverse Cuthill-McKee numbering of the free markers on the co-observation graph.
// cam_ptr/cam_slots: per camera the free-marker slots (original numbering 0..nfree) it observes.
// Returns new_of_old[slot].
void rcm_marker_order(int64_t nfree, int64_t Nc, const std::vector<int64_t>& cam_ptr, const std::vector<int64_t>& cam_slots,
                      std::vector<int64_t>& new_of_old) {
    std::vector<int64_t> mk_ptr(nfree + 1, 0), mk_cams(cam_slots.size());
    for (int64_t v : cam_slots) mk_ptr[v + 1]++;
    for (int64_t s = 0; s < nfree; ++s) mk_ptr[s + 1] += mk_ptr[s];
    {
        std::vector<int64_t> fill(mk_ptr.begin(), mk_ptr.end() - 1);
        for (int64_t i = 0; i < Nc; ++i)
            for (int64_t e = cam_ptr[i]; e < cam_ptr[i + 1]; ++e) mk_cams[fill[cam_slots[e]]++] = i;
    }
    std::vector<int64_t> degree(nfree, 0);   // co-observations counted with multiplicity: a cheap stand-in for the vertex degree
    for (int64_t s = 0; s < nfree; ++s)
        for (int64_t e = mk_ptr[s]; e < mk_ptr[s + 1]; ++e) degree[s] += cam_ptr[mk_cams[e] + 1] - cam_ptr[mk_cams[e]];
    std::vector<char> seen_m(nfree, 0), seen_c(Nc, 0);
    std::vector<int64_t> order;
    order.reserve(nfree);
    std::vector<int64_t> nbrs;
    // breadth-first sweep from `start`, appending to `out`; neighbours of a vertex enter by increasing degree
    auto bfs = [&](int64_t start, std::vector<int64_t>& out) {
        const size_t first = out.size();
        seen_m[start] = 1;
        out.push_back(start);
        for (size_t head = first; head < out.size(); ++head) {
            const int64_t u = out[head];
            nbrs.clear();
            for (int64_t e = mk_ptr[u]; e < mk_ptr[u + 1]; ++e) {
                const int64_t i = mk_cams[e];
                if (seen_c[i]) continue;     // all markers of an expanded camera are already queued
                seen_c[i] = 1;
                for (int64_t f = cam_ptr[i]; f < cam_ptr[i + 1]; ++f) {
                    const int64_t v = cam_slots[f];
                    if (!seen_m[v]) { seen_m[v] = 1; nbrs.push_back(v); }
                }
            }
            std::sort(nbrs.begin(), nbrs.end(), [&](int64_t x, int64_t y) { return degree[x] < degree[y] || (degree[x] == degree[y] && x < y); });
            out.insert(out.end(), nbrs.begin(), nbrs.end());
        }
    };
    for (int64_t s0 = 0; s0 < nfree; ++s0) {
        if (seen_m[s0]) continue;
        // pseudo-peripheral start vertex of this component: the last vertex of two successive sweeps
        int64_t start = s0;
        for (int sweep = 0; sweep < 2; ++sweep) {
            std::vector<int64_t> probe;
            bfs(start, probe);
            start = probe.back();
            for (int64_t v : probe) {
                seen_m[v] = 0;
                for (int64_t e = mk_ptr[v]; e < mk_ptr[v + 1]; ++e) seen_c[mk_cams[e]] = 0;
            }
        }
        bfs(start, order);
    }
    new_of_old.assign(nfree, 0);
    for (int64_t p = 0; p < nfree; ++p) new_of_old[order[p]] = nfree - 1 - p;   // reversed
}

struct Linearization {
    // per tag-observation 8x6 camera and marker Jacobians (local, corrected, UNSCALED) and residuals(8)
    std::vector<double> Jc, Jm, r;
    std::vector<double> grad;   // 6Nc + 6Nm, zero for inactive blocks
    double cost;
};

struct State {
    std::vector<double> cam_q, cam_t, tag_q, tag_t;
};

void evaluate_all(const Problem& P, const State& S, int loss_kind, double loss_a, bool want_jac,
                  Linearization* lin, double* cost_out, double* res_out, double* jl_out, double* jg_out) {
    const int64_t Nto = P.Nto;
    double cost = 0.0;
    if (lin && want_jac) {
        lin->Jc.assign((size_t)Nto * 48, 0.0);
        lin->Jm.assign((size_t)Nto * 48, 0.0);
        lin->r.assign((size_t)Nto * 8, 0.0);
    }
#pragma omp parallel for schedule(static) reduction(+ : cost)
    for (int64_t k = 0; k < Nto; ++k) {
        const int64_t ci = P.to_cam[k], mi = P.to_tag[k];
        const double* cq = &S.cam_q[4 * ci];
        const double* ct = &S.cam_t[3 * ci];
        const double* mq = &S.tag_q[4 * mi];
        const double* mt = &S.tag_t[3 * mi];
        for (int c = 0; c < 4; ++c) {
            double r[2], Jl[24], Jg[28];
            const double obs[2] = { P.to_uv[8 * k + 2 * c], P.to_uv[8 * k + 2 * c + 1] };
            cost += eval_corner(P, cq, ct, mq, mt, &P.tag_wh[2 * mi], c, obs, loss_kind, loss_a, r, Jl,
                                jg_out ? Jg : nullptr, want_jac);
            if (res_out) { res_out[8 * k + 2 * c] = r[0]; res_out[8 * k + 2 * c + 1] = r[1]; }
            if (want_jac) {
                if (jl_out) std::memcpy(jl_out + (size_t)(4 * k + c) * 24, Jl, 24 * sizeof(double));
                if (jg_out) std::memcpy(jg_out + (size_t)(4 * k + c) * 28, Jg, 28 * sizeof(double));
                if (lin) {
                    for (int row = 0; row < 2; ++row) {
                        lin->r[8 * k + 2 * c + row] = r[row];
                        for (int j = 0; j < 6; ++j) {
                            lin->Jc[(size_t)k * 48 + (2 * c + row) * 6 + j] = Jl[row * 12 + j];
                            lin->Jm[(size_t)k * 48 + (2 * c + row) * 6 + j] = Jl[row * 12 + 6 + j];
                        }
                    }
                }
            }
        }
    }
    if (cost_out) *cost_out = cost;
    if (lin) lin->cost = cost;
}

// gradient g = J^T r over free blocks (layout: camera i at 6i, marker j at 6Nc+6j)
void compute_gradient(const Problem& P, const std::vector<double>& Jc, const std::vector<double>& Jm,
                      const std::vector<double>& r, std::vector<double>& g) {
    g.assign((size_t)6 * (P.Nc + P.Nm), 0.0);
    for (int64_t k = 0; k < P.Nto; ++k) {
        const int64_t ci = P.to_cam[k], mi = P.to_tag[k];
        for (int row = 0; row < 8; ++row) {
            const double rv = r[8 * k + row];
            for (int j = 0; j < 6; ++j) {
                g[6 * ci + j] += Jc[(size_t)k * 48 + row * 6 + j] * rv;
                if (mi != P.fixed_tag) g[6 * P.Nc + 6 * mi + j] += Jm[(size_t)k * 48 + row * 6 + j] * rv;
            }
        }
    }
}

// The marker-reduced system of one LM step in its sparse form (see EnvelopeCholesky): structure once per solve,
// numbers once per step.  Everything is in Jacobi-scaled coordinates like the rest of vmo_solve.
struct SparseReduced {
    int64_t nfree = 0;
    std::vector<int64_t> slot_new;         // per marker: row block (6 rows) in the reverse Cuthill-McKee order, -1 = not free
    std::vector<int64_t> mk_ptr, mk_obs;   // marker -> its tag observations
    EnvelopeCholesky chol;
    std::vector<double> Ek, Wk;            // per tag observation: E_k = Jc_s^T Jm_s and W_k = (B_i + D^2)^-1 E_k

    void setup(const Problem& P, const std::vector<char>& tag_active, const std::vector<int64_t>& cam_ptr,
               const std::vector<int64_t>& cam_obs) {
        const int64_t Nc = P.Nc, Nm = P.Nm, Nto = P.Nto;
        std::vector<int64_t> slot0(Nm, -1);
        nfree = 0;
        for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) slot0[j] = nfree++;
        std::vector<int64_t> cs_ptr(Nc + 1, 0), cs;
        cs.reserve(Nto);
        for (int64_t i = 0; i < Nc; ++i) {
            for (int64_t e = cam_ptr[i]; e < cam_ptr[i + 1]; ++e) {
                const int64_t s = slot0[P.to_tag[cam_obs[e]]];
                if (s >= 0) cs.push_back(s);
            }
            cs_ptr[i + 1] = (int64_t)cs.size();
        }
        std::vector<int64_t> new_of_old;
        rcm_marker_order(nfree, Nc, cs_ptr, cs, new_of_old);
        slot_new.assign(Nm, -1);
        for (int64_t j = 0; j < Nm; ++j) if (slot0[j] >= 0) slot_new[j] = new_of_old[slot0[j]];
        std::vector<int64_t> slot_lo(nfree);
        for (int64_t s = 0; s < nfree; ++s) slot_lo[s] = s;
        for (int64_t i = 0; i < Nc; ++i) {
            int64_t m = nfree;
            for (int64_t e = cs_ptr[i]; e < cs_ptr[i + 1]; ++e) m = std::min(m, new_of_old[cs[e]]);
            for (int64_t e = cs_ptr[i]; e < cs_ptr[i + 1]; ++e) { int64_t& l = slot_lo[new_of_old[cs[e]]]; l = std::min(l, m); }
        }
        chol.set_structure(nfree, slot_lo);
        mk_ptr.assign(Nm + 1, 0);
        mk_obs.resize(Nto);
        for (int64_t k = 0; k < Nto; ++k) mk_ptr[P.to_tag[k] + 1]++;
        for (int64_t j = 0; j < Nm; ++j) mk_ptr[j + 1] += mk_ptr[j];
        std::vector<int64_t> fill(mk_ptr.begin(), mk_ptr.end() - 1);
        for (int64_t k = 0; k < Nto; ++k) mk_obs[fill[P.to_tag[k]]++] = k;
    }

    // ym (6 per marker, zeros for markers without columns) = solution of the reduced system
    //   (C + D^2 - E^T (B + D^2)^-1 E) ym = g_m - E^T (B + D^2)^-1 g_c
    bool solve(const Problem& P, const Linearization& lin, const std::vector<double>& scale, const std::vector<double>& Dsq,
               const std::vector<double>& gs, const std::vector<double>& Binv, const std::vector<int64_t>& cam_ptr,
               const std::vector<int64_t>& cam_obs, std::vector<double>& ym) {
        const int64_t Nc = P.Nc, Nm = P.Nm, Nto = P.Nto;
        Ek.assign((size_t)Nto * 36, 0.0);
        Wk.assign((size_t)Nto * 36, 0.0);
#pragma omp parallel for schedule(static)
        for (int64_t k = 0; k < Nto; ++k) {
            const int64_t i = P.to_cam[k], mi = P.to_tag[k];
            if (slot_new[mi] < 0) continue;
            const double* Jc = &lin.Jc[(size_t)k * 48];
            const double* Jm = &lin.Jm[(size_t)k * 48];
            double* E = &Ek[(size_t)k * 36];
            for (int row = 0; row < 8; ++row)
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 6; ++b)
                        E[a * 6 + b] += Jc[row * 6 + a] * scale[6 * i + a] * Jm[row * 6 + b] * scale[6 * Nc + 6 * mi + b];
            const double* Bi = &Binv[(size_t)i * 36];
            double* W = &Wk[(size_t)k * 36];
            for (int a = 0; a < 6; ++a)
                for (int b = 0; b < 6; ++b) {
                    double s = 0.0;
                    for (int c = 0; c < 6; ++c) s += Bi[a * 6 + c] * E[c * 6 + b];
                    W[a * 6 + b] = s;
                }
        }
        chol.clear(nfree);
        std::vector<double> rhs((size_t)chol.n, 0.0);
        // every marker owns its six rows of the matrix: no two threads write the same entry
#pragma omp parallel for schedule(dynamic, 16)
        for (int64_t j = 0; j < Nm; ++j) {
            const int64_t s1 = slot_new[j];
            if (s1 < 0) continue;
            double Cj[36] = { 0 }, rj[6];
            for (int a = 0; a < 6; ++a) rj[a] = gs[6 * Nc + 6 * j + a];
            const double* sc = &scale[6 * Nc + 6 * j];
            for (int64_t e1 = mk_ptr[j]; e1 < mk_ptr[j + 1]; ++e1) {
                const int64_t k1 = mk_obs[e1], i = P.to_cam[k1];
                const double* J = &lin.Jm[(size_t)k1 * 48];
                for (int row = 0; row < 8; ++row)
                    for (int a = 0; a < 6; ++a)
                        for (int b = 0; b <= a; ++b) Cj[a * 6 + b] += J[row * 6 + a] * sc[a] * J[row * 6 + b] * sc[b];
                const double* E1 = &Ek[(size_t)k1 * 36];
                double bg[6];   // (B_i + D^2)^-1 g_c
                for (int c = 0; c < 6; ++c) {
                    bg[c] = 0.0;
                    for (int d = 0; d < 6; ++d) bg[c] += Binv[(size_t)i * 36 + c * 6 + d] * gs[6 * i + d];
                }
                for (int a = 0; a < 6; ++a)
                    for (int c = 0; c < 6; ++c) rj[a] -= E1[c * 6 + a] * bg[c];
                for (int64_t e2 = cam_ptr[i]; e2 < cam_ptr[i + 1]; ++e2) {
                    const int64_t k2 = cam_obs[e2];
                    const int64_t s2 = slot_new[P.to_tag[k2]];
                    if (s2 < 0 || s2 > s1) continue;
                    const double* W2 = &Wk[(size_t)k2 * 36];
                    double* dst = chol.entry(6 * s1, 6 * s2);
                    for (int a = 0; a < 6; ++a)
                        for (int b = 0; b < 6; ++b) {
                            if (s1 == s2 && b > a) continue;
                            double s = 0.0;
                            for (int c = 0; c < 6; ++c) s += E1[c * 6 + a] * W2[c * 6 + b];
                            dst[(size_t)a * kNB + b] -= s;
                        }
                }
            }
            double* dd = chol.entry(6 * s1, 6 * s1);
            for (int a = 0; a < 6; ++a) {
                for (int b = 0; b <= a; ++b) dd[(size_t)a * kNB + b] += Cj[a * 6 + b];
                dd[(size_t)a * kNB + a] += Dsq[6 * Nc + 6 * j + a];
                rhs[6 * s1 + a] = rj[a];
            }
        }
        if (!chol.factor()) return false;
        chol.solve(rhs.data());
        ym.assign((size_t)6 * Nm, 0.0);
        for (int64_t j = 0; j < Nm; ++j)
            if (slot_new[j] >= 0)
                for (int a = 0; a < 6; ++a) ym[6 * j + a] = rhs[6 * slot_new[j] + a];
        return true;
    }
};

}  // namespace

// =====================================================================================
// C interface (ctypes)
// =====================================================================================
extern "C" {

typedef struct {
    int max_num_iterations;
    int loss_kind;
    double loss_scale;
    double function_tolerance;
    double gradient_tolerance;
    double parameter_tolerance;
    double initial_trust_region_radius;
    double max_trust_region_radius;
    double min_trust_region_radius;
    double min_relative_decrease;
    double min_lm_diagonal;
    double max_lm_diagonal;
    int max_num_consecutive_invalid_steps;
    int jacobi_scaling;
    int linear_solver;            // 0 = block-envelope Cholesky under a reverse Cuthill-McKee marker ordering (default),
                                  // 1 = dense Cholesky of the marker-reduced system (the round-1 path, kept as a cross-check)
    double max_solver_time_s;     // Ceres' max_solver_time_in_seconds (default 1e9): checked once per iteration
} vmo_options;

typedef struct {
    int termination_type;   // 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE (ceres::TerminationType order)
    double initial_cost;
    double final_cost;
    int num_successful_steps;
    int num_unsuccessful_steps;
    int num_iterations;     // number of iteration records incl. iteration 0
    double time_total_s;
    double time_evaluate_s;      // residual+Jacobian evaluations
    double time_linear_solver_s;
    int64_t num_jacobian_evals;
    int64_t num_residual_evals;
    char message[256];
} vmo_summary;

void vmo_options_init(vmo_options* o) {
    // Ceres defaults in force at TagReconstructor.cpp:725-735
    o->max_num_iterations = 50;
    o->loss_kind = LOSS_NONE;
    o->loss_scale = 1.0;
    o->function_tolerance = 1e-6;
    o->gradient_tolerance = 1e-10;
    o->parameter_tolerance = 1e-8;
    o->initial_trust_region_radius = 1e4;
    o->max_trust_region_radius = 1e16;
    o->min_trust_region_radius = 1e-32;
    o->min_relative_decrease = 1e-3;
    o->min_lm_diagonal = 1e-6;
    o->max_lm_diagonal = 1e32;
    o->max_num_consecutive_invalid_steps = 5;
    o->jacobi_scaling = 1;
    o->linear_solver = 0;
    o->max_solver_time_s = 1e9;
}

int vmo_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void vmo_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

static Problem make_problem(const double* intr, int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm,
                            const double* tag_q, const double* tag_t, const double* tag_wh, int64_t fixed_tag,
                            int64_t Nto, const int32_t* to_cam, const int32_t* to_tag, const double* to_uv) {
    Problem P;
    P.in = Intrinsics{ intr[0], intr[1], intr[2], intr[3], intr[4], intr[5], intr[6], intr[7], intr[8] };
    P.Nc = Nc; P.Nm = Nm; P.Nto = Nto;
    P.cam_q = cam_q; P.cam_t = cam_t; P.tag_q = tag_q; P.tag_t = tag_t; P.tag_wh = tag_wh;
    P.fixed_tag = fixed_tag; P.to_cam = to_cam; P.to_tag = to_tag; P.to_uv = to_uv;
    return P;
}

// Evaluate cost (= sum 0.5*rho), corrected residuals [2*4Nto], local Jacobians [4Nto][2][12]
// (columns t_c, delta_c, t_m, delta_m), global Jacobians [4Nto][2][14] (t_c q_c t_m q_m; uncorrected),
// gradient [6Nc+6Nm] (fixed tag: zeros).  Any output may be NULL.
int vmo_evaluate(const double* intr, int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm,
                 const double* tag_q, const double* tag_t, const double* tag_wh, int64_t fixed_tag, int64_t Nto,
                 const int32_t* to_cam, const int32_t* to_tag, const double* to_uv, int loss_kind, double loss_scale,
                 double* cost, double* residuals, double* jac_local, double* jac_global, double* gradient) {
    Problem P = make_problem(intr, Nc, cam_q, cam_t, Nm, tag_q, tag_t, tag_wh, fixed_tag, Nto, to_cam, to_tag, to_uv);
    State S;
    S.cam_q.assign(cam_q, cam_q + 4 * Nc); S.cam_t.assign(cam_t, cam_t + 3 * Nc);
    S.tag_q.assign(tag_q, tag_q + 4 * Nm); S.tag_t.assign(tag_t, tag_t + 3 * Nm);
    const bool want_jac = jac_local || jac_global || gradient;
    Linearization lin;
    evaluate_all(P, S, loss_kind, loss_scale, want_jac, gradient ? &lin : nullptr, cost, residuals, jac_local, jac_global);
    if (gradient) {
        std::vector<double> g;
        compute_gradient(P, lin.Jc, lin.Jm, lin.r, g);
        std::memcpy(gradient, g.data(), g.size() * sizeof(double));
    }
    return 0;
}

// Plus on every block: x (+) delta, delta laid out like the gradient.
void vmo_plus(int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm, const double* tag_q,
              const double* tag_t, const double* delta, double* cam_q_out, double* cam_t_out, double* tag_q_out,
              double* tag_t_out) {
    for (int64_t i = 0; i < Nc; ++i) {
        for (int k = 0; k < 3; ++k) cam_t_out[3 * i + k] = cam_t[3 * i + k] + delta[6 * i + k];
        quat_plus(cam_q + 4 * i, delta + 6 * i + 3, cam_q_out + 4 * i);
    }
    for (int64_t j = 0; j < Nm; ++j) {
        const double* d = delta + 6 * Nc + 6 * j;
        for (int k = 0; k < 3; ++k) tag_t_out[3 * j + k] = tag_t[3 * j + k] + d[k];
        quat_plus(tag_q + 4 * j, d + 3, tag_q_out + 4 * j);
    }
}

// Reprojection statistics, TagReconstructor.cpp:340-428 with CameraModel::projectPoint
// (CameraModel.cpp:6-26).  NOTE the reference quirk reproduced here: projectPoint overwrites pt.x()
// with the distorted x BEFORE computing pt.y(), so the tangential terms of y use the distorted x
// (CameraModel.cpp:20-23).  Rotation uses Eigen's Quaterniond::toRotationMatrix (no normalisation).
// Outputs: per-tag-observation sum of the 4 corner Euclidean errors.
int vmo_reprojection_error_sums(const double* intr, int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm,
                                const double* tag_q, const double* tag_t, const double* tag_wh, int64_t Nto,
                                const int32_t* to_cam, const int32_t* to_tag, const double* to_uv, double* sums) {
    (void)Nc; (void)Nm;
    const Intrinsics in{ intr[0], intr[1], intr[2], intr[3], intr[4], intr[5], intr[6], intr[7], intr[8] };
    auto rotmat = [](const double* q, double R[9]) {
        // Eigen::Quaterniond(w,x,y,z).toRotationMatrix()
        const double w = q[0], x = q[1], y = q[2], z = q[3];
        const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
        const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
                     tyz = tz * y, tzz = tz * z;
        R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
        R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
        R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
    };
#pragma omp parallel for schedule(static)
    for (int64_t k = 0; k < Nto; ++k) {
        const int64_t ci = to_cam[k], mi = to_tag[k];
        double Rm[9], Rc[9];
        rotmat(tag_q + 4 * mi, Rm);
        rotmat(cam_q + 4 * ci, Rc);
        double sum = 0.0;
        for (int c = 0; c < 4; ++c) {
            double p[3];
            local_corner(tag_wh[2 * mi], tag_wh[2 * mi + 1], c, p);
            double pw[3], pc[3];
            for (int i = 0; i < 3; ++i) pw[i] = Rm[3 * i] * p[0] + Rm[3 * i + 1] * p[1] + Rm[3 * i + 2] * p[2] + tag_t[3 * mi + i];
            for (int i = 0; i < 3; ++i) pc[i] = Rc[3 * i] * pw[0] + Rc[3 * i + 1] * pw[1] + Rc[3 * i + 2] * pw[2] + cam_t[3 * ci + i];
            double x = pc[0] / pc[2], y = pc[1] / pc[2];
            const double r2 = x * x + y * y;
            x = x * (1 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3))) + 2 * in.p1 * x * y + in.p2 * (r2 + 2 * x * x);
            y = y * (1 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3))) + 2 * in.p2 * x * y + in.p1 * (r2 + 2 * y * y);
            const double du = in.fx * x + in.cx - to_uv[8 * k + 2 * c];
            const double dv = in.fy * y + in.cy - to_uv[8 * k + 2 * c + 1];
            sum += std::sqrt(du * du + dv * dv);
        }
        sums[k] = sum;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Levenberg-Marquardt with Ceres trust-region semantics and an EXACT linear solve:
// cameras are eliminated block-wise (exact Schur complement), the reduced marker system is
// factorised by dense Cholesky.  Poses are updated in place (TagReconstructor.cpp:665-666,692-693).
// trace (optional, [max_num_iterations+1][6]): cost, gradient_max_norm, step_norm, relative_decrease,
// trust_region_radius, step_is_successful.
// ---------------------------------------------------------------------------------------------
int vmo_solve(const double* intr, int64_t Nc, double* cam_q, double* cam_t, int64_t Nm, double* tag_q, double* tag_t,
              const double* tag_wh, int64_t fixed_tag, int64_t Nto, const int32_t* to_cam, const int32_t* to_tag,
              const double* to_uv, const vmo_options* opt, vmo_summary* sum, double* trace) {
    using clk = std::chrono::steady_clock;
    const auto t_start = clk::now();
    auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
    std::memset(sum, 0, sizeof(*sum));
    Problem P = make_problem(intr, Nc, cam_q, cam_t, Nm, tag_q, tag_t, tag_wh, fixed_tag, Nto, to_cam, to_tag, to_uv);
    State X;
    X.cam_q.assign(cam_q, cam_q + 4 * Nc); X.cam_t.assign(cam_t, cam_t + 3 * Nc);
    X.tag_q.assign(tag_q, tag_q + 4 * Nm); X.tag_t.assign(tag_t, tag_t + 3 * Nm);

    // Active blocks: Ceres' preprocessor drops constant blocks and blocks without residuals.
    std::vector<char> cam_active(Nc, 0), tag_active(Nm, 0);
    for (int64_t k = 0; k < Nto; ++k) { cam_active[to_cam[k]] = 1; tag_active[to_tag[k]] = 1; }
    if (fixed_tag >= 0 && fixed_tag < Nm) tag_active[fixed_tag] = 0;
    std::vector<int64_t> tag_slot(Nm, -1);
    int64_t nfree = 0;
    for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) tag_slot[j] = nfree++;
    const int64_t nred = 6 * nfree;
    const int64_t ntan = 6 * (Nc + Nm);

    // camera -> tag-observation CSR
    std::vector<int64_t> cam_ptr(Nc + 1, 0), cam_obs(Nto);
    for (int64_t k = 0; k < Nto; ++k) cam_ptr[to_cam[k] + 1]++;
    for (int64_t i = 0; i < Nc; ++i) cam_ptr[i + 1] += cam_ptr[i];
    { std::vector<int64_t> fill(cam_ptr.begin(), cam_ptr.end() - 1);
      for (int64_t k = 0; k < Nto; ++k) cam_obs[fill[to_cam[k]]++] = k; }

    SparseReduced sparse;
    if (opt->linear_solver != 1) sparse.setup(P, tag_active, cam_ptr, cam_obs);

    auto x_norm_of = [&](const State& S) {
        double s = 0.0;
        for (int64_t i = 0; i < Nc; ++i) if (cam_active[i]) {
            for (int k = 0; k < 4; ++k) s += S.cam_q[4 * i + k] * S.cam_q[4 * i + k];
            for (int k = 0; k < 3; ++k) s += S.cam_t[3 * i + k] * S.cam_t[3 * i + k]; }
        for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) {
            for (int k = 0; k < 4; ++k) s += S.tag_q[4 * j + k] * S.tag_q[4 * j + k];
            for (int k = 0; k < 3; ++k) s += S.tag_t[3 * j + k] * S.tag_t[3 * j + k]; }
        return std::sqrt(s);
    };
    auto plus_state = [&](const State& S, const std::vector<double>& delta, State& out) {
        out = S;
        for (int64_t i = 0; i < Nc; ++i) if (cam_active[i]) {
            for (int k = 0; k < 3; ++k) out.cam_t[3 * i + k] = S.cam_t[3 * i + k] + delta[6 * i + k];
            quat_plus(&S.cam_q[4 * i], &delta[6 * i + 3], &out.cam_q[4 * i]); }
        for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) {
            const double* d = &delta[6 * Nc + 6 * j];
            for (int k = 0; k < 3; ++k) out.tag_t[3 * j + k] = S.tag_t[3 * j + k] + d[k];
            quat_plus(&S.tag_q[4 * j], d + 3, &out.tag_q[4 * j]); }
    };
    auto state_diff_norm = [&](const State& A, const State& B, bool inf_norm) {
        double s = 0.0;
        auto acc = [&](double d) { if (inf_norm) s = std::max(s, std::fabs(d)); else s += d * d; };
        for (int64_t i = 0; i < Nc; ++i) if (cam_active[i]) {
            for (int k = 0; k < 4; ++k) acc(A.cam_q[4 * i + k] - B.cam_q[4 * i + k]);
            for (int k = 0; k < 3; ++k) acc(A.cam_t[3 * i + k] - B.cam_t[3 * i + k]); }
        for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) {
            for (int k = 0; k < 4; ++k) acc(A.tag_q[4 * j + k] - B.tag_q[4 * j + k]);
            for (int k = 0; k < 3; ++k) acc(A.tag_t[3 * j + k] - B.tag_t[3 * j + k]); }
        return inf_norm ? s : std::sqrt(s);
    };

    Linearization lin;
    std::vector<double> grad, scale(ntan, 1.0), diag(ntan, 0.0), colsq(ntan, 0.0);
    double t_eval = 0.0, t_lin = 0.0;
    int64_t n_jac = 0, n_res = 0;

    auto linearize = [&](const State& S) {
        const auto t0 = clk::now();
        evaluate_all(P, S, opt->loss_kind, opt->loss_scale, true, &lin, nullptr, nullptr, nullptr, nullptr);
        compute_gradient(P, lin.Jc, lin.Jm, lin.r, grad);
        t_eval += secs(t0, clk::now());
        ++n_jac;
    };
    auto squared_column_norms = [&](bool scaled) {
        std::fill(colsq.begin(), colsq.end(), 0.0);
        for (int64_t k = 0; k < Nto; ++k) {
            const int64_t ci = to_cam[k], mi = to_tag[k];
            for (int row = 0; row < 8; ++row)
                for (int j = 0; j < 6; ++j) {
                    const double a = lin.Jc[(size_t)k * 48 + row * 6 + j] * (scaled ? scale[6 * ci + j] : 1.0);
                    colsq[6 * ci + j] += a * a;
                    if (tag_active[mi]) {
                        const double b = lin.Jm[(size_t)k * 48 + row * 6 + j] * (scaled ? scale[6 * Nc + 6 * mi + j] : 1.0);
                        colsq[6 * Nc + 6 * mi + j] += b * b;
                    }
                }
        }
    };
    auto gradient_max_norm = [&](const State& S) {
        std::vector<double> neg(ntan);
        for (int64_t i = 0; i < ntan; ++i) neg[i] = -grad[i];
        State tmp;
        plus_state(S, neg, tmp);
        return state_diff_norm(S, tmp, true);
    };

    // ---------------- iteration 0 ----------------
    linearize(X);
    double x_cost = lin.cost;
    double x_norm = x_norm_of(X);
    sum->initial_cost = x_cost;
    if (opt->jacobi_scaling) {
        squared_column_norms(false);
        for (int64_t i = 0; i < ntan; ++i) scale[i] = 1.0 / (1.0 + std::sqrt(colsq[i]));
    }
    double gmax = gradient_max_norm(X);
    double radius = opt->initial_trust_region_radius;
    double decrease_factor = 2.0;
    bool reuse_diagonal = false;
    int num_consecutive_invalid = 0;
    int iteration = 0;
    int n_success = 0, n_unsuccess = 0;
    int termination = 1;
    const char* msg = "";
    auto record = [&](int it, double cost, double g, double stepn, double rel, double rad, int ok) {
        if (!trace) return;
        double* t = trace + (size_t)it * 6;
        t[0] = cost; t[1] = g; t[2] = stepn; t[3] = rel; t[4] = rad; t[5] = ok;
    };
    record(0, x_cost, gmax, 0.0, 0.0, radius, 1);

    std::vector<double> S_red, rhs_red, Binv((size_t)Nc * 36), step(ntan), delta(ntan), Dsq(ntan);
    State cand;

    while (true) {
        // FinalizeIterationAndCheckIfMinimizerCanContinue
        if (iteration >= opt->max_num_iterations) { termination = 1; msg = "Maximum number of iterations reached."; break; }
        if (secs(t_start, clk::now()) >= opt->max_solver_time_s) { termination = 1; msg = "Maximum solver time reached."; break; }
        if (gmax <= opt->gradient_tolerance) { termination = 0; msg = "Gradient tolerance reached."; break; }
        if (radius < opt->min_trust_region_radius) { termination = 0; msg = "Minimum trust region radius reached."; break; }
        ++iteration;

        // ---- LevenbergMarquardtStrategy::ComputeStep ----
        const auto tl0 = clk::now();
        if (!reuse_diagonal) {
            squared_column_norms(true);
            for (int64_t i = 0; i < ntan; ++i)
                diag[i] = std::min(std::max(colsq[i], opt->min_lm_diagonal), opt->max_lm_diagonal);
        }
        for (int64_t i = 0; i < ntan; ++i) Dsq[i] = diag[i] / radius;   // D^2, D = sqrt(diag/radius)
        bool solve_ok = true;
        // camera blocks (B_i + D^2)^-1 in scaled space
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < Nc; ++i) {
            if (!cam_active[i]) continue;
            double B[36] = { 0 };
            for (int64_t e = cam_ptr[i]; e < cam_ptr[i + 1]; ++e) {
                const double* J = &lin.Jc[(size_t)cam_obs[e] * 48];
                for (int row = 0; row < 8; ++row)
                    for (int a = 0; a < 6; ++a)
                        for (int b = 0; b < 6; ++b) B[a * 6 + b] += J[row * 6 + a] * scale[6 * i + a] * J[row * 6 + b] * scale[6 * i + b];
            }
            for (int a = 0; a < 6; ++a) B[a * 6 + a] += Dsq[6 * i + a];
            if (!spd_inverse6(B, &Binv[(size_t)i * 36])) {
#pragma omp atomic write
                solve_ok = false;
            }
        }
        std::vector<double> gs(ntan);   // scaled gradient J_s^T r
        for (int64_t i = 0; i < ntan; ++i) gs[i] = grad[i] * scale[i];
        std::fill(step.begin(), step.end(), 0.0);
        std::vector<double> ym_m;   // reduced solution, six per marker
        if (solve_ok && opt->linear_solver != 1) {
            solve_ok = sparse.solve(P, lin, scale, Dsq, gs, Binv, cam_ptr, cam_obs, ym_m);
        } else if (solve_ok) {
            S_red.assign((size_t)nred * nred, 0.0);
            rhs_red.assign(nred, 0.0);
            // marker diagonal blocks C_j + D^2 and rhs g_m
            for (int64_t k = 0; k < Nto; ++k) {
                const int64_t mi = to_tag[k];
                if (!tag_active[mi]) continue;
                const int64_t s = tag_slot[mi];
                const double* J = &lin.Jm[(size_t)k * 48];
                const double* sc = &scale[6 * Nc + 6 * mi];
                for (int row = 0; row < 8; ++row)
                    for (int a = 0; a < 6; ++a)
                        for (int b = 0; b <= a; ++b)
                            S_red[(size_t)(6 * s + a) * nred + 6 * s + b] += J[row * 6 + a] * sc[a] * J[row * 6 + b] * sc[b];
            }
            for (int64_t j = 0; j < Nm; ++j) if (tag_active[j]) {
                const int64_t s = tag_slot[j];
                for (int a = 0; a < 6; ++a) {
                    S_red[(size_t)(6 * s + a) * nred + 6 * s + a] += Dsq[6 * Nc + 6 * j + a];
                    rhs_red[6 * s + a] = gs[6 * Nc + 6 * j + a];
                }
            }
            // subtract E^T Binv E and E^T Binv g_c, camera by camera (lower triangle only)
#pragma omp parallel
            {
                std::vector<double> Ek, W;
#pragma omp for schedule(dynamic, 4)
                for (int64_t i = 0; i < Nc; ++i) {
                    if (!cam_active[i]) continue;
                    const int64_t n_i = cam_ptr[i + 1] - cam_ptr[i];
                    Ek.assign((size_t)n_i * 36, 0.0);   // E_k = Jc_s^T Jm_s (6x6, camera rows)
                    W.assign((size_t)n_i * 36, 0.0);    // W_k = Binv E_k
                    for (int64_t e = 0; e < n_i; ++e) {
                        const int64_t k = cam_obs[cam_ptr[i] + e];
                        const int64_t mi = to_tag[k];
                        if (!tag_active[mi]) continue;
                        const double* Jc = &lin.Jc[(size_t)k * 48];
                        const double* Jm = &lin.Jm[(size_t)k * 48];
                        double* E = &Ek[(size_t)e * 36];
                        for (int row = 0; row < 8; ++row)
                            for (int a = 0; a < 6; ++a)
                                for (int b = 0; b < 6; ++b)
                                    E[a * 6 + b] += Jc[row * 6 + a] * scale[6 * i + a] * Jm[row * 6 + b] * scale[6 * Nc + 6 * mi + b];
                        double* Wk = &W[(size_t)e * 36];
                        const double* Bi = &Binv[(size_t)i * 36];
                        for (int a = 0; a < 6; ++a)
                            for (int b = 0; b < 6; ++b) {
                                double s = 0.0;
                                for (int c = 0; c < 6; ++c) s += Bi[a * 6 + c] * E[c * 6 + b];
                                Wk[a * 6 + b] = s;
                            }
                    }
                    for (int64_t e1 = 0; e1 < n_i; ++e1) {
                        const int64_t m1 = to_tag[cam_obs[cam_ptr[i] + e1]];
                        if (!tag_active[m1]) continue;
                        const int64_t s1 = tag_slot[m1];
                        const double* E1 = &Ek[(size_t)e1 * 36];
                        // rhs: -= E1^T Binv g_c
                        for (int a = 0; a < 6; ++a) {
                            double s = 0.0;
                            for (int c = 0; c < 6; ++c) {
                                double bg = 0.0;
                                for (int d = 0; d < 6; ++d) bg += Binv[(size_t)i * 36 + c * 6 + d] * gs[6 * i + d];
                                s += E1[c * 6 + a] * bg;
                            }
#pragma omp atomic
                            rhs_red[6 * s1 + a] -= s;
                        }
                        for (int64_t e2 = 0; e2 < n_i; ++e2) {
                            const int64_t m2 = to_tag[cam_obs[cam_ptr[i] + e2]];
                            if (!tag_active[m2]) continue;
                            const int64_t s2 = tag_slot[m2];
                            if (s2 > s1) continue;
                            const double* W2 = &W[(size_t)e2 * 36];
                            for (int a = 0; a < 6; ++a)
                                for (int b = 0; b < 6; ++b) {
                                    if (s1 == s2 && b > a) continue;
                                    double s = 0.0;
                                    for (int c = 0; c < 6; ++c) s += E1[c * 6 + a] * W2[c * 6 + b];
#pragma omp atomic
                                    S_red[(size_t)(6 * s1 + a) * nred + 6 * s2 + b] -= s;
                                }
                        }
                    }
                }
            }
            std::vector<double> ym(rhs_red);
            if (nred > 0) {
                if (!dense_cholesky(S_red.data(), nred)) solve_ok = false;
                else dense_cholesky_solve(S_red.data(), nred, ym.data());
            }
            ym_m.assign((size_t)6 * Nm, 0.0);
            for (int64_t j = 0; j < Nm; ++j) if (tag_active[j])
                for (int a = 0; a < 6; ++a) ym_m[6 * j + a] = ym[6 * tag_slot[j] + a];
        }
        {
            if (solve_ok) {
                // back-substitute cameras: y_c = Binv (g_c - sum_k E_k y_m)
#pragma omp parallel for schedule(static)
                for (int64_t i = 0; i < Nc; ++i) {
                    if (!cam_active[i]) continue;
                    double t[6];
                    for (int a = 0; a < 6; ++a) t[a] = gs[6 * i + a];
                    for (int64_t e = cam_ptr[i]; e < cam_ptr[i + 1]; ++e) {
                        const int64_t k = cam_obs[e];
                        const int64_t mi = to_tag[k];
                        if (!tag_active[mi]) continue;
                        const double* Jc = &lin.Jc[(size_t)k * 48];
                        const double* Jm = &lin.Jm[(size_t)k * 48];
                        for (int row = 0; row < 8; ++row) {
                            double jm = 0.0;
                            for (int b = 0; b < 6; ++b) jm += Jm[row * 6 + b] * scale[6 * Nc + 6 * mi + b] * ym_m[6 * mi + b];
                            for (int a = 0; a < 6; ++a) t[a] -= Jc[row * 6 + a] * scale[6 * i + a] * jm;
                        }
                    }
                    for (int a = 0; a < 6; ++a) {
                        double s = 0.0;
                        for (int c = 0; c < 6; ++c) s += Binv[(size_t)i * 36 + a * 6 + c] * t[c];
                        step[6 * i + a] = -s;   // step = -(J^T J + D^2)^-1 J^T r
                    }
                }
                for (int64_t j = 0; j < Nm; ++j) if (tag_active[j])
                    for (int a = 0; a < 6; ++a) step[6 * Nc + 6 * j + a] = -ym_m[6 * j + a];
                for (int64_t i = 0; i < ntan; ++i) if (!std::isfinite(step[i])) solve_ok = false;
            }
        }
        reuse_diagonal = false;
        t_lin += secs(tl0, clk::now());

        // model cost change = -(J_s step).(r + J_s step / 2)
        double model_cost_change = 0.0;
        if (solve_ok) {
            double acc = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : acc)
            for (int64_t k = 0; k < Nto; ++k) {
                const int64_t ci = to_cam[k], mi = to_tag[k];
                for (int row = 0; row < 8; ++row) {
                    double m = 0.0;
                    for (int a = 0; a < 6; ++a) {
                        m += lin.Jc[(size_t)k * 48 + row * 6 + a] * scale[6 * ci + a] * step[6 * ci + a];
                        if (tag_active[mi]) m += lin.Jm[(size_t)k * 48 + row * 6 + a] * scale[6 * Nc + 6 * mi + a] * step[6 * Nc + 6 * mi + a];
                    }
                    acc += m * (lin.r[8 * k + row] + 0.5 * m);
                }
            }
            model_cost_change = -acc;
        }
        const bool step_valid = solve_ok && model_cost_change > 0.0;
        if (!step_valid) {
            // HandleInvalidStep
            if (++num_consecutive_invalid >= opt->max_num_consecutive_invalid_steps) {
                termination = 2; msg = "Number of consecutive invalid steps more than max_num_consecutive_invalid_steps.";
                record(iteration, x_cost, gmax, 0.0, 0.0, radius, 0);
                break;
            }
            radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;   // StepIsInvalid -> StepRejected
            ++n_unsuccess;
            record(iteration, x_cost, gmax, 0.0, 0.0, radius, 0);
            continue;
        }
        num_consecutive_invalid = 0;
        for (int64_t i = 0; i < ntan; ++i) delta[i] = step[i] * scale[i];

        // candidate point, cost only
        plus_state(X, delta, cand);
        double cand_cost;
        { const auto t0 = clk::now();
          evaluate_all(P, cand, opt->loss_kind, opt->loss_scale, false, nullptr, &cand_cost, nullptr, nullptr, nullptr);
          t_eval += secs(t0, clk::now()); ++n_res; }
        if (!std::isfinite(cand_cost)) cand_cost = std::numeric_limits<double>::max();

        // ParameterToleranceReached
        const double step_norm = state_diff_norm(X, cand, false);
        if (step_norm <= opt->parameter_tolerance * (x_norm + opt->parameter_tolerance)) {
            termination = 0; msg = "Parameter tolerance reached.";
            record(iteration, x_cost, gmax, step_norm, 0.0, radius, 0);
            break;
        }
        // FunctionToleranceReached
        const double cost_change = x_cost - cand_cost;
        if (std::fabs(cost_change) <= opt->function_tolerance * x_cost) {
            termination = 0; msg = "Function tolerance reached.";
            record(iteration, x_cost, gmax, step_norm, 0.0, radius, 0);
            break;
        }
        const double relative_decrease = cost_change / model_cost_change;
        if (relative_decrease > opt->min_relative_decrease) {
            // HandleSuccessfulStep
            X = cand;
            x_norm = x_norm_of(X);
            linearize(X);
            x_cost = lin.cost;
            gmax = gradient_max_norm(X);
            radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * relative_decrease - 1.0, 3));
            radius = std::min(opt->max_trust_region_radius, radius);
            decrease_factor = 2.0; reuse_diagonal = false;
            ++n_success;
            record(iteration, x_cost, gmax, step_norm, relative_decrease, radius, 1);
        } else {
            radius = radius / decrease_factor; decrease_factor *= 2.0; reuse_diagonal = true;
            ++n_unsuccess;
            record(iteration, cand_cost, gmax, step_norm, relative_decrease, radius, 0);
        }
    }

    std::memcpy(cam_q, X.cam_q.data(), sizeof(double) * 4 * Nc);
    std::memcpy(cam_t, X.cam_t.data(), sizeof(double) * 3 * Nc);
    std::memcpy(tag_q, X.tag_q.data(), sizeof(double) * 4 * Nm);
    std::memcpy(tag_t, X.tag_t.data(), sizeof(double) * 3 * Nm);
    sum->termination_type = termination;
    sum->final_cost = x_cost;
    sum->num_successful_steps = n_success;
    sum->num_unsuccessful_steps = n_unsuccess;
    sum->num_iterations = iteration + 1;
    sum->time_total_s = secs(t_start, clk::now());
    sum->time_evaluate_s = t_eval;
    sum->time_linear_solver_s = t_lin;
    sum->num_jacobian_evals = n_jac;
    sum->num_residual_evals = n_res;
    std::snprintf(sum->message, sizeof(sum->message), "%s", msg);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Check of ONE Levenberg-Marquardt step computed elsewhere (the CUDA engine), matrix-free and without any
// factorisation: with J, r from this file's dual-number evaluation at the given poses, the Jacobi scaling
// of iteration 0 (scale = 1 / (1 + ||column||)) and D^2 = clamp(diag(J_s^T J_s), lo, hi) / radius,
// an exact step satisfies  (J_s^T J_s + D^2) step_s + J_s^T r = 0,  step_s = delta / scale.
// delta: tangent step [6Nc + 6Nm] in unscaled coordinates (what Plus is applied with).
// out[0] = ||(J_s^T J_s + D^2) step_s + g_s||_2,  out[1] = ||g_s||_2,  out[2] = model cost change
// -(J_s step_s).(r + J_s step_s / 2),  out[3] = cost at the given poses.
// Two passes over J; linear in the number of corners, so it runs at C3/C4 scale.
// ---------------------------------------------------------------------------------------------
int vmo_lm_step_residual(const double* intr, int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm,
                         const double* tag_q, const double* tag_t, const double* tag_wh, int64_t fixed_tag, int64_t Nto,
                         const int32_t* to_cam, const int32_t* to_tag, const double* to_uv, int loss_kind, double loss_scale,
                         int jacobi_scaling, double min_lm_diagonal, double max_lm_diagonal, double radius,
                         const double* delta, double* out) {
    Problem P = make_problem(intr, Nc, cam_q, cam_t, Nm, tag_q, tag_t, tag_wh, fixed_tag, Nto, to_cam, to_tag, to_uv);
    State S;
    S.cam_q.assign(cam_q, cam_q + 4 * Nc); S.cam_t.assign(cam_t, cam_t + 3 * Nc);
    S.tag_q.assign(tag_q, tag_q + 4 * Nm); S.tag_t.assign(tag_t, tag_t + 3 * Nm);
    Linearization lin;
    evaluate_all(P, S, loss_kind, loss_scale, true, &lin, nullptr, nullptr, nullptr, nullptr);
    std::vector<double> grad;
    compute_gradient(P, lin.Jc, lin.Jm, lin.r, grad);
    const int64_t ntan = 6 * (Nc + Nm);
    std::vector<char> tag_free(Nm, 0), cam_free(Nc, 0);
    for (int64_t k = 0; k < Nto; ++k) { tag_free[to_tag[k]] = 1; cam_free[to_cam[k]] = 1; }
    if (fixed_tag >= 0 && fixed_tag < Nm) tag_free[fixed_tag] = 0;
    std::vector<double> colsq(ntan, 0.0), scale(ntan, 1.0);
    auto column_norms = [&](bool scaled) {
        std::fill(colsq.begin(), colsq.end(), 0.0);
        for (int64_t k = 0; k < Nto; ++k) {
            const int64_t ci = to_cam[k], mi = to_tag[k];
            for (int row = 0; row < 8; ++row)
                for (int j = 0; j < 6; ++j) {
                    const double a = lin.Jc[(size_t)k * 48 + row * 6 + j] * (scaled ? scale[6 * ci + j] : 1.0);
                    colsq[6 * ci + j] += a * a;
                    if (tag_free[mi]) {
                        const double b = lin.Jm[(size_t)k * 48 + row * 6 + j] * (scaled ? scale[6 * Nc + 6 * mi + j] : 1.0);
                        colsq[6 * Nc + 6 * mi + j] += b * b;
                    }
                }
        }
    };
    if (jacobi_scaling) {
        column_norms(false);
        for (int64_t i = 0; i < ntan; ++i) scale[i] = 1.0 / (1.0 + std::sqrt(colsq[i]));
    }
    column_norms(true);
    std::vector<double> res(ntan, 0.0), step(ntan, 0.0);
    for (int64_t i = 0; i < Nc; ++i) if (cam_free[i]) for (int a = 0; a < 6; ++a) step[6 * i + a] = delta[6 * i + a] / scale[6 * i + a];
    for (int64_t j = 0; j < Nm; ++j) if (tag_free[j]) for (int a = 0; a < 6; ++a) step[6 * (Nc + j) + a] = delta[6 * (Nc + j) + a] / scale[6 * (Nc + j) + a];
    double model = 0.0;
    for (int64_t k = 0; k < Nto; ++k) {
        const int64_t ci = to_cam[k], mi = to_tag[k];
        for (int row = 0; row < 8; ++row) {
            double m = 0.0;
            for (int a = 0; a < 6; ++a) {
                m += lin.Jc[(size_t)k * 48 + row * 6 + a] * scale[6 * ci + a] * step[6 * ci + a];
                if (tag_free[mi]) m += lin.Jm[(size_t)k * 48 + row * 6 + a] * scale[6 * Nc + 6 * mi + a] * step[6 * Nc + 6 * mi + a];
            }
            model += m * (lin.r[8 * k + row] + 0.5 * m);
            for (int a = 0; a < 6; ++a) {
                res[6 * ci + a] += lin.Jc[(size_t)k * 48 + row * 6 + a] * scale[6 * ci + a] * m;
                if (tag_free[mi]) res[6 * Nc + 6 * mi + a] += lin.Jm[(size_t)k * 48 + row * 6 + a] * scale[6 * Nc + 6 * mi + a] * m;
            }
        }
    }
    double rn = 0.0, gn = 0.0;
    for (int64_t i = 0; i < ntan; ++i) {
        const bool free_col = i < 6 * Nc ? cam_free[i / 6] : tag_free[(i - 6 * Nc) / 6];
        if (!free_col) continue;
        const double d2 = std::min(std::max(colsq[i], min_lm_diagonal), max_lm_diagonal) / radius;
        const double gsi = grad[i] * scale[i];
        const double e = res[i] + d2 * step[i] + gsi;
        rn += e * e;
        gn += gsi * gsi;
    }
    out[0] = std::sqrt(rn);
    out[1] = std::sqrt(gn);
    out[2] = -model;
    out[3] = lin.cost;
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Marker-translation covariance for SELECTED markers at site scale (src/TagReconstructor.cpp:744-783):
// the 3x3 translation block of (J^T J)^-1 = the translation block of the inverse of the marker-reduced
// matrix C - E^T B^-1 E (no damping, no scaling), from the same envelope factorisation as the solver: three
// right-hand sides per requested marker.  cov_out: [n_sel][3][3]; markers without columns get zeros.
// Returns 1 when the normal equations are rank deficient.
// ---------------------------------------------------------------------------------------------
int vmo_marker_translation_covariance_sparse(const double* intr, int64_t Nc, const double* cam_q, const double* cam_t, int64_t Nm,
                                             const double* tag_q, const double* tag_t, const double* tag_wh, int64_t fixed_tag,
                                             int64_t Nto, const int32_t* to_cam, const int32_t* to_tag, const double* to_uv,
                                             int loss_kind, double loss_scale, int64_t n_sel, const int64_t* sel, double* cov_out) {
    Problem P = make_problem(intr, Nc, cam_q, cam_t, Nm, tag_q, tag_t, tag_wh, fixed_tag, Nto, to_cam, to_tag, to_uv);
    State S;
    S.cam_q.assign(cam_q, cam_q + 4 * Nc); S.cam_t.assign(cam_t, cam_t + 3 * Nc);
    S.tag_q.assign(tag_q, tag_q + 4 * Nm); S.tag_t.assign(tag_t, tag_t + 3 * Nm);
    Linearization lin;
    evaluate_all(P, S, loss_kind, loss_scale, true, &lin, nullptr, nullptr, nullptr, nullptr);
    std::vector<char> cam_active(Nc, 0), tag_active(Nm, 0);
    for (int64_t k = 0; k < Nto; ++k) { cam_active[to_cam[k]] = 1; tag_active[to_tag[k]] = 1; }
    if (fixed_tag >= 0 && fixed_tag < Nm) tag_active[fixed_tag] = 0;
    std::vector<int64_t> cam_ptr(Nc + 1, 0), cam_obs(Nto);
    for (int64_t k = 0; k < Nto; ++k) cam_ptr[to_cam[k] + 1]++;
    for (int64_t i = 0; i < Nc; ++i) cam_ptr[i + 1] += cam_ptr[i];
    { std::vector<int64_t> fill(cam_ptr.begin(), cam_ptr.end() - 1);
      for (int64_t k = 0; k < Nto; ++k) cam_obs[fill[to_cam[k]]++] = k; }
    const int64_t ntan = 6 * (Nc + Nm);
    std::vector<double> scale(ntan, 1.0), zero(ntan, 0.0), Binv((size_t)Nc * 36, 0.0);
    bool ok = true;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < Nc; ++i) {
        if (!cam_active[i]) continue;
        double B[36] = { 0 };
        for (int64_t e = cam_ptr[i]; e < cam_ptr[i + 1]; ++e) {
            const double* J = &lin.Jc[(size_t)cam_obs[e] * 48];
            for (int row = 0; row < 8; ++row)
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 6; ++b) B[a * 6 + b] += J[row * 6 + a] * J[row * 6 + b];
        }
        if (!spd_inverse6(B, &Binv[(size_t)i * 36])) {
#pragma omp atomic write
            ok = false;
        }
    }
    if (!ok) return 1;
    SparseReduced sp;
    sp.setup(P, tag_active, cam_ptr, cam_obs);
    std::vector<double> ym;
    if (!sp.solve(P, lin, scale, zero, zero, Binv, cam_ptr, cam_obs, ym)) return 1;   // factorises; the zero right-hand side is discarded
    std::vector<double> rhs((size_t)sp.chol.n);
    for (int64_t q = 0; q < n_sel; ++q) {
        double* c = cov_out + 9 * q;
        for (int a = 0; a < 9; ++a) c[a] = 0.0;
        const int64_t s = sp.slot_new[sel[q]];
        if (s < 0) continue;
        for (int col = 0; col < 3; ++col) {
            std::fill(rhs.begin(), rhs.end(), 0.0);
            rhs[6 * s + col] = 1.0;
            sp.chol.solve(rhs.data());
            for (int row = 0; row < 3; ++row) c[row * 3 + col] = rhs[6 * s + row];
        }
    }
    return 0;
}

// Structure report of the sparse path: out[0] = free markers, out[1] = scalar order (padded), out[2] = blocks stored,
// out[3] = bytes of the factor, out[4] = flops of the factorisation (2 x multiply-adds of the trailing updates).
int vmo_sparse_structure(int64_t Nc, int64_t Nm, int64_t fixed_tag, int64_t Nto, const int32_t* to_cam, const int32_t* to_tag, double* out) {
    Problem P{};
    P.Nc = Nc; P.Nm = Nm; P.Nto = Nto; P.to_cam = to_cam; P.to_tag = to_tag; P.fixed_tag = fixed_tag;
    std::vector<char> tag_active(Nm, 0);
    for (int64_t k = 0; k < Nto; ++k) tag_active[to_tag[k]] = 1;
    if (fixed_tag >= 0 && fixed_tag < Nm) tag_active[fixed_tag] = 0;
    std::vector<int64_t> cam_ptr(Nc + 1, 0), cam_obs(Nto);
    for (int64_t k = 0; k < Nto; ++k) cam_ptr[to_cam[k] + 1]++;
    for (int64_t i = 0; i < Nc; ++i) cam_ptr[i + 1] += cam_ptr[i];
    { std::vector<int64_t> fill(cam_ptr.begin(), cam_ptr.end() - 1);
      for (int64_t k = 0; k < Nto; ++k) cam_obs[fill[to_cam[k]]++] = k; }
    SparseReduced sp;
    sp.setup(P, tag_active, cam_ptr, cam_obs);
    double flops = 0.0;
    for (int64_t K = 0; K < sp.chol.nb; ++K) {
        int64_t hi = K;
        while (hi + 1 < sp.chol.nb && sp.chol.lo[hi + 1] <= K) ++hi;
        const double rows = (double)(hi - K);
        flops += rows * (rows + 1) / 2 * 2.0 * kNB * kNB * kNB + rows * (double)kNB * kNB * kNB;
    }
    out[0] = (double)sp.nfree; out[1] = (double)sp.chol.n; out[2] = (double)sp.chol.blocks();
    out[3] = (double)sp.chol.blocks() * kNB * kNB * 8.0; out[4] = flops;
    return 0;
}

}  // extern "C"
