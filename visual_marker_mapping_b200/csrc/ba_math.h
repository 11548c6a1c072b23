// ba_math.h — closed-form per-corner maths of the BA path, shared by every kernel.
//
// What it computes follows TagReconstructionCostFunction::operator()
// (include/visual_marker_mapping/TagReconstructionCostFunction.h:101-159) and the Ceres pieces the
// reference plugs around it (QuaternionParameterization, HuberLoss/CauchyLoss + Corrector,
// src/TagReconstructor.cpp:661,721); how it computes is new: rotation matrices are prepared once per
// pose block instead of per corner, and the 2x12 local Jacobian is analytic instead
// of 14-lane forward autodiff.
//
// All functions are VMM_HD so that the identical code can be compiled by g++ for the CPU-side
// arithmetic tests (tests/test_math_host.py); the product only ever runs them inside CUDA kernels.
#ifndef VMM_BA_MATH_H_
#define VMM_BA_MATH_H_

#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define VMM_HD __host__ __device__ __forceinline__
#else
#define VMM_HD inline
#endif

namespace vmm {

struct Intrinsics {
    double fx, fy, cx, cy, k1, k2, p1, p2, k3;
};

enum { LOSS_NONE = 0, LOSS_HUBER = 1, LOSS_CAUCHY = 2 };

// Prepared camera block: R (row-major, world->camera, from the NORMALISED quaternion as
// ceres::QuaternionRotatePoint does) and t.  12 doubles = 96 B.
constexpr int CAM_PREP = 12;
// Prepared marker block: a = (w/2) R_m[:,0], b = (h/2) R_m[:,1], t_m  (+3 pad) so that the world
// offset of corner c is sx[c]*a + sy[c]*b (local corners have z = 0, TagReconstructor.h:44-52).
constexpr int MARK_PREP = 12;

VMM_HD void quat_to_rot(const double q[4], double R[9]) {
    const double n = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    const double w = q[0] * n, x = q[1] * n, y = q[2] * n, z = q[3] * n;
    R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - w * z);       R[2] = 2.0 * (x * z + w * y);
    R[3] = 2.0 * (x * y + w * z);       R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - w * x);
    R[6] = 2.0 * (x * z - w * y);       R[7] = 2.0 * (y * z + w * x);       R[8] = 1.0 - 2.0 * (x * x + y * y);
}

VMM_HD void prep_camera(const double q[4], const double t[3], double out[CAM_PREP]) {
    quat_to_rot(q, out);
    out[9] = t[0]; out[10] = t[1]; out[11] = t[2];
}

VMM_HD void prep_marker(const double q[4], const double t[3], double w, double h, double out[MARK_PREP]) {
    double R[9];
    quat_to_rot(q, R);
    const double hw = 0.5 * w, hh = 0.5 * h;
    out[0] = hw * R[0]; out[1] = hw * R[3]; out[2] = hw * R[6];
    out[3] = hh * R[1]; out[4] = hh * R[4]; out[5] = hh * R[7];
    out[6] = t[0]; out[7] = t[1]; out[8] = t[2];
    out[9] = 0.0; out[10] = 0.0; out[11] = 0.0;
}

VMM_HD double corner_sx(int c) { return (c == 1 || c == 2) ? 1.0 : -1.0; }
VMM_HD double corner_sy(int c) { return (c >= 2) ? 1.0 : -1.0; }

// rho'(s) of the loss and rho(s) (loss_function.cc semantics)
VMM_HD void loss_eval(int kind, double a, double s, double* rho0, double* rho1) {
    if (kind == LOSS_HUBER) {
        const double b = a * a;
        if (s > b) {
            const double r = sqrt(s);
            *rho0 = 2.0 * a * r - b;
            *rho1 = fmax(DBL_MIN, a / r);
        } else { *rho0 = s; *rho1 = 1.0; }
    } else if (kind == LOSS_CAUCHY) {
        const double b = a * a, c = 1.0 / b;
        const double sum = 1.0 + s * c;
        *rho0 = b * log(sum);
        *rho1 = fmax(DBL_MIN, 1.0 / sum);
    } else { *rho0 = s; *rho1 = 1.0; }
}

// Geometry shared by the four Jacobian blocks of one corner.
struct CornerGeom {
    double r[2];    // residual (uncorrected)
    double JX[6];   // d r / d X_c, 2x3 row-major
    double vc[3];   // R_c X_w
    double vw[3];   // R_m p_local
};

// Residual only (candidate-cost evaluations).
VMM_HD void corner_residual(const Intrinsics& in, const double* cam, const double* mk, int c, double u, double v,
                            double r[2]) {
    const double sx = corner_sx(c), sy = corner_sy(c);
    const double Xw0 = sx * mk[0] + sy * mk[3] + mk[6];
    const double Xw1 = sx * mk[1] + sy * mk[4] + mk[7];
    const double Xw2 = sx * mk[2] + sy * mk[5] + mk[8];
    const double X = cam[0] * Xw0 + cam[1] * Xw1 + cam[2] * Xw2 + cam[9];
    const double Y = cam[3] * Xw0 + cam[4] * Xw1 + cam[5] * Xw2 + cam[10];
    const double Z = cam[6] * Xw0 + cam[7] * Xw1 + cam[8] * Xw2 + cam[11];
    const double iz = 1.0 / Z;
    const double x = X * iz, y = Y * iz;
    const double r2 = x * x + y * y;
    const double cr = 1.0 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3));
    const double xd = x * cr + 2.0 * in.p1 * x * y + in.p2 * (r2 + 2.0 * x * x);
    const double yd = y * cr + 2.0 * in.p2 * x * y + in.p1 * (r2 + 2.0 * y * y);
    r[0] = in.fx * xd + in.cx - u;
    r[1] = in.fy * yd + in.cy - v;
}

// Residual + the pieces of the analytic Jacobian.
VMM_HD void corner_geometry(const Intrinsics& in, const double* cam, const double* mk, int c, double u, double v,
                            CornerGeom& g) {
    const double sx = corner_sx(c), sy = corner_sy(c);
    g.vw[0] = sx * mk[0] + sy * mk[3];
    g.vw[1] = sx * mk[1] + sy * mk[4];
    g.vw[2] = sx * mk[2] + sy * mk[5];
    const double Xw0 = g.vw[0] + mk[6], Xw1 = g.vw[1] + mk[7], Xw2 = g.vw[2] + mk[8];
    g.vc[0] = cam[0] * Xw0 + cam[1] * Xw1 + cam[2] * Xw2;
    g.vc[1] = cam[3] * Xw0 + cam[4] * Xw1 + cam[5] * Xw2;
    g.vc[2] = cam[6] * Xw0 + cam[7] * Xw1 + cam[8] * Xw2;
    const double X = g.vc[0] + cam[9], Y = g.vc[1] + cam[10], Z = g.vc[2] + cam[11];
    const double iz = 1.0 / Z;
    const double x = X * iz, y = Y * iz;
    const double r2 = x * x + y * y;
    const double cr = 1.0 + r2 * (in.k1 + r2 * (in.k2 + r2 * in.k3));
    const double dcr = in.k1 + r2 * (2.0 * in.k2 + 3.0 * in.k3 * r2);
    const double xd = x * cr + 2.0 * in.p1 * x * y + in.p2 * (r2 + 2.0 * x * x);
    const double yd = y * cr + 2.0 * in.p2 * x * y + in.p1 * (r2 + 2.0 * y * y);
    g.r[0] = in.fx * xd + in.cx - u;
    g.r[1] = in.fy * yd + in.cy - v;
    // Delta = d(xd,yd)/d(x,y), symmetric off-diagonal
    const double d00 = cr + 2.0 * x * x * dcr + 2.0 * in.p1 * y + 6.0 * in.p2 * x;
    const double d01 = 2.0 * x * y * dcr + 2.0 * in.p1 * x + 2.0 * in.p2 * y;
    const double d11 = cr + 2.0 * y * y * dcr + 2.0 * in.p2 * x + 6.0 * in.p1 * y;
    // JX = diag(fx,fy) * Delta * [[iz,0,-x iz],[0,iz,-y iz]]
    const double a00 = in.fx * d00 * iz, a01 = in.fx * d01 * iz;
    const double a10 = in.fy * d01 * iz, a11 = in.fy * d11 * iz;
    g.JX[0] = a00; g.JX[1] = a01; g.JX[2] = -(a00 * x + a01 * y);
    g.JX[3] = a10; g.JX[4] = a11; g.JX[5] = -(a10 * x + a11 * y);
}

// One row (length 12) of the local Jacobian from one row j (length 3) of JX:
//   [ j | 2 (vc x j) | j R_c | 2 (vw x (j R_c)) ]
// delta is the QuaternionParameterization tangent (left perturbation, half-angle: R' = (I + 2[d]x) R).
VMM_HD void jacobian_row(const double* j, const double* cam, const double* vc, const double* vw, double* out) {
    out[0] = j[0]; out[1] = j[1]; out[2] = j[2];
    out[3] = 2.0 * (vc[1] * j[2] - vc[2] * j[1]);
    out[4] = 2.0 * (vc[2] * j[0] - vc[0] * j[2]);
    out[5] = 2.0 * (vc[0] * j[1] - vc[1] * j[0]);
    const double m0 = j[0] * cam[0] + j[1] * cam[3] + j[2] * cam[6];
    const double m1 = j[0] * cam[1] + j[1] * cam[4] + j[2] * cam[7];
    const double m2 = j[0] * cam[2] + j[1] * cam[5] + j[2] * cam[8];
    out[6] = m0; out[7] = m1; out[8] = m2;
    out[9]  = 2.0 * (vw[1] * m2 - vw[2] * m1);
    out[10] = 2.0 * (vw[2] * m0 - vw[0] * m2);
    out[11] = 2.0 * (vw[0] * m1 - vw[1] * m0);
}

// Full corner: corrected residual (2), corrected local Jacobian (2x12 row-major), block cost.
// Corrector semantics (corrector.cc): rho'' <= 0 for Huber and Cauchy, so both the
// residual and the Jacobian are scaled by sqrt(rho'); the cost uses rho itself.
VMM_HD double corner_full(const Intrinsics& in, const double* cam, const double* mk, int c, double u, double v,
                          int loss_kind, double loss_a, double r[2], double J[24]) {
    CornerGeom g;
    corner_geometry(in, cam, mk, c, u, v, g);
    jacobian_row(g.JX, cam, g.vc, g.vw, J);
    jacobian_row(g.JX + 3, cam, g.vc, g.vw, J + 12);
    const double s = g.r[0] * g.r[0] + g.r[1] * g.r[1];
    double rho0 = s, rho1 = 1.0;
    if (loss_kind != LOSS_NONE) {
        loss_eval(loss_kind, loss_a, s, &rho0, &rho1);
        const double w = sqrt(rho1);
        r[0] = w * g.r[0]; r[1] = w * g.r[1];
        for (int k = 0; k < 24; ++k) J[k] *= w;
    } else {
        r[0] = g.r[0]; r[1] = g.r[1];
    }
    return 0.5 * rho0;
}

// Camera-side half of corner_full: corrected residual and the six camera columns (t_c, delta_c) of
// both rows, Jc[row*6 + k].  Same arithmetic, the marker columns are simply not formed.
VMM_HD double corner_camera_side(const Intrinsics& in, const double* cam, const double* mk, int c, double u, double v,
                                 int loss_kind, double loss_a, double r[2], double Jc[12]) {
    CornerGeom g;
    corner_geometry(in, cam, mk, c, u, v, g);
#if defined(__CUDACC__)
#pragma unroll
#endif
    for (int row = 0; row < 2; ++row) {
        const double* j = g.JX + 3 * row;
        double* o = Jc + 6 * row;
        o[0] = j[0]; o[1] = j[1]; o[2] = j[2];
        o[3] = 2.0 * (g.vc[1] * j[2] - g.vc[2] * j[1]);
        o[4] = 2.0 * (g.vc[2] * j[0] - g.vc[0] * j[2]);
        o[5] = 2.0 * (g.vc[0] * j[1] - g.vc[1] * j[0]);
    }
    const double s = g.r[0] * g.r[0] + g.r[1] * g.r[1];
    double rho0 = s, rho1 = 1.0;
    if (loss_kind != LOSS_NONE) {
        loss_eval(loss_kind, loss_a, s, &rho0, &rho1);
        const double w = sqrt(rho1);
        r[0] = w * g.r[0]; r[1] = w * g.r[1];
        for (int k = 0; k < 12; ++k) Jc[k] *= w;
    } else {
        r[0] = g.r[0]; r[1] = g.r[1];
    }
    return 0.5 * rho0;
}

VMM_HD double corner_cost(const Intrinsics& in, const double* cam, const double* mk, int c, double u, double v,
                          int loss_kind, double loss_a) {
    double r[2];
    corner_residual(in, cam, mk, c, u, v, r);
    const double s = r[0] * r[0] + r[1] * r[1];
    double rho0 = s, rho1 = 1.0;
    if (loss_kind != LOSS_NONE) loss_eval(loss_kind, loss_a, s, &rho0, &rho1);
    return 0.5 * rho0;
}

// ceres::QuaternionParameterization::Plus: x+ = (cos|d|, sin|d|/|d| d) (x) x   
VMM_HD void quat_plus(const double x[4], const double d[3], double out[4]) {
    const double n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    if (n > 0.0) {
        const double s = sin(n) / n;
        const double q0 = cos(n), q1 = s * d[0], q2 = s * d[1], q3 = s * d[2];
        out[0] = q0 * x[0] - q1 * x[1] - q2 * x[2] - q3 * x[3];
        out[1] = q0 * x[1] + q1 * x[0] + q2 * x[3] - q3 * x[2];
        out[2] = q0 * x[2] - q1 * x[3] + q2 * x[0] + q3 * x[1];
        out[3] = q0 * x[3] + q1 * x[2] - q2 * x[1] + q3 * x[0];
    } else {
        out[0] = x[0]; out[1] = x[1]; out[2] = x[2]; out[3] = x[3];
    }
}

// In-place inverse of a symmetric positive definite 6x6 (row-major, full storage) by Cholesky.
// Returns false (and leaves A unspecified) if a pivot is not positive.
// If Linv_out != nullptr it receives G^-1 (G = Cholesky factor of the input, lower triangular, row-major
// 6x6 with zeros above the diagonal): A^-1 = Linv^T Linv.
VMM_HD bool spd_inverse6(double* A, double* Linv_out = nullptr) {
    double L[21];   // packed lower: L[i*(i+1)/2 + j]
#define VMM_L(i, j) L[(i) * ((i) + 1) / 2 + (j)]
    for (int j = 0; j < 6; ++j) {
        double d = A[j * 6 + j];
        for (int p = 0; p < j; ++p) d -= VMM_L(j, p) * VMM_L(j, p);
        if (!(d > 0.0)) return false;
        d = sqrt(d);
        VMM_L(j, j) = d;
        const double id = 1.0 / d;
        for (int i = j + 1; i < 6; ++i) {
            double s = A[i * 6 + j];
            for (int p = 0; p < j; ++p) s -= VMM_L(i, p) * VMM_L(j, p);
            VMM_L(i, j) = s * id;
        }
    }
    // invert L in place (lower triangular)
    for (int j = 0; j < 6; ++j) {
        VMM_L(j, j) = 1.0 / VMM_L(j, j);
        for (int i = j + 1; i < 6; ++i) {
            double s = 0.0;
            for (int p = j; p < i; ++p) s -= VMM_L(i, p) * VMM_L(p, j);
            VMM_L(i, j) = s / VMM_L(i, i);
        }
    }
    if (Linv_out) {
        for (int i = 0; i < 6; ++i)
            for (int j = 0; j < 6; ++j) Linv_out[i * 6 + j] = (j <= i) ? VMM_L(i, j) : 0.0;
    }
    // A^-1 = L^-T L^-1
    for (int i = 0; i < 6; ++i)
        for (int j = 0; j <= i; ++j) {
            double s = 0.0;
            for (int p = i; p < 6; ++p) s += VMM_L(p, i) * VMM_L(p, j);
            A[i * 6 + j] = s;
            A[j * 6 + i] = s;
        }
#undef VMM_L
    return true;
}

}  // namespace vmm
#endif  // VMM_BA_MATH_H_
