// PnP.cpp — see PnP.h.  Small dense linear algebra is written out here (Jacobi eigen-solver, 6x6 Cholesky):
// the problems have 6 unknowns and at most a few hundred correspondences.
#include "visual_marker_mapping/PnP.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

namespace visual_marker_mapping
{
namespace
{
// Cyclic Jacobi for a symmetric n x n matrix (row-major, n <= 12).  On return d holds the eigenvalues and the
// COLUMNS of V the eigenvectors.
void jacobiEigen(int n, double* A, double* V, double* d)
{
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) V[i * n + j] = (i == j) ? 1.0 : 0.0;
    for (int sweep = 0; sweep < 100; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) (i == j ? diag : off) += A[i * n + j] * A[i * n + j];
        if (off <= 1e-30 * (diag + 1e-300)) break;
        for (int p = 0; p < n - 1; ++p)
            for (int q = p + 1; q < n; ++q) {
                const double apq = A[p * n + q];
                if (std::fabs(apq) < 1e-300) continue;
                const double theta = (A[q * n + q] - A[p * n + p]) / (2.0 * apq);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < n; ++k) {   // columns p, q
                    const double akp = A[k * n + p], akq = A[k * n + q];
                    A[k * n + p] = c * akp - s * akq;
                    A[k * n + q] = s * akp + c * akq;
                }
                for (int k = 0; k < n; ++k) {   // rows p, q
                    const double apk = A[p * n + k], aqk = A[q * n + k];
                    A[p * n + k] = c * apk - s * aqk;
                    A[q * n + k] = s * apk + c * aqk;
                }
                for (int k = 0; k < n; ++k) {
                    const double vkp = V[k * n + p], vkq = V[k * n + q];
                    V[k * n + p] = c * vkp - s * vkq;
                    V[k * n + q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < n; ++i) d[i] = A[i * n + i];
}

int argMin(const double* d, int n)
{
    int k = 0;
    for (int i = 1; i < n; ++i) if (d[i] < d[k]) k = i;
    return k;
}

double det3(const double* M)
{
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

void mul3(const double* A, const double* B, double* C)
{
    double T[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) T[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
    std::memcpy(C, T, sizeof(T));
}

// nearest rotation to M (polar factor): R = M (M^T M)^(-1/2), determinant forced to +1
Matrix3d nearestRotation(const double* M)
{
    double S[9], V[9], d[3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) S[i * 3 + j] = M[i] * M[j] + M[3 + i] * M[3 + j] + M[6 + i] * M[6 + j];   // M^T M
    jacobiEigen(3, S, V, d);
    double P[9] = { 0, 0, 0, 0, 0, 0, 0, 0, 0 };   // (M^T M)^(-1/2) = V diag(1/sqrt(d)) V^T
    for (int k = 0; k < 3; ++k) {
        const double w = 1.0 / std::sqrt(std::max(d[k], 1e-300));
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) P[i * 3 + j] += w * V[i * 3 + k] * V[j * 3 + k];
    }
    Matrix3d R;
    mul3(M, P, R.m);
    if (det3(R.m) < 0) {   // reflect along the weakest direction
        const int k = argMin(d, 3);
        double F[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) F[i * 3 + j] = (i == j ? 1.0 : 0.0) - 2.0 * V[i * 3 + k] * V[j * 3 + k];
        mul3(R.m, F, R.m);
    }
    return R;
}

Matrix3d rodrigues(const double* w)
{
    const double th = std::sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    Matrix3d R;
    const double a = th < 1e-8 ? 1.0 - th * th / 6.0 : std::sin(th) / th;
    const double b = th < 1e-8 ? 0.5 - th * th / 24.0 : (1.0 - std::cos(th)) / (th * th);
    const double K[9] = { 0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0 };
    double K2[9];
    mul3(K, K, K2);
    for (int i = 0; i < 9; ++i) R.m[i] = ((i % 4 == 0) ? 1.0 : 0.0) + a * K[i] + b * K2[i];
    return R;
}

struct Pose { Matrix3d R; Vector3d t; };

// residual (pixels) and its 2x3 Jacobian with respect to the camera-frame point
void projectWithJacobian(const CameraModel& cam, const Vector3d& p, double* uv, double* J)
{
    const double* d = cam.distortionCoefficients;
    const double iz = 1.0 / p(2), x = p(0) * iz, y = p(1) * iz;
    const double r2 = x * x + y * y;
    const double c = 1.0 + r2 * (d[0] + r2 * (d[1] + r2 * d[4]));
    const double dc = d[0] + r2 * (2.0 * d[1] + 3.0 * d[4] * r2);
    uv[0] = cam.fx * (x * c + 2.0 * d[2] * x * y + d[3] * (r2 + 2.0 * x * x)) + cam.cx;
    uv[1] = cam.fy * (y * c + d[2] * (r2 + 2.0 * y * y) + 2.0 * d[3] * x * y) + cam.cy;
    if (!J) return;
    const double xdx = c + 2.0 * x * x * dc + 2.0 * d[2] * y + 6.0 * d[3] * x;
    const double xdy = 2.0 * x * y * dc + 2.0 * d[2] * x + 2.0 * d[3] * y;
    const double ydx = xdy;
    const double ydy = c + 2.0 * y * y * dc + 6.0 * d[2] * y + 2.0 * d[3] * x;
    // d(x,y)/d(X,Y,Z)
    const double xX = iz, xZ = -x * iz, yY = iz, yZ = -y * iz;
    J[0] = cam.fx * xdx * xX; J[1] = cam.fx * xdy * yY; J[2] = cam.fx * (xdx * xZ + xdy * yZ);
    J[3] = cam.fy * ydx * xX; J[4] = cam.fy * ydy * yY; J[5] = cam.fy * (ydx * xZ + ydy * yZ);
}

double reprojectionCost(const CameraModel& cam, const Pose& T, const std::vector<Vector3d>& X, const std::vector<Vector2d>& obs,
                        const std::vector<unsigned char>* use)
{
    double s = 0.0;
    for (size_t i = 0; i < X.size(); ++i) {
        if (use && !(*use)[i]) continue;
        const Vector3d p = T.R * X[i] + T.t;
        if (!(p(2) > 1e-9)) return std::numeric_limits<double>::infinity();
        double uv[2];
        projectWithJacobian(cam, p, uv, nullptr);
        s += (uv[0] - obs[i](0)) * (uv[0] - obs[i](0)) + (uv[1] - obs[i](1)) * (uv[1] - obs[i](1));
    }
    return s;
}

bool solve6(double* A, double* b)   // Cholesky, A symmetric positive definite 6x6 (destroyed), solution in b
{
    const int n = 6;
    for (int j = 0; j < n; ++j) {
        double s = A[j * n + j];
        for (int k = 0; k < j; ++k) s -= A[j * n + k] * A[j * n + k];
        if (!(s > 0)) return false;
        A[j * n + j] = std::sqrt(s);
        for (int i = j + 1; i < n; ++i) {
            double t = A[i * n + j];
            for (int k = 0; k < j; ++k) t -= A[i * n + k] * A[j * n + k];
            A[i * n + j] = t / A[j * n + j];
        }
    }
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < i; ++k) b[i] -= A[i * n + k] * b[k];
        b[i] /= A[i * n + i];
    }
    for (int i = n - 1; i >= 0; --i) {
        for (int k = i + 1; k < n; ++k) b[i] -= A[k * n + i] * b[k];
        b[i] /= A[i * n + i];
    }
    return true;
}

// Levenberg-Marquardt on the pixel reprojection error; update R <- exp(dw) R, t <- t + dt
int refinePose(const CameraModel& cam, Pose& T, const std::vector<Vector3d>& X, const std::vector<Vector2d>& obs,
               const std::vector<unsigned char>* use, double* finalCost)
{
    double cost = reprojectionCost(cam, T, X, obs, use);
    double lambda = 1e-3;
    int it = 0;
    for (; it < 100 && std::isfinite(cost); ++it) {
        double H[36] = { 0 }, g[6] = { 0 };
        for (size_t i = 0; i < X.size(); ++i) {
            if (use && !(*use)[i]) continue;
            const Vector3d q = T.R * X[i];
            const Vector3d p = q + T.t;
            double uv[2], Jp[6];
            projectWithJacobian(cam, p, uv, Jp);
            const double r[2] = { uv[0] - obs[i](0), uv[1] - obs[i](1) };
            // dp/d(dw) = -[q]x , dp/d(dt) = I
            const double S[9] = { 0, q(2), -q(1), -q(2), 0, q(0), q(1), -q(0), 0 };
            double J[12];
            for (int a = 0; a < 2; ++a) {
                for (int b = 0; b < 3; ++b) {
                    J[a * 6 + b] = Jp[a * 3] * S[b] + Jp[a * 3 + 1] * S[3 + b] + Jp[a * 3 + 2] * S[6 + b];
                    J[a * 6 + 3 + b] = Jp[a * 3 + b];
                }
            }
            for (int a = 0; a < 6; ++a) {
                g[a] += J[a] * r[0] + J[6 + a] * r[1];
                for (int b = 0; b < 6; ++b) H[a * 6 + b] += J[a] * J[b] + J[6 + a] * J[6 + b];
            }
        }
        bool improved = false;
        for (int tries = 0; tries < 12; ++tries) {
            double A[36], b[6];
            std::memcpy(A, H, sizeof(A));
            for (int a = 0; a < 6; ++a) { A[a * 6 + a] += lambda * (H[a * 6 + a] + 1e-12); b[a] = -g[a]; }
            if (!solve6(A, b)) { lambda *= 10.0; continue; }
            Pose C;
            const Matrix3d dR = rodrigues(b);
            mul3(dR.m, T.R.m, C.R.m);
            C.t = Vector3d(T.t(0) + b[3], T.t(1) + b[4], T.t(2) + b[5]);
            const double c2 = reprojectionCost(cam, C, X, obs, use);
            if (c2 < cost) {
                const double step = std::sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2] + b[3] * b[3] + b[4] * b[4] + b[5] * b[5]);
                const double gain = cost - c2;
                T = C;
                cost = c2;
                lambda = std::max(lambda * 0.2, 1e-12);
                improved = true;
                if (step < 1e-13 || gain <= 1e-16 * cost) { improved = false; }
                break;
            }
            lambda *= 10.0;
        }
        if (!improved) break;
    }
    // re-orthonormalise the accumulated rotation
    T.R = nearestRotation(T.R.m);
    if (finalCost) *finalCost = reprojectionCost(cam, T, X, obs, use);
    return it;
}

// closed-form start from normalised image points; false when the configuration is degenerate
bool closedFormStart(const std::vector<Vector3d>& X, const std::vector<Vector2d>& xn, Pose& T)
{
    const size_t n = X.size();
    if (n < 4) return false;
    Vector3d c;
    for (const auto& p : X) c = c + p;
    for (int a = 0; a < 3; ++a) c(a) /= (double)n;
    double C[9] = { 0 }, E[9], ev[3];
    for (const auto& p : X) {
        const Vector3d d = p - c;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) C[i * 3 + j] += d(i) * d(j);
    }
    jacobiEigen(3, C, E, ev);
    int order[3] = { 0, 1, 2 };
    std::sort(order, order + 3, [&](int a, int b) { return ev[a] > ev[b]; });
    const double scale = std::sqrt(std::max(ev[order[0]], 1e-300) / (double)n);
    const bool planar = n < 6 || ev[order[2]] < 1e-8 * ev[order[0]];
    if (planar) {
        if (ev[order[1]] < 1e-10 * ev[order[0]]) return false;   // collinear
        Vector3d e1(E[order[0]], E[3 + order[0]], E[6 + order[0]]), e2(E[order[1]], E[3 + order[1]], E[6 + order[1]]);
        const Vector3d e3(e1(1) * e2(2) - e1(2) * e2(1), e1(2) * e2(0) - e1(0) * e2(2), e1(0) * e2(1) - e1(1) * e2(0));
        // homography (a, b, 1) -> (x, y, 1), plane coordinates scaled to unit spread
        double A[81] = { 0 }, V[81], d[9];
        for (size_t i = 0; i < n; ++i) {
            const Vector3d q = X[i] - c;
            const double a = (q(0) * e1(0) + q(1) * e1(1) + q(2) * e1(2)) / scale, b = (q(0) * e2(0) + q(1) * e2(1) + q(2) * e2(2)) / scale;
            const double x = xn[i](0), y = xn[i](1);
            const double r1[9] = { a, b, 1, 0, 0, 0, -x * a, -x * b, -x };
            const double r2[9] = { 0, 0, 0, a, b, 1, -y * a, -y * b, -y };
            for (int p = 0; p < 9; ++p)
                for (int q2 = 0; q2 < 9; ++q2) A[p * 9 + q2] += r1[p] * r1[q2] + r2[p] * r2[q2];
        }
        jacobiEigen(9, A, V, d);
        const int k = argMin(d, 9);
        double h[9];
        for (int p = 0; p < 9; ++p) h[p] = V[p * 9 + k];
        // columns: h1 = (h0,h3,h6), h2 = (h1,h4,h7), h3 = (h2,h5,h8)
        const double n1 = std::sqrt(h[0] * h[0] + h[3] * h[3] + h[6] * h[6]), n2 = std::sqrt(h[1] * h[1] + h[4] * h[4] + h[7] * h[7]);
        if (!(n1 > 0) || !(n2 > 0)) return false;
        double s = 2.0 / (n1 + n2);
        if (h[8] * s < 0) s = -s;   // the plane origin (centroid) must lie in front of the camera
        const double r1[3] = { h[0] * s, h[3] * s, h[6] * s }, r2[3] = { h[1] * s, h[4] * s, h[7] * s };
        const double r3[3] = { r1[1] * r2[2] - r1[2] * r2[1], r1[2] * r2[0] - r1[0] * r2[2], r1[0] * r2[1] - r1[1] * r2[0] };
        const double M[9] = { r1[0], r2[0], r3[0], r1[1], r2[1], r3[1], r1[2], r2[2], r3[2] };
        const Matrix3d Rp = nearestRotation(M);   // plane frame -> camera
        const double Et[9] = { e1(0), e1(1), e1(2), e2(0), e2(1), e2(2), e3(0), e3(1), e3(2) };   // object -> plane frame
        mul3(Rp.m, Et, T.R.m);
        const Vector3d tp(h[2] * s * scale, h[5] * s * scale, h[8] * s * scale);   // camera position of the centroid
        const Vector3d Rc = T.R * c;
        T.t = tp - Rc;
        return true;
    }
    // DLT for general position: x = (P1.X~)/(P3.X~), y = (P2.X~)/(P3.X~), object points centred and scaled
    double A[144] = { 0 }, V[144], d[12];
    for (size_t i = 0; i < n; ++i) {
        const Vector3d q = X[i] - c;
        const double Xh[4] = { q(0) / scale, q(1) / scale, q(2) / scale, 1.0 };
        const double x = xn[i](0), y = xn[i](1);
        double r1[12], r2[12];
        for (int a = 0; a < 4; ++a) { r1[a] = Xh[a]; r1[4 + a] = 0; r1[8 + a] = -x * Xh[a]; r2[a] = 0; r2[4 + a] = Xh[a]; r2[8 + a] = -y * Xh[a]; }
        for (int p = 0; p < 12; ++p)
            for (int q2 = 0; q2 < 12; ++q2) A[p * 12 + q2] += r1[p] * r1[q2] + r2[p] * r2[q2];
    }
    jacobiEigen(12, A, V, d);
    const int k = argMin(d, 12);
    double P[12];
    for (int p = 0; p < 12; ++p) P[p] = V[p * 12 + k];
    double M[9] = { P[0], P[1], P[2], P[4], P[5], P[6], P[8], P[9], P[10] };
    double dt = det3(M);
    if (std::fabs(dt) < 1e-300) return false;
    const double s = (dt > 0 ? 1.0 : -1.0) / std::cbrt(std::fabs(dt));
    for (double& m : M) m *= s;
    T.R = nearestRotation(M);
    const Vector3d tc(P[3] * s * scale, P[7] * s * scale, P[11] * s * scale);   // camera position of the centroid
    const Vector3d Rc = T.R * c;
    T.t = tc - Rc;
    return true;
}

PnPResult finish(const CameraModel& cam, const Pose& T, const std::vector<Vector3d>& X, const std::vector<Vector2d>& obs,
                 const std::vector<unsigned char>* use, int iterations)
{
    PnPResult r;
    r.R = T.R;
    r.t = T.t;
    size_t m = 0;
    for (size_t i = 0; i < X.size(); ++i) m += (!use || (*use)[i]) ? 1 : 0;
    const double c = reprojectionCost(cam, T, X, obs, use);
    r.rmsError = m ? std::sqrt(c / (double)m) : 0.0;
    r.iterations = iterations;
    r.ok = std::isfinite(c);
    return r;
}
}   // namespace

Vector2d projectPinhole(const CameraModel& cam, const Vector3d& p)
{
    double uv[2];
    projectWithJacobian(cam, p, uv, nullptr);
    return Vector2d(uv[0], uv[1]);
}

Vector2d undistortPixel(const CameraModel& cam, const Vector2d& uv)
{
    const double* d = cam.distortionCoefficients;
    const double x0 = (uv(0) - cam.cx) / cam.fx, y0 = (uv(1) - cam.cy) / cam.fy;
    double x = x0, y = y0;
    for (int it = 0; it < 100; ++it) {
        const double r2 = x * x + y * y;
        const double c = 1.0 + r2 * (d[0] + r2 * (d[1] + r2 * d[4]));
        const double dx = 2.0 * d[2] * x * y + d[3] * (r2 + 2.0 * x * x), dy = d[2] * (r2 + 2.0 * y * y) + 2.0 * d[3] * x * y;
        const double xn = (x0 - dx) / c, yn = (y0 - dy) / c;
        const double e = std::fabs(xn - x) + std::fabs(yn - y);
        x = xn; y = yn;
        if (e < 1e-14) break;
    }
    return Vector2d(x, y);
}

PnPResult solvePnP(const std::vector<Vector3d>& X, const std::vector<Vector2d>& obs, const CameraModel& cam)
{
    PnPResult bad;
    if (X.size() != obs.size() || X.size() < 4) return bad;
    std::vector<Vector2d> xn(obs.size());
    for (size_t i = 0; i < obs.size(); ++i) xn[i] = undistortPixel(cam, obs[i]);
    Pose T;
    if (!closedFormStart(X, xn, T)) return bad;
    const int it = refinePose(cam, T, X, obs, nullptr, nullptr);
    return finish(cam, T, X, obs, nullptr, it);
}

PnPResult solvePnPRobust(const std::vector<Vector3d>& X, const std::vector<Vector2d>& obs, const CameraModel& cam, double thr,
                         std::vector<unsigned char>* inliersOut)
{
    PnPResult bad;
    const size_t n = X.size();
    if (n != obs.size() || n < 4) return bad;
    std::vector<Vector2d> xn(n);
    for (size_t i = 0; i < n; ++i) xn[i] = undistortPixel(cam, obs[i]);
    auto classify = [&](const Pose& T, std::vector<unsigned char>& in, double& err) {
        size_t cnt = 0;
        err = 0.0;
        in.assign(n, 0);
        for (size_t i = 0; i < n; ++i) {
            const Vector3d p = T.R * X[i] + T.t;
            if (!(p(2) > 1e-9)) continue;
            const Vector2d uv = projectPinhole(cam, p);
            const double e = std::hypot(uv(0) - obs[i](0), uv(1) - obs[i](1));
            if (e < thr) { in[i] = 1; ++cnt; err += e * e; }
        }
        return cnt;
    };
    // hypotheses: all correspondences at once, and the four corners of every tag
    std::vector<Pose> hyp;
    {
        Pose T;
        if (closedFormStart(X, xn, T)) { refinePose(cam, T, X, obs, nullptr, nullptr); hyp.push_back(T); }
    }
    if (n % 4 == 0 && n > 4) {
        for (size_t g = 0; g < n / 4; ++g) {
            std::vector<Vector3d> Xg(X.begin() + 4 * g, X.begin() + 4 * g + 4);
            std::vector<Vector2d> og(obs.begin() + 4 * g, obs.begin() + 4 * g + 4), xg(xn.begin() + 4 * g, xn.begin() + 4 * g + 4);
            Pose T;
            if (!closedFormStart(Xg, xg, T)) continue;
            refinePose(cam, T, Xg, og, nullptr, nullptr);
            hyp.push_back(T);
        }
    }
    if (hyp.empty()) return bad;
    size_t bestCount = 0;
    double bestErr = std::numeric_limits<double>::infinity();
    Pose best = hyp[0];
    std::vector<unsigned char> in, bestIn(n, 1);
    for (const Pose& T : hyp) {
        double err;
        const size_t cnt = classify(T, in, err);
        if (cnt > bestCount || (cnt == bestCount && err < bestErr)) { bestCount = cnt; bestErr = err; best = T; bestIn = in; }
    }
    if (bestCount < 4) {   // nothing agrees within the threshold: the least-squares answer over everything (what RANSAC's fallback amounts to)
        best = hyp[0];
        bestIn.assign(n, 1);
    }
    int its = 0;
    for (int round = 0; round < 2; ++round) {   // refit on the inliers, re-classify, refit
        its += refinePose(cam, best, X, obs, &bestIn, nullptr);
        double err;
        std::vector<unsigned char> again;
        if (classify(best, again, err) >= 4) bestIn = again;
    }
    if (inliersOut) *inliersOut = bestIn;
    return finish(cam, best, X, obs, &bestIn, its);
}

Vector4d rotationMatrixToQuaternion(const Matrix3d& R)
{
    const double* m = R.m;
    const double tr = m[0] + m[4] + m[8];
    double w, x, y, z;
    if (tr > 0) {
        const double s = std::sqrt(tr + 1.0) * 2.0;
        w = 0.25 * s; x = (m[7] - m[5]) / s; y = (m[2] - m[6]) / s; z = (m[3] - m[1]) / s;
    } else if (m[0] > m[4] && m[0] > m[8]) {
        const double s = std::sqrt(1.0 + m[0] - m[4] - m[8]) * 2.0;
        w = (m[7] - m[5]) / s; x = 0.25 * s; y = (m[1] + m[3]) / s; z = (m[2] + m[6]) / s;
    } else if (m[4] > m[8]) {
        const double s = std::sqrt(1.0 + m[4] - m[0] - m[8]) * 2.0;
        w = (m[2] - m[6]) / s; x = (m[1] + m[3]) / s; y = 0.25 * s; z = (m[5] + m[7]) / s;
    } else {
        const double s = std::sqrt(1.0 + m[8] - m[0] - m[4]) * 2.0;
        w = (m[3] - m[1]) / s; x = (m[2] + m[6]) / s; y = (m[5] + m[7]) / s; z = 0.25 * s;
    }
    const double nrm = std::sqrt(w * w + x * x + y * y + z * z);
    return Vector4d(w / nrm, x / nrm, y / nrm, z / nrm);
}
}   // namespace visual_marker_mapping
