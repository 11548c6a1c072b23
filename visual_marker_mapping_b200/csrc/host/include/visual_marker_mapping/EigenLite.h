// EigenLite.h — the few fixed-size vector types the TagReconstructor API exposes.
// The reference uses Eigen::Vector{2,3,4}d (Camera.h:13-14, TagReconstructor.h:19-20, DetectionResults.h:14);
// Eigen is not part of this build, so the facade ships look-alikes with the members the reference's
// callers touch: operator()(i), x()/y()/z(), Zero(), UnitX(), norm().
#ifndef VISUAL_MARKER_MAPPING_EIGENLITE_H_
#define VISUAL_MARKER_MAPPING_EIGENLITE_H_
#include <cmath>
#include <cstddef>

namespace visual_marker_mapping
{
template <int N> struct VectorNd
{
    double v[N];
    VectorNd() { for (int i = 0; i < N; ++i) v[i] = 0.0; }
    VectorNd(double a, double b) { static_assert(N == 2, "size"); v[0] = a; v[1] = b; }
    VectorNd(double a, double b, double c) { static_assert(N == 3, "size"); v[0] = a; v[1] = b; v[2] = c; }
    VectorNd(double a, double b, double c, double d) { static_assert(N == 4, "size"); v[0] = a; v[1] = b; v[2] = c; v[3] = d; }
    double& operator()(int i) { return v[i]; }
    double operator()(int i) const { return v[i]; }
    double& operator[](int i) { return v[i]; }
    double operator[](int i) const { return v[i]; }
    double& x() { return v[0]; }
    double& y() { return v[1]; }
    double& z() { static_assert(N >= 3, "size"); return v[2]; }
    double x() const { return v[0]; }
    double y() const { return v[1]; }
    double z() const { static_assert(N >= 3, "size"); return v[2]; }
    double* data() { return v; }
    const double* data() const { return v; }
    static VectorNd Zero() { return VectorNd(); }
    static VectorNd UnitX() { VectorNd r; r.v[0] = 1.0; return r; }   // identity quaternion in (w,x,y,z) order
    double norm() const { double s = 0; for (int i = 0; i < N; ++i) s += v[i] * v[i]; return std::sqrt(s); }
    VectorNd operator-(const VectorNd& o) const { VectorNd r; for (int i = 0; i < N; ++i) r.v[i] = v[i] - o.v[i]; return r; }
    VectorNd operator+(const VectorNd& o) const { VectorNd r; for (int i = 0; i < N; ++i) r.v[i] = v[i] + o.v[i]; return r; }
};
typedef VectorNd<2> Vector2d;
typedef VectorNd<3> Vector3d;
typedef VectorNd<4> Vector4d;

// 3x3 rotation from a quaternion (w,x,y,z) WITHOUT normalisation — what Eigen::Quaterniond::toRotationMatrix
// does (used by computeMarkerCorners3D, TagReconstructor.h:33-43, and the reprojection statistics).
struct Matrix3d
{
    double m[9];
    Vector3d operator*(const Vector3d& p) const
    {
        return Vector3d(m[0] * p(0) + m[1] * p(1) + m[2] * p(2), m[3] * p(0) + m[4] * p(1) + m[5] * p(2),
            m[6] * p(0) + m[7] * p(1) + m[8] * p(2));
    }
};
inline Matrix3d quaternionToRotationMatrix(const Vector4d& q)
{
    const double w = q(0), x = q(1), y = q(2), z = q(3);
    const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
    const double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y,
                 tyz = tz * y, tzz = tz * z;
    Matrix3d R;
    R.m[0] = 1 - (tyy + tzz); R.m[1] = txy - twz; R.m[2] = txz + twy;
    R.m[3] = txy + twz; R.m[4] = 1 - (txx + tzz); R.m[5] = tyz - twx;
    R.m[6] = txz - twy; R.m[7] = tyz + twx; R.m[8] = 1 - (txx + tyy);
    return R;
}
}
#endif
