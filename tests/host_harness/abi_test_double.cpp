// abi_test_double.cpp — TEST INFRASTRUCTURE ONLY: a stand-in for the handful of C-ABI calls the C++ facade makes
// (include/vmm_ba.h), so that the facade's HOST logic — flattening the maps, aggregating the per-observation error
// sums per image / per tag, culling, write-back, the closing sequence of the CLI — runs in the `-m "not gpu"` suite.
// It is never linked into or loaded by the product: tests put it in front of libvmm_ba.so with LD_PRELOAD.
//   vmm_ba_solve is the identity (the poses stay as uploaded, termination "convergence"),
//   vmm_ba_reprojection_error_sums restates k_reprojection_sums (csrc/ba_kernels.cuh; reference
//   src/TagReconstructor.cpp:340-428 with CameraModel::projectPoint, src/CameraModel.cpp:6-26) in plain C++.
// Nothing here says anything about the CUDA engine; the GPU tests do that.
#include "../../include/vmm_ba.h"

#include <cmath>
#include <cstring>
#include <vector>

struct vmm_ba_engine
{
    double fx = 0, fy = 0, cx = 0, cy = 0, d[5] = { 0, 0, 0, 0, 0 };
    std::vector<double> camQ, camT, tagQ, tagT, tagWH, uv;
    std::vector<int32_t> toCam, toTag;
};

namespace
{
void eigenRotation(const double* q, double R[9])   // Eigen::Quaterniond::toRotationMatrix, q = (w, x, y, z), not normalised
{
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - w * z); R[2] = 2 * (x * z + w * y);
    R[3] = 2 * (x * y + w * z); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - w * x);
    R[6] = 2 * (x * z - w * y); R[7] = 2 * (y * z + w * x); R[8] = 1 - 2 * (x * x + y * y);
}
}

extern "C" {
int vmm_ba_create(int, vmm_ba_handle* out)
{
    *out = new vmm_ba_engine();
    return VMM_BA_OK;
}
int vmm_ba_destroy(vmm_ba_handle h)
{
    delete h;
    return VMM_BA_OK;
}
const char* vmm_ba_last_error(vmm_ba_handle) { return "abi test double"; }
int vmm_ba_set_intrinsics(vmm_ba_handle h, double fx, double fy, double cx, double cy, const double dist[5])
{
    h->fx = fx; h->fy = fy; h->cx = cx; h->cy = cy;
    std::memcpy(h->d, dist, sizeof(h->d));
    return VMM_BA_OK;
}
int vmm_ba_set_problem(vmm_ba_handle h, int64_t nc, const double* cq, const double* ct, int64_t nm, const double* tq, const double* tt,
    const double* wh, int64_t, int64_t nto, const int32_t* toCam, const int32_t* toTag, const double* uv)
{
    h->camQ.assign(cq, cq + 4 * nc); h->camT.assign(ct, ct + 3 * nc);
    h->tagQ.assign(tq, tq + 4 * nm); h->tagT.assign(tt, tt + 3 * nm); h->tagWH.assign(wh, wh + 2 * nm);
    h->toCam.assign(toCam, toCam + nto); h->toTag.assign(toTag, toTag + nto); h->uv.assign(uv, uv + 8 * nto);
    for (int64_t k = 0; k < nto; ++k)
        if (toCam[k] < 0 || toCam[k] >= nc || toTag[k] < 0 || toTag[k] >= nm) return VMM_BA_ERR_INVALID_ARGUMENT;
    return VMM_BA_OK;
}
int vmm_ba_solve(vmm_ba_handle, const vmm_ba_options*, vmm_ba_summary* s)
{
    std::memset(s, 0, sizeof(*s));
    s->termination_type = 0;
    std::strcpy(s->message, "abi test double: identity solve");
    return VMM_BA_OK;
}
int vmm_ba_get_poses(vmm_ba_handle h, double* cq, double* ct, double* tq, double* tt)
{
    std::memcpy(cq, h->camQ.data(), sizeof(double) * h->camQ.size()); std::memcpy(ct, h->camT.data(), sizeof(double) * h->camT.size());
    std::memcpy(tq, h->tagQ.data(), sizeof(double) * h->tagQ.size()); std::memcpy(tt, h->tagT.data(), sizeof(double) * h->tagT.size());
    return VMM_BA_OK;
}
int vmm_ba_marker_translation_covariance(vmm_ba_handle h, int, double, double* cov)
{
    for (size_t i = 0; i < 9 * h->tagWH.size() / 2; ++i) cov[i] = 0.0;
    return VMM_BA_OK;
}
int vmm_ba_reprojection_error_sums(vmm_ba_handle h, double* sums)
{
    static const double sx[4] = { -1, 1, 1, -1 }, sy[4] = { -1, -1, 1, 1 };   // LL, LR, UR, UL (TagReconstructor.h:44-52)
    const double k1 = h->d[0], k2 = h->d[1], p1 = h->d[2], p2 = h->d[3], k3 = h->d[4];
    for (size_t k = 0; k < h->toCam.size(); ++k) {
        const int ci = h->toCam[k], mi = h->toTag[k];
        double Rm[9], Rc[9], s = 0.0;
        eigenRotation(&h->tagQ[4 * mi], Rm);
        eigenRotation(&h->camQ[4 * ci], Rc);
        for (int c = 0; c < 4; ++c) {
            const double px = sx[c] * h->tagWH[2 * mi] / 2.0, py = sy[c] * h->tagWH[2 * mi + 1] / 2.0;
            double pw[3], pc[3];
            for (int i = 0; i < 3; ++i) pw[i] = Rm[3 * i] * px + Rm[3 * i + 1] * py + h->tagT[3 * mi + i];
            for (int i = 0; i < 3; ++i) pc[i] = Rc[3 * i] * pw[0] + Rc[3 * i + 1] * pw[1] + Rc[3 * i + 2] * pw[2] + h->camT[3 * ci + i];
            double x = pc[0] / pc[2], y = pc[1] / pc[2];
            const double r2 = x * x + y * y;
            x = x * (1 + r2 * (k1 + r2 * (k2 + r2 * k3))) + 2 * p1 * x * y + p2 * (r2 + 2 * x * x);   // x is overwritten first, as in the reference
            y = y * (1 + r2 * (k1 + r2 * (k2 + r2 * k3))) + 2 * p2 * x * y + p1 * (r2 + 2 * y * y);
            const double du = h->fx * x + h->cx - h->uv[8 * k + 2 * c], dv = h->fy * y + h->cy - h->uv[8 * k + 2 * c + 1];
            s += std::sqrt(du * du + dv * dv);
        }
        sums[k] = s;
    }
    return VMM_BA_OK;
}
}
