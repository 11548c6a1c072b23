// Test-only: compiles the product's device maths (csrc/ba_math.h) for the host so that the
// closed-form Jacobian, loss, Plus and 6x6 inverse can be checked against the oracle without a GPU.
#include "ba_math.h"
#include <cstdint>
extern "C" {
double hm_corner_full(const double* intr, const double* cam_q, const double* cam_t, const double* tag_q,
                      const double* tag_t, const double* wh, int c, double u, double v, int loss, double a,
                      double* r, double* J) {
    vmm::Intrinsics in{intr[0], intr[1], intr[2], intr[3], intr[4], intr[5], intr[6], intr[7], intr[8]};
    double cam[vmm::CAM_PREP], mk[vmm::MARK_PREP];
    vmm::prep_camera(cam_q, cam_t, cam);
    vmm::prep_marker(tag_q, tag_t, wh[0], wh[1], mk);
    return vmm::corner_full(in, cam, mk, c, u, v, loss, a, r, J);
}
double hm_corner_cost(const double* intr, const double* cam_q, const double* cam_t, const double* tag_q,
                      const double* tag_t, const double* wh, int c, double u, double v, int loss, double a) {
    vmm::Intrinsics in{intr[0], intr[1], intr[2], intr[3], intr[4], intr[5], intr[6], intr[7], intr[8]};
    double cam[vmm::CAM_PREP], mk[vmm::MARK_PREP];
    vmm::prep_camera(cam_q, cam_t, cam);
    vmm::prep_marker(tag_q, tag_t, wh[0], wh[1], mk);
    return vmm::corner_cost(in, cam, mk, c, u, v, loss, a);
}
int hm_spd_inverse6(double* A) { return vmm::spd_inverse6(A) ? 1 : 0; }
void hm_quat_plus(const double* x, const double* d, double* out) { vmm::quat_plus(x, d, out); }
}
