"""Pins the CPU oracle (and the product's device maths compiled for the host) against the committed
known-answer vectors tests/golden/corner_kats.json — cv2.projectPoints (OpenCV 4.13) and 50-digit
mpmath restatements of TagReconstructionCostFunction.h:101-159 + ceres::QuaternionParameterization.
The reference itself holds no fixtures for this path (parity unpinned)."""
import ctypes as C
import json
import os
from types import SimpleNamespace

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def kats():
    with open(os.path.join(HERE, "golden", "corner_kats.json")) as f:
        return json.load(f)


def case_scene(intr, c):
    """1 camera, 1 marker, 1 tag observation whose corner `corner_index` carries the observation."""
    uv = np.zeros((1, 8))
    uv[0, 2 * c["corner_index"]:2 * c["corner_index"] + 2] = c["obs"]
    return SimpleNamespace(
        intr=np.array(intr), cam_q=np.array([c["cam_q"]]), cam_t=np.array([c["cam_t"]]), tag_q=np.array([c["tag_q"]]),
        tag_t=np.array([c["tag_t"]]), tag_wh=np.array([c["tag_wh"]]), fixed_tag=-1, to_cam=np.array([0], dtype=np.int32),
        to_tag=np.array([0], dtype=np.int32), to_uv=uv)


def rel(a, b):
    return np.linalg.norm(np.asarray(a) - np.asarray(b)) / max(np.linalg.norm(b), 1.0)


def test_oracle_matches_mpmath_and_cv2(kats, oracle):
    assert len(kats["cases"]) >= 24
    worst_r = worst_j = 0.0
    for c in kats["cases"]:
        ev = oracle.evaluate(case_scene(kats["intrinsics"], c))
        i = c["corner_index"]
        r_mp = np.array([float(v) for v in c["mp_residual"]])
        J_mp = np.array([[float(v) for v in row] for row in c["mp_jac_local"]])
        worst_r = max(worst_r, rel(ev.residuals[i], r_mp))
        worst_j = max(worst_j, rel(ev.jac_local[i], J_mp))
        # cv2: projection and d(pixel)/d(tvec) = d r / d t_c ; d(pixel)/d(rvec at 0) is the full-angle
        # tangent, Ceres' QuaternionParameterization is half-angle: factor 2
        proj = ev.residuals[i] + np.array(c["obs"])
        assert rel(proj, c["cv2_projection"]) < 1e-12
        assert rel(ev.jac_local[i][:, 0:3], c["cv2_dp_dtvec"]) < 1e-11
        assert rel(ev.jac_local[i][:, 3:6], 2.0 * np.array(c["cv2_dp_drvec_at_0"])) < 1e-11
    assert worst_r < 1e-13 and worst_j < 1e-12, (worst_r, worst_j)


def test_survey_probe_point(kats, oracle):
    """X=(0.3,-0.2,2.5) -> (3982.03977278, 1351.89242311)."""
    c = dict(corner_index=0, obs=[0.0, 0.0], cam_q=[1, 0, 0, 0], cam_t=[0, 0, 0], tag_q=[1, 0, 0, 0],
             tag_t=[0.3 + 0.05, -0.2 + 0.05, 2.5], tag_wh=[0.1, 0.1])   # corner 0 = (-w/2,-h/2,0)
    ev = oracle.evaluate(case_scene(kats["intrinsics"], c))
    assert np.allclose(ev.residuals[0], kats["probe_cv2"], rtol=0, atol=1e-9)
    assert np.allclose(ev.residuals[0], [3982.03977278, 1351.89242311], atol=1e-7)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def harness_corner(h, intr, c, loss=0, a=1.0):
    r, J = np.zeros(2), np.zeros(24)
    arrs = [np.array(x, dtype=np.float64) for x in (intr, c["cam_q"], c["cam_t"], c["tag_q"], c["tag_t"], c["tag_wh"])]
    cost = h.hm_corner_full(*[_dp(x) for x in arrs], C.c_int(c["corner_index"]), C.c_double(c["obs"][0]), C.c_double(c["obs"][1]),
                            C.c_int(loss), C.c_double(a), _dp(r), _dp(J))
    cost2 = h.hm_corner_cost(*[_dp(x) for x in arrs], C.c_int(c["corner_index"]), C.c_double(c["obs"][0]), C.c_double(c["obs"][1]),
                             C.c_int(loss), C.c_double(a))
    return cost, cost2, r, J.reshape(2, 12)


def test_device_maths_matches_golden(kats, math_harness):
    """csrc/ba_math.h (closed-form Jacobian, the code the kernels run) against the same vectors."""
    for c in kats["cases"]:
        cost, cost2, r, J = harness_corner(math_harness, kats["intrinsics"], c)
        r_mp = np.array([float(v) for v in c["mp_residual"]])
        J_mp = np.array([[float(v) for v in row] for row in c["mp_jac_local"]])
        assert rel(r, r_mp) < 1e-13
        assert rel(J, J_mp) < 1e-12
        assert abs(cost - 0.5 * r_mp @ r_mp) <= 1e-12 * cost and cost == cost2


@pytest.mark.parametrize("loss,a", [(1, 1.0), (1, 25.0), (2, 3.0), (2, 0.5)])
def test_device_maths_matches_oracle_with_loss(kats, oracle, math_harness, loss, a):
    for c in kats["cases"]:
        obs = np.array(c["obs"])
        for shift in (0.0, 0.3, 5.0, 200.0):     # inlier and outlier branches of the loss
            ev0 = oracle.evaluate(case_scene(kats["intrinsics"], c))
            cc = dict(c)
            cc["obs"] = (obs + ev0.residuals[c["corner_index"]] - shift).tolist()   # residual = (shift, shift)
            ev = oracle.evaluate(case_scene(kats["intrinsics"], cc), loss, a)
            cost, cost2, r, J = harness_corner(math_harness, kats["intrinsics"], cc, loss, a)
            i = c["corner_index"]
            # residuals are differences of ~1e3 px quantities: absolute 1e-9 px = 1e-12 relative to the pixel
            assert np.abs(r - ev.residuals[i]).max() < 1e-9 and rel(J, ev.jac_local[i]) < 1e-11   # gate: 1e-10
            assert abs(cost - cost2) <= 1e-15 * max(cost, 1.0)


def test_loss_functions_closed_form(oracle):
    """Huber / Cauchy per Ceres loss_function.cc: block cost = 0.5*rho(s)."""
    from visual_marker_mapping_b200 import synth

    s = synth.make_scene(n_markers=8, n_cameras=10, markers_per_camera=5, seed=2, outlier_frac=0.3)
    r = oracle.evaluate(s, 0, 1.0).residuals
    sq = (r * r).sum(axis=1)
    huber = np.where(sq <= 1.0, sq, 2.0 * np.sqrt(sq) - 1.0)
    assert np.isclose(oracle.evaluate(s, oracle.LOSS_HUBER, 1.0).cost, 0.5 * huber.sum(), rtol=1e-13)
    cauchy = 9.0 * np.log1p(sq / 9.0)
    assert np.isclose(oracle.evaluate(s, oracle.LOSS_CAUCHY, 3.0).cost, 0.5 * cauchy.sum(), rtol=1e-13)
    # Corrector: residual and Jacobian scaled by sqrt(rho')
    rh = oracle.evaluate(s, oracle.LOSS_HUBER, 1.0).residuals
    w = np.where(sq <= 1.0, 1.0, 1.0 / np.sqrt(np.sqrt(sq)))
    assert np.allclose(rh, r * w[:, None], rtol=1e-13, atol=0)


def test_quat_plus_and_spd_inverse(oracle, math_harness):
    rng = np.random.default_rng(0)
    for _ in range(50):
        x = rng.normal(size=4)
        d = rng.normal(size=3) * rng.choice([1e-9, 1e-3, 0.5])
        out = np.zeros(4)
        math_harness.hm_quat_plus(_dp(x), _dp(d), _dp(out))
        n = np.linalg.norm(d)
        qd = np.concatenate([[np.cos(n)], np.sin(n) / n * d])
        w0, x0, y0, z0 = qd
        w1, x1, y1, z1 = x
        ref = np.array([w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1, w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1,
                        w0 * y1 - x0 * z1 + y0 * w1 + z0 * x1, w0 * z1 + x0 * y1 - y0 * x1 + z0 * w1])
        assert np.allclose(out, ref, rtol=1e-14, atol=1e-16)
        A = rng.normal(size=(8, 6))
        M = A.T @ A + 1e-3 * np.eye(6)
        Mi = M.copy()
        assert math_harness.hm_spd_inverse6(_dp(Mi)) == 1
        assert np.allclose(Mi @ M, np.eye(6), atol=1e-9)
    bad = -np.eye(6)
    assert math_harness.hm_spd_inverse6(_dp(bad)) == 0
