"""Properties of the CPU oracle's Ceres-semantics LM (oracle/vmm_oracle.cpp) that
do not need Ceres to check: convergence to ground truth on zero-noise scenes, invariances, the loss
limits, the analytic/dual-number Jacobian against central differences through Plus."""
import numpy as np
import pytest

from visual_marker_mapping_b200 import synth


def small_scene(**kw):
    args = dict(n_markers=10, n_cameras=24, markers_per_camera=8, seed=42)
    args.update(kw)
    return synth.make_scene(**args)


def test_zero_noise_returns_to_ground_truth(oracle):
    s = small_scene(noise_px=0.0, float32_round=False)
    assert oracle.evaluate(s.at_ground_truth()).cost < 1e-12
    r = oracle.solve(s, max_num_iterations=100, function_tolerance=1e-16, parameter_tolerance=1e-14)
    assert r.summary.final_cost < 1e-14
    obs = np.unique(s.to_tag)
    assert np.abs(r.tag_t - s.tag_t_gt)[obs].max() < 1e-8
    assert np.abs(r.cam_t - s.cam_t_gt).max() < 1e-8
    assert synth.quat_angle(r.cam_q, s.cam_q_gt).max() < 1e-8
    # the origin tag is constant (src/TagReconstructor.cpp:669-673)
    assert np.array_equal(r.tag_q[0], s.tag_q[0]) and np.array_equal(r.tag_t[0], s.tag_t[0])


def test_default_options_are_the_ceres_defaults(oracle):
    o = oracle.default_options()
    assert (o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance) == (1e-6, 1e-10, 1e-8)
    assert o.initial_trust_region_radius == 1e4 and o.min_relative_decrease == 1e-3 and o.jacobi_scaling == 1


def test_trace_is_monotone_and_terminates_by_function_tolerance(oracle, scene_c1):
    r = oracle.solve(scene_c1, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    assert r.summary.termination_type == 0 and b"tolerance" in r.summary.message
    costs = r.trace[r.trace[:, 5] == 1, 0]
    assert np.all(np.diff(costs) < 0)
    assert r.summary.final_cost == costs[-1]
    assert r.summary.num_iterations == r.summary.num_successful_steps + r.summary.num_unsuccessful_steps + 1 or \
        r.summary.num_iterations == r.summary.num_successful_steps + r.summary.num_unsuccessful_steps + 2


def test_huber_with_huge_scale_is_no_loss(oracle, scene_c1):
    a = oracle.solve(scene_c1, max_num_iterations=30)
    b = oracle.solve(scene_c1, max_num_iterations=30, loss_kind=oracle.LOSS_HUBER, loss_scale=1e9)
    assert abs(a.summary.final_cost - b.summary.final_cost) <= 1e-12 * a.summary.final_cost
    assert np.abs(a.cam_t - b.cam_t).max() < 1e-12


def test_observation_order_does_not_matter(oracle, scene_c1):
    s = scene_c1.copy()
    perm = np.random.default_rng(1).permutation(s.n_tag_obs)
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[perm].copy(), s.to_tag[perm].copy(), s.to_uv[perm].copy()
    a = oracle.solve(scene_c1, max_num_iterations=30, loss_kind=oracle.LOSS_HUBER)
    b = oracle.solve(s, max_num_iterations=30, loss_kind=oracle.LOSS_HUBER)
    assert abs(a.summary.final_cost - b.summary.final_cost) <= 1e-9 * a.summary.final_cost
    assert np.abs(a.tag_t - b.tag_t).max() < 1e-8


def test_quaternion_scale_invariance(oracle, scene_c1):
    """QuaternionRotatePoint normalises: cost and local Jacobian do not depend on |q|."""
    s = scene_c1.copy()
    s.cam_q = s.cam_q * 1.7
    s.tag_q = s.tag_q * 0.4
    a, b = oracle.evaluate(scene_c1), oracle.evaluate(s)
    assert abs(a.cost - b.cost) <= 1e-12 * a.cost
    # local Jacobian of a scaled quaternion: Plus keeps |q|, tangent columns scale with nothing
    assert np.abs(a.jac_local[:, :, :3] - b.jac_local[:, :, :3]).max() <= 1e-9 * np.abs(a.jac_local).max()


@pytest.mark.parametrize("loss,scale", [(0, 1.0), (1, 1.0), (2, 3.0)])
def test_gradient_matches_central_differences_through_plus(oracle, loss, scale):
    s = small_scene(n_markers=6, n_cameras=8, markers_per_camera=5, outlier_frac=0.2)
    ev = oracle.evaluate(s, loss, scale)
    n = 6 * (s.n_cameras + s.n_markers)
    rng = np.random.default_rng(3)
    for idx in rng.choice(n, 25, replace=False):
        blk = idx // 6
        if blk >= s.n_cameras and blk - s.n_cameras == s.fixed_tag:
            continue
        h = 1e-6
        d = np.zeros(n)
        d[idx] = h
        sp, sm = s.copy(), s.copy()
        sp.cam_q, sp.cam_t, sp.tag_q, sp.tag_t = oracle.plus(s, d)
        sm.cam_q, sm.cam_t, sm.tag_q, sm.tag_t = oracle.plus(s, -d)
        fd = (oracle.evaluate(sp, loss, scale).cost - oracle.evaluate(sm, loss, scale).cost) / (2 * h)
        assert abs(fd - ev.gradient[idx]) <= 2e-5 * max(abs(ev.gradient[idx]), 1.0), (idx, fd, ev.gradient[idx])


def test_fixed_tag_and_unobserved_blocks_have_zero_gradient(oracle):
    s = small_scene()
    keep = s.to_tag != 3
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[keep].copy(), s.to_tag[keep].copy(), s.to_uv[keep].copy()
    g = oracle.evaluate(s).gradient.reshape(-1, 6)
    assert not g[s.n_cameras + 0].any() and not g[s.n_cameras + 3].any()
    r = oracle.solve(s, max_num_iterations=20)
    assert np.array_equal(r.tag_t[3], s.tag_t[3]) and np.array_equal(r.tag_q[3], s.tag_q[3])


def test_reprojection_error_sums_follow_cameramodel_quirk(oracle):
    """CameraModel::projectPoint reuses the distorted x in the y formula (src/CameraModel.cpp:20-23):
    the statistic differs slightly from the functor's residual norm, and matches a numpy restatement."""
    s = small_scene()
    sums = oracle.reprojection_error_sums(s)
    fx, fy, cx, cy, k1, k2, p1, p2, k3 = s.intr
    Rm = synth.quat_to_rot(s.tag_q)[s.to_tag]
    Rc = synth.quat_to_rot(s.cam_q)[s.to_cam]
    ref = np.zeros(s.n_tag_obs)
    for c in range(4):
        p = np.zeros((s.n_tag_obs, 3))
        p[:, 0] = synth.CORNER_SIGNS[c, 0] * s.tag_wh[s.to_tag, 0] / 2
        p[:, 1] = synth.CORNER_SIGNS[c, 1] * s.tag_wh[s.to_tag, 1] / 2
        pw = np.einsum("nij,nj->ni", Rm, p) + s.tag_t[s.to_tag]
        pc = np.einsum("nij,nj->ni", Rc, pw) + s.cam_t[s.to_cam]
        x, y = pc[:, 0] / pc[:, 2], pc[:, 1] / pc[:, 2]
        r2 = x * x + y * y
        rad = 1 + r2 * (k1 + r2 * (k2 + r2 * k3))
        xd = x * rad + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
        yd = y * rad + 2 * p2 * xd * y + p1 * (r2 + 2 * y * y)      # quirk: distorted x
        ref += np.hypot(fx * xd + cx - s.to_uv[:, 2 * c], fy * yd + cy - s.to_uv[:, 2 * c + 1])
    assert np.abs(sums - ref).max() < 1e-9


# ---- covariance restatement (src/TagReconstructor.cpp:744-783) and scene tiling ---------------------------
def test_oracle_covariance_equals_schur_form_and_scales_with_noise(oracle):
    """The oracle inverts the full tangent-space normal matrix; the engine eliminates the cameras first.  Both are the
    same blocks of the same inverse: check the identity in numpy, independent of any GPU code."""
    from visual_marker_mapping_b200 import synth

    s = synth.make_scene(n_markers=10, n_cameras=24, markers_per_camera=6, seed=5)
    cov = oracle.marker_translation_covariance(s)
    assert cov.shape == (s.n_markers, 3, 3) and not cov[s.fixed_tag].any()
    ev = oracle.evaluate(s, oracle.LOSS_NONE, 1.0, residuals=False, gradient=False)
    J = ev.jac_local.reshape(-1, 8, 12)
    Nc, Nm = s.n_cameras, s.n_markers
    B = np.zeros((Nc, 6, 6)); C = np.zeros((Nm, 6, 6))
    np.add.at(B, s.to_cam, np.einsum("kra,krb->kab", J[:, :, :6], J[:, :, :6]))
    np.add.at(C, s.to_tag, np.einsum("kra,krb->kab", J[:, :, 6:], J[:, :, 6:]))
    E = np.einsum("kra,krb->kab", J[:, :, :6], J[:, :, 6:])
    R = np.zeros((6 * Nm, 6 * Nm))
    for j in range(Nm):
        R[6 * j:6 * j + 6, 6 * j:6 * j + 6] = C[j]
    Binv = np.linalg.inv(B)
    for i in range(Nc):
        obs = np.nonzero(s.to_cam == i)[0]
        for a in obs:
            for b in obs:
                ja, jb = s.to_tag[a], s.to_tag[b]
                R[6 * ja:6 * ja + 6, 6 * jb:6 * jb + 6] -= E[a].T @ Binv[i] @ E[b]
    keep = np.ones(6 * Nm, dtype=bool)
    keep[6 * s.fixed_tag:6 * s.fixed_tag + 6] = False
    Rinv = np.linalg.inv(R[np.ix_(keep, keep)])
    idx = -np.ones(6 * Nm, dtype=int)
    idx[keep] = np.arange(keep.sum())
    for j in range(Nm):
        if j == s.fixed_tag:
            continue
        p = idx[6 * j:6 * j + 3]
        blk = Rinv[np.ix_(p, p)]
        assert np.abs(blk - cov[j]).max() <= 1e-9 * np.abs(cov[j]).max()
        assert np.all(np.linalg.eigvalsh(cov[j]) > 0)
    # markers far from the fixed origin tag are less certain than its neighbours
    d = np.linalg.norm(s.tag_t_gt - s.tag_t_gt[s.fixed_tag], axis=1)
    tr = np.trace(cov, axis1=1, axis2=2)
    far, near = np.argsort(d)[-3:], np.argsort(d)[1:4]
    assert tr[far].mean() > tr[near].mean()


def test_tile_scene_replicas_are_disjoint_and_evaluate_alike(oracle):
    from visual_marker_mapping_b200 import synth

    s = synth.make_scene(n_markers=8, n_cameras=10, markers_per_camera=5, seed=2)
    t = synth.tile_scene(s, 3)
    assert (t.n_cameras, t.n_markers, t.n_tag_obs) == (3 * s.n_cameras, 3 * s.n_markers, 3 * s.n_tag_obs)
    assert t.to_cam.dtype == np.int32 and t.to_cam.max() == t.n_cameras - 1 and t.to_tag.max() == t.n_markers - 1
    # replica r only refers to its own cameras and markers
    for r in range(3):
        sl = slice(r * s.n_tag_obs, (r + 1) * s.n_tag_obs)
        assert np.array_equal(t.to_cam[sl], s.to_cam + r * s.n_cameras) and np.array_equal(t.to_tag[sl], s.to_tag + r * s.n_markers)
    e1 = oracle.evaluate(s, oracle.LOSS_HUBER, 1.0, gradient=False)
    e3 = oracle.evaluate(t, oracle.LOSS_HUBER, 1.0, gradient=False)
    assert abs(e3.cost - 3 * e1.cost) <= 1e-12 * e3.cost
    assert np.array_equal(e3.residuals[:e1.residuals.shape[0]], e1.residuals)


# ---- the sparse exact solve (block-envelope Cholesky under a reverse Cuthill-McKee marker order) -------------
def test_envelope_and_dense_solvers_take_the_same_trajectory(oracle, scene_c2):
    """Two factorisations of the same marker-reduced system (dense, round 1; block envelope in RCM order): every
    iteration's cost, the iteration counts and the final poses agree to round-off."""
    a = oracle.solve(scene_c2, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0, linear_solver=oracle.SOLVER_DENSE)
    b = oracle.solve(scene_c2, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0, linear_solver=oracle.SOLVER_ENVELOPE)
    assert a.summary.num_iterations == b.summary.num_iterations
    assert np.abs(a.trace[:, 0] - b.trace[:, 0]).max() <= 1e-12 * a.trace[0, 0]
    assert np.abs(a.cam_t - b.cam_t).max() < 1e-11 and np.abs(a.tag_t - b.tag_t).max() < 1e-11


def test_envelope_is_sparse_on_a_site_and_dense_on_a_room(oracle, scene_c2):
    site = synth.make_config("c3/8")
    st = oracle.sparse_structure(site)
    dense_blocks = (st["order"] // 96) * (st["order"] // 96 + 1) // 2
    assert st["blocks"] < 0.35 * dense_blocks          # local co-visibility: the envelope stays narrow (C3: 0.096)
    room = oracle.sparse_structure(scene_c2)
    assert room["free_markers"] > 400


def test_lm_step_residual_vanishes_for_the_oracles_own_step(oracle, scene_c1):
    """The matrix-free normal-equation check used on the GPU step at C3/C4 scale, applied to a step this file knows:
    one LM iteration of the oracle itself.  delta is recovered from the poses (inverse of Plus)."""
    s = scene_c1
    r = oracle.solve(s, max_num_iterations=1, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    Nc, Nm = s.n_cameras, s.n_markers

    def log_q(q_new, q_old):
        # q_delta = q_new * conj(q_old) = (cos n, sin(n)/n d)
        w0, x0, y0, z0 = q_old.T
        c = np.stack([w0, -x0, -y0, -z0], axis=1) / np.sum(q_old * q_old, axis=1, keepdims=True)
        a, b = q_new, c
        w = a[:, 0] * b[:, 0] - a[:, 1] * b[:, 1] - a[:, 2] * b[:, 2] - a[:, 3] * b[:, 3]
        v = np.stack([a[:, 0] * b[:, 1] + a[:, 1] * b[:, 0] + a[:, 2] * b[:, 3] - a[:, 3] * b[:, 2],
                      a[:, 0] * b[:, 2] - a[:, 1] * b[:, 3] + a[:, 2] * b[:, 0] + a[:, 3] * b[:, 1],
                      a[:, 0] * b[:, 3] + a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1] + a[:, 3] * b[:, 0]], axis=1)
        n = np.linalg.norm(v, axis=1)
        ang = np.arctan2(n, w)
        return v * np.where(n > 0, ang / np.maximum(n, 1e-300), 1.0)[:, None]

    delta = np.zeros((Nc + Nm, 6))
    delta[:Nc, :3] = r.cam_t - s.cam_t
    delta[:Nc, 3:] = log_q(r.cam_q, s.cam_q)
    delta[Nc:, :3] = r.tag_t - s.tag_t
    delta[Nc:, 3:] = log_q(r.tag_q, s.tag_q)
    chk = oracle.lm_step_residual(s, delta.reshape(-1), radius=1e4, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    assert chk["relative"] < 1e-9, chk
    wrong = oracle.lm_step_residual(s, 1.01 * delta.reshape(-1), radius=1e4, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    assert wrong["relative"] > 1e-3           # the check does notice a 1 % error


def test_sparse_covariance_equals_the_dense_inverse(oracle):
    s = synth.make_scene(n_markers=40, n_cameras=90, markers_per_camera=8, seed=21)
    ref = oracle.marker_translation_covariance(s)
    sel = np.arange(s.n_markers)
    cov = oracle.marker_translation_covariance_sparse(s, sel)
    assert np.abs(cov - ref).max() <= 1e-8 * np.abs(ref).max()
