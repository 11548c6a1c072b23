"""GPU parity tests: the CUDA path through the C ABI (include/vmm_ba.h) against the CPU oracle.

Tolerances are BASELINE.json's: residuals / Jacobians 1e-10 relative (per-block Frobenius norm,
denominator max(|ref block|, 1)), final cost 1e-6 relative, translations 1e-6 m, rotations 1e-6 rad.
"""
import numpy as np
import pytest

import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

pytestmark = pytest.mark.gpu

RJ_TOL = 1e-10
COST_TOL = 1e-6
TRANS_TOL = 1e-6
ROT_TOL = 1e-6
LOSSES = [(pkg.LOSS_NONE, 1.0), (pkg.LOSS_HUBER, 1.0), (pkg.LOSS_CAUCHY, 3.0)]


def block_rel_err(a, ref):
    a = a.reshape(ref.shape[0], -1)
    r = ref.reshape(ref.shape[0], -1)
    den = np.maximum(np.linalg.norm(r, axis=1), 1.0)
    return float((np.linalg.norm(a - r, axis=1) / den).max())


def check_eval(scene, oracle, loss, a):
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene)
        ev = ba.evaluate(loss, a)
    ref = oracle.evaluate(scene, loss, a)
    assert block_rel_err(ev.residuals, ref.residuals) < RJ_TOL
    assert block_rel_err(ev.jacobians, ref.jac_local) < RJ_TOL
    assert abs(ev.cost - ref.cost) <= 1e-12 * max(ref.cost, 1.0)
    gden = max(np.abs(ref.gradient).max(), 1.0)
    assert np.abs(ev.gradient - ref.gradient).max() / gden < 1e-11
    return ev, ref


@pytest.mark.parametrize("loss,a", LOSSES)
def test_evaluate_c1(scene_c1, oracle, loss, a):
    check_eval(scene_c1, oracle, loss, a)


@pytest.mark.parametrize("loss,a", LOSSES)
def test_evaluate_c2(scene_c2, oracle, loss, a):
    check_eval(scene_c2, oracle, loss, a)


def test_evaluate_outliers(oracle):
    s = synth.make_scene(n_markers=60, n_cameras=150, markers_per_camera=10, seed=77, outlier_frac=0.2)
    for loss, a in LOSSES:
        check_eval(s, oracle, loss, a)


def test_evaluate_unnormalised_quaternions(scene_c1, oracle):
    """ceres::QuaternionRotatePoint normalises: the residual is invariant to |q| and the local
    Jacobian has the same closed form."""
    s = scene_c1.copy()
    rng = np.random.default_rng(5)
    s.cam_q *= rng.uniform(0.5, 2.0, (s.n_cameras, 1))
    s.tag_q *= rng.uniform(0.5, 2.0, (s.n_markers, 1))
    check_eval(s, oracle, pkg.LOSS_HUBER, 1.0)


def test_evaluate_shuffled_and_duplicate_observations(scene_c1, oracle):
    """Caller order is arbitrary and (image, tag) is not a unique key (MIT detector):
    outputs come back in the caller's order."""
    s = scene_c1.copy()
    rng = np.random.default_rng(11)
    dup = rng.integers(0, s.n_tag_obs, 25)
    s.to_cam = np.concatenate([s.to_cam, s.to_cam[dup]])
    s.to_tag = np.concatenate([s.to_tag, s.to_tag[dup]])
    s.to_uv = np.concatenate([s.to_uv, s.to_uv[dup] + 0.25])
    perm = rng.permutation(s.n_tag_obs)
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[perm].copy(), s.to_tag[perm].copy(), s.to_uv[perm].copy()
    check_eval(s, oracle, pkg.LOSS_HUBER, 1.0)


def test_blocks_without_observations(oracle):
    """Cameras whose tags were all culled and tags nobody sees keep their pose and get no columns
    (src/TagReconstructor.cpp:686-690)."""
    s = synth.make_scene(n_markers=20, n_cameras=30, markers_per_camera=8, seed=3)
    keep = (s.to_cam != 4) & (s.to_cam != 29) & (s.to_tag != 7)
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[keep].copy(), s.to_tag[keep].copy(), s.to_uv[keep].copy()
    ev, ref = check_eval(s, oracle, pkg.LOSS_NONE, 1.0)
    assert np.all(ev.gradient[6 * 4:6 * 5] == 0) and np.all(ev.gradient[6 * 29:6 * 30] == 0)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        summ = ba.solve(max_num_iterations=30, parity_mode=1)
        cq, ct, tq, tt = ba.get_poses()
    o = oracle.solve(s, max_num_iterations=30)
    assert abs(summ.final_cost - o.summary.final_cost) <= COST_TOL * o.summary.final_cost
    for i in (4, 29):
        assert np.array_equal(cq[i], s.cam_q[i]) and np.array_equal(ct[i], s.cam_t[i])
    assert np.array_equal(tq[7], s.tag_q[7]) and np.array_equal(tt[7], s.tag_t[7])
    assert np.array_equal(tq[0], s.tag_q[0]) and np.array_equal(tt[0], s.tag_t[0])   # fixed origin tag


def test_empty_observations():
    s = synth.make_scene(n_markers=5, n_cameras=4, markers_per_camera=3, seed=9)
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[:0].copy(), s.to_tag[:0].copy(), s.to_uv[:0].copy()
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        ev = ba.evaluate()
        assert ev.cost == 0.0 and ev.residuals.shape == (0, 2) and not ev.gradient.any()
        summ = ba.solve(max_num_iterations=5)
        assert summ.termination_type == pkg.CONVERGENCE and summ.final_cost == 0.0
        cq, ct, tq, tt = ba.get_poses()
    assert np.array_equal(cq, s.cam_q) and np.array_equal(tt, s.tag_t)


def test_reprojection_error_sums(scene_c2, oracle):
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene_c2)
        got = ba.reprojection_error_sums()
    ref = oracle.reprojection_error_sums(scene_c2)
    assert np.abs(got - ref).max() <= 1e-10 * max(ref.max(), 1.0)


def pose_errors(cq, ct, tq, tt, o):
    return (np.abs(ct - o.cam_t).max(), np.abs(tt - o.tag_t).max(),
            synth.quat_angle(cq, o.cam_q).max(), synth.quat_angle(tq, o.tag_q).max())


@pytest.mark.parametrize("loss,a", LOSSES)
def test_solve_parity_c1(scene_c1, oracle, loss, a):
    """Parity mode (PCG to 1e-13) tracks the exact-solve trajectory step for step."""
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene_c1)
        summ = ba.solve(max_num_iterations=50, loss_kind=loss, loss_scale=a, parity_mode=1)
        poses = ba.get_poses()
    o = oracle.solve(scene_c1, max_num_iterations=50, loss_kind=loss, loss_scale=a)
    assert summ.termination_type == o.summary.termination_type
    assert summ.num_iterations == o.summary.num_iterations - 1
    assert summ.num_successful_steps == o.summary.num_successful_steps
    assert abs(summ.initial_cost - o.summary.initial_cost) <= 1e-12 * o.summary.initial_cost
    assert abs(summ.final_cost - o.summary.final_cost) <= COST_TOL * o.summary.final_cost
    ect, ett, ecq, etq = pose_errors(*poses, o)
    assert ect < TRANS_TOL and ett < TRANS_TOL and ecq < ROT_TOL and etq < ROT_TOL


def test_solve_parity_c2_huber(scene_c2, oracle):
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene_c2)
        summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)
        poses = ba.get_poses()
    o = oracle.solve(scene_c2, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    assert summ.num_iterations == o.summary.num_iterations - 1
    assert abs(summ.final_cost - o.summary.final_cost) <= COST_TOL * o.summary.final_cost
    ect, ett, ecq, etq = pose_errors(*poses, o)
    assert ect < TRANS_TOL and ett < TRANS_TOL and ecq < ROT_TOL and etq < ROT_TOL


def test_solve_inexact_reaches_reference_cost(scene_c2, oracle):
    """Product mode (inexact Newton, loose PCG) run to a tight function tolerance lands on the same
    minimum as the exact-solve oracle run to the same tolerance: cost 1e-6, poses 1e-6."""
    kw = dict(max_num_iterations=100, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, function_tolerance=1e-14,
              parameter_tolerance=1e-12)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene_c2)
        summ = ba.solve(**kw)
        poses = ba.get_poses()
    o = oracle.solve(scene_c2, **kw)
    assert abs(summ.final_cost - o.summary.final_cost) <= COST_TOL * o.summary.final_cost
    ect, ett, ecq, etq = pose_errors(*poses, o)
    assert ect < TRANS_TOL and ett < TRANS_TOL and ecq < ROT_TOL and etq < ROT_TOL


def test_zero_noise_scene_returns_to_ground_truth():
    s = synth.make_scene(n_markers=40, n_cameras=120, markers_per_camera=10, seed=21, noise_px=0.0, float32_round=False)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        summ = ba.solve(max_num_iterations=100, function_tolerance=1e-16, parameter_tolerance=1e-14, parity_mode=1)
        cq, ct, tq, tt = ba.get_poses()
    assert summ.final_cost < 1e-12
    assert np.abs(tt - s.tag_t_gt).max() < 1e-8 and np.abs(ct - s.cam_t_gt).max() < 1e-8
    assert synth.quat_angle(tq, s.tag_q_gt).max() < 1e-8


def test_properties_large_scene():
    """Size-independent properties on a scene too large for the oracle's dense solve (~1M corners):
    the fused Schur builder's gradient equals J^T r of the materialised pass, the cost agrees
    between the materialising and the cost-only kernels, and the solve decreases the cost."""
    s = synth.make_scene(n_markers=2500, n_cameras=12500, markers_per_camera=20, seed=1234)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        ev = ba.evaluate(pkg.LOSS_HUBER, 1.0)
        J, r = ev.jacobians, ev.residuals
        g = np.zeros_like(ev.gradient)
        cam = np.repeat(s.to_cam, 4)
        tag = np.repeat(s.to_tag, 4)
        gc = np.einsum("nrk,nr->nk", J[:, :, :6], r)
        gm = np.einsum("nrk,nr->nk", J[:, :, 6:], r)
        np.add.at(g.reshape(-1, 6), cam, gc)
        gm[tag == s.fixed_tag] = 0.0
        np.add.at(g.reshape(-1, 6), s.n_cameras + tag, gm)
        assert np.abs(g - ev.gradient).max() <= 1e-10 * np.abs(g).max()
        summ = ba.solve(max_num_iterations=25, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        assert abs(summ.initial_cost - ev.cost) <= 1e-12 * ev.cost
        assert summ.final_cost < 1e-3 * summ.initial_cost
        ev2 = ba.evaluate(pkg.LOSS_HUBER, 1.0, jacobians=False, gradient=False)
        assert abs(ev2.cost - summ.final_cost) <= 1e-12 * summ.final_cost
        assert summ.kernel_launches > 0


def _reference_schur_product(s, oracle, x, loss, a, radius):
    """S x from the oracle's (loss-corrected) local Jacobian in numpy: the normal-equation blocks of
    src/TagReconstructor.cpp:699-724, LM damping as in SURVEY A.5, marker blocks eliminated."""
    ev = oracle.evaluate(s, loss, a, gradient=False)
    J = ev.jac_local.reshape(-1, 8, 12)
    Jc, Jm = J[:, :, :6], J[:, :, 6:]
    Nc, Nm = s.n_cameras, s.n_markers
    B = np.zeros((Nc, 6, 6)); C = np.zeros((Nm, 6, 6))
    np.add.at(B, s.to_cam, np.einsum("kra,krb->kab", Jc, Jc))
    np.add.at(C, s.to_tag, np.einsum("kra,krb->kab", Jm, Jm))
    E = np.einsum("kra,krb->kab", Jc, Jm)
    hc, hm = np.einsum("kaa->ka", B), np.einsum("kaa->ka", C)
    sc, sm = 1 / (1 + np.sqrt(hc)), 1 / (1 + np.sqrt(hm))
    lam_c = np.clip(hc * sc * sc, 1e-6, 1e32) / (radius * sc * sc)
    lam_m = np.clip(hm * sm * sm, 1e-6, 1e32) / (radius * sm * sm)
    active = np.bincount(s.to_tag, minlength=Nm) > 0
    active[s.fixed_tag] = False
    Cinv = np.zeros_like(C)
    Cinv[active] = np.linalg.inv(C[active] + np.einsum("ka,ab->kab", lam_m[active], np.eye(6)))
    xb = x.reshape(Nc, 6)
    w = np.zeros((Nm, 6))
    np.add.at(w, s.to_tag, np.einsum("kab,ka->kb", E, xb[s.to_cam]))
    z = np.einsum("jab,jb->ja", Cinv, w)
    y = np.einsum("iab,ib->ia", B, xb) + lam_c * xb
    np.subtract.at(y, s.to_cam, np.einsum("kab,kb->ka", E, z[s.to_tag]))
    return y.reshape(-1)


SCHUR_SCENES = {
    # markers_per_camera == n_markers: every camera sees (nearly) every marker
    "registers": dict(n_markers=40, n_cameras=150, markers_per_camera=12),      # < 224 observations per marker
    "ring_slot": dict(n_markers=24, n_cameras=420, markers_per_camera=24),      # > one register segment per marker
    "split": dict(n_markers=12, n_cameras=900, markers_per_camera=12),          # > kSegMaxObs: gather / solve / scatter
}


@pytest.mark.parametrize("name", list(SCHUR_SCENES))
@pytest.mark.parametrize("path", [0, 1])
def test_schur_product_matches_reference(oracle, name, path):
    """The streamed Schur product (TMA ring, both consumer paths, split markers) against the product
    formed explicitly from the oracle's Jacobian: 1e-10 relative."""
    s = synth.make_scene(seed=77, **SCHUR_SCENES[name])
    counts = np.bincount(s.to_tag, minlength=s.n_markers)
    if name == "ring_slot":
        assert counts.max() > 256
    if name == "split":
        assert counts.max() > 512
    rng = np.random.default_rng(5)
    x = rng.normal(size=6 * s.n_cameras)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        for loss, a in LOSSES:
            y = ba.schur_multiply(x, loss, a, radius=1e4, path=path)
            ref = _reference_schur_product(s, oracle, x, loss, a, 1e4)
            assert np.abs(y - ref).max() <= 1e-10 * np.abs(ref).max(), (name, path, loss)


# ---- marker-translation covariance (src/TagReconstructor.cpp:744-783) ----------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("loss_kind,loss_scale", [(pkg.LOSS_NONE, 1.0), (pkg.LOSS_HUBER, 1.0)])
def test_marker_translation_covariance_matches_dense_inverse(oracle, loss_kind, loss_scale):
    """Engine: cameras eliminated, dense marker system, Gram matrices of L^-1.  Oracle: the full tangent-space
    normal matrix inverted densely.  Relative 1e-8 on every 3x3 block (the two routes differ in conditioning)."""
    s = synth.make_scene(n_markers=40, n_cameras=90, markers_per_camera=8, seed=21, outlier_frac=0.1)
    ref = oracle.marker_translation_covariance(s, loss_kind, loss_scale)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        cov = ba.marker_translation_covariance(loss_kind, loss_scale)
    assert cov.shape == ref.shape
    assert not cov[s.fixed_tag].any()                         # constant origin tag: zero block, as in Ceres
    scale = np.abs(ref).reshape(len(ref), -1).max(1)
    ok = scale > 0
    observed = np.bincount(s.to_tag, minlength=s.n_markers) > 0
    observed[s.fixed_tag] = False
    assert np.array_equal(ok, observed)                       # unobserved markers: zero blocks on both sides
    err = np.abs(cov - ref).reshape(len(ref), -1).max(1)
    assert (err[ok] <= 1e-8 * scale[ok]).all(), (err[ok] / scale[ok]).max()
    assert np.allclose(cov, np.transpose(cov, (0, 2, 1)))


@pytest.mark.gpu
def test_marker_translation_covariance_c1_and_limits(oracle):
    s = synth.make_config("c1")
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_NONE, parity_mode=1)
        cq, ct, tq, tt = ba.get_poses()
        cov = ba.marker_translation_covariance()
    s.cam_q, s.cam_t, s.tag_q, s.tag_t = cq, ct, tq, tt
    ref = oracle.marker_translation_covariance(s)
    assert np.abs(cov - ref).max() <= 1e-8 * np.abs(ref).max()
    # the dense route is refused beyond its documented limit instead of allocating (6 Nm)^2 doubles
    big = synth.make_scene(n_markers=4200, n_cameras=20, markers_per_camera=10, seed=3)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(big)
        with pytest.raises(pkg.VmmBaError, match="limited to 4096 markers"):
            ba.marker_translation_covariance(method=pkg.COV_DENSE)


def test_block_tridiagonal_covariance_equals_the_dense_route():
    """The site-scale route (breadth-first levels -> block-tridiagonal two-sweep recursion) against the dense inversion
    of the same marker-reduced matrix on a scene with several levels: every 3x3 block to 1e-8 relative."""
    s = synth.make_config("c3/16")           # 1250 markers, a site: ~8 levels
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        ba.solve(max_num_iterations=30, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        dense = ba.marker_translation_covariance(method=pkg.COV_DENSE)
        block = ba.marker_translation_covariance(method=pkg.COV_BLOCK_TRIDIAGONAL)
    scale = np.abs(dense).reshape(len(dense), -1).max(1)
    ok = scale > 0
    assert ok.sum() >= s.n_markers - 2
    err = np.abs(block - dense).reshape(len(dense), -1).max(1)
    assert (err[ok] <= 1e-8 * scale[ok]).all(), (err[ok] / scale[ok]).max()
    assert not block[~ok].any()


# ---- coarse level of the two-level preconditioner -------------------------------------------------------------
@pytest.mark.parametrize("cams_per_cluster", [100, 192, 400])
def test_coarse_inverse_matches_numpy(cams_per_cluster):
    """The hand-written in-place inversion of A_c (blocked sweep operator, csrc/ba_coarse_inv.cuh) against
    numpy.linalg.inv of the matrix the device assembled, for orders that are and are not multiples of its block size."""
    s = synth.make_config("c3/16")
    with pkg.BundleAdjuster(0) as ba:
        ba.set_coarse_cluster_size(cams_per_cluster)
        ba.set_scene(s)
        a, inv, n = ba.debug_coarse_matrix(pkg.LOSS_HUBER, 1.0, 1e4)
    assert n > 0 and n % 6 == 0, getattr(ba, "coarse_unavailable_reason", "")
    low = np.tril(a)
    full = low + np.tril(a, -1).T                       # the kernel reads the lower triangle only
    assert np.allclose(full[:n, :n], a[:n, :n], rtol=1e-9, atol=1e-9 * np.abs(a).max())    # the assembly is symmetric
    ref = np.linalg.inv(full)
    assert np.abs(inv - ref).max() <= 1e-8 * np.abs(ref).max()
    assert np.allclose(inv, inv.T, rtol=0, atol=1e-12 * np.abs(inv).max())
    pad = a.shape[0] - n
    if pad:
        assert np.array_equal(inv[n:, n:], np.eye(pad)) and not inv[n:, :n].any()


@pytest.mark.parametrize("shuffle", [False, True])
def test_set_problem_with_unchanged_structure_refreshes_observations_and_poses(scene_c2, oracle, shuffle):
    """A second vmm_ba_set_problem with the same sizes and index arrays keeps the plans and copies only observations,
    marker sizes and poses (straight from a camera-sorted input, through the cached permutation otherwise): the
    evaluation must be that of the NEW data, and a changed index array must rebuild."""
    s = scene_c2.copy()
    if shuffle:
        perm = np.random.default_rng(3).permutation(s.n_tag_obs)
        s.to_cam, s.to_tag, s.to_uv = s.to_cam[perm], s.to_tag[perm], s.to_uv[perm]
    rng = np.random.default_rng(4)
    s2 = s.copy()
    s2.to_uv = s.to_uv + rng.normal(scale=0.7, size=s.to_uv.shape)
    s2.cam_t = s.cam_t + rng.normal(scale=0.01, size=s.cam_t.shape)
    s2.tag_t = s.tag_t + rng.normal(scale=0.01, size=s.tag_t.shape)
    s2.tag_wh = s.tag_wh * 1.01
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        b0 = ba.h2d_bytes()
        ba.set_scene(s2)                                   # same structure: refresh
        assert ba.h2d_bytes() - b0 < s.n_tag_obs * 72      # no index arrays travelled
        ev = ba.evaluate(pkg.LOSS_HUBER, 1.0)
        ref = oracle.evaluate(s2, oracle.LOSS_HUBER, 1.0)
        assert block_rel_err(ev.residuals, ref.residuals) < RJ_TOL and block_rel_err(ev.jacobians, ref.jac_local) < RJ_TOL
        assert np.abs(ev.gradient - ref.gradient).max() / max(np.abs(ref.gradient).max(), 1.0) < 1e-11
        summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)
        oref = oracle.solve(s2, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
        assert abs(summ.final_cost - oref.summary.final_cost) <= COST_TOL * oref.summary.final_cost
        s3 = s2.copy()
        s3.to_tag = s2.to_tag.copy()
        s3.to_tag[5] = (s3.to_tag[5] + 1) % s3.n_markers   # one index differs: everything is rebuilt
        ba.set_scene(s3)
        ev3 = ba.evaluate(pkg.LOSS_NONE, 1.0, jacobians=False, gradient=False)
        assert abs(ev3.cost - oracle.evaluate(s3, oracle.LOSS_NONE, 1.0, jac_local=False, gradient=False).cost) <= 1e-12 * ev3.cost


def test_solve_with_split_markers_matches_oracle(oracle):
    """Markers with more observations than one segment of the streamed product holds go through the boundary path
    (gather / local sum / scatter inside the persistent kernel): the solve must still be the oracle's."""
    s = synth.make_scene(seed=77, **SCHUR_SCENES["split"])
    assert np.bincount(s.to_tag, minlength=s.n_markers).max() > 512
    oref = oracle.solve(s, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)
        cq, ct, tq, tt = ba.get_poses()
    assert abs(summ.final_cost - oref.summary.final_cost) <= COST_TOL * oref.summary.final_cost
    assert summ.num_iterations + 1 == oref.summary.num_iterations
    assert np.abs(ct - oref.cam_t).max() < TRANS_TOL and np.abs(tt - oref.tag_t).max() < TRANS_TOL
