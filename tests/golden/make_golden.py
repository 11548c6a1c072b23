"""Generates tests/golden/corner_kats.json — independent known-answer vectors for the reprojection
functor (TagReconstructionCostFunction.h:101-159) and its local Jacobian.

The reference has no tests or fixtures for this path, and Ceres is not available, so
the pins come from two independent statements of the same maths, both run in the build container:
  * cv2.projectPoints (OpenCV 4.13): projected pixel + d(pixel)/d(tvec), d(pixel)/d(rvec at 0)
  * mpmath at 50 digits: a literal restatement of the functor (quaternion normalise + rotate, pinhole,
    radial-tangential) and central differences through Ceres' QuaternionParameterization::Plus
    (half-angle, left multiplication) for the full 2x12 local Jacobian.
Run:  python tests/golden/make_golden.py     (needs cv2 + mpmath; neither is needed to run the tests)
"""
import json
import os

import cv2
import mpmath as mp
import numpy as np

mp.mp.dps = 50
HERE = os.path.dirname(os.path.abspath(__file__))

INTR = [8075.2937867635346, 8083.1676114192869, 3016.3896805084278, 1996.2896554785455,
        -0.1861818326266976, 0.37018092365577054, -2.9390604003594177e-4, 4.1533180829908799e-4,
        5.7043887874185996e-2]


def mp_rotate(q, p):
    n = mp.sqrt(sum(x * x for x in q))
    w, x, y, z = [c / n for c in q]
    R = [[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
         [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
         [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]]
    return [sum(R[i][j] * p[j] for j in range(3)) for i in range(3)]


def mp_functor(intr, cam_t, cam_q, tag_t, tag_q, corner, obs):
    fx, fy, cx, cy, k1, k2, p1, p2, k3 = [mp.mpf(v) for v in intr]
    pw = mp_rotate(tag_q, corner)
    pw = [pw[i] + tag_t[i] for i in range(3)]
    pc = mp_rotate(cam_q, pw)
    pc = [pc[i] + cam_t[i] for i in range(3)]
    x, y = pc[0] / pc[2], pc[1] / pc[2]
    r2 = x * x + y * y
    c = 1 + r2 * (k1 + r2 * (k2 + r2 * k3))
    xd = x * c + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
    yd = y * c + 2 * p2 * x * y + p1 * (r2 + 2 * y * y)
    return [fx * xd + cx - obs[0], fy * yd + cy - obs[1]]


def mp_plus(q, d):
    n = mp.sqrt(sum(x * x for x in d))
    if n == 0:
        return list(q)
    s = mp.sin(n) / n
    a = [mp.cos(n), s * d[0], s * d[1], s * d[2]]
    return [a[0] * q[0] - a[1] * q[1] - a[2] * q[2] - a[3] * q[3],
            a[0] * q[1] + a[1] * q[0] + a[2] * q[3] - a[3] * q[2],
            a[0] * q[2] - a[1] * q[3] + a[2] * q[0] + a[3] * q[1],
            a[0] * q[3] + a[1] * q[2] - a[2] * q[1] + a[3] * q[0]]


def mp_local_jacobian(intr, cam_t, cam_q, tag_t, tag_q, corner, obs):
    h = mp.mpf(10) ** -20
    J = [[None] * 12 for _ in range(2)]

    def f(delta):
        ct = [cam_t[i] + delta[i] for i in range(3)]
        cq = mp_plus(cam_q, delta[3:6])
        tt = [tag_t[i] + delta[6 + i] for i in range(3)]
        tq = mp_plus(tag_q, delta[9:12])
        return mp_functor(intr, ct, cq, tt, tq, corner, obs)

    for k in range(12):
        dp = [mp.mpf(0)] * 12
        dm = [mp.mpf(0)] * 12
        dp[k] = h
        dm[k] = -h
        fp, fm = f(dp), f(dm)
        for r in range(2):
            J[r][k] = (fp[r] - fm[r]) / (2 * h)
    return J


def main():
    rng = np.random.default_rng(20261104)
    cases = []
    K = np.array([[INTR[0], 0, INTR[2]], [0, INTR[1], INTR[3]], [0, 0, 1]])
    dist = np.array(INTR[4:9])
    for i in range(24):
        w, h = 0.1285, 0.1295
        ci = i % 4
        sx, sy = [(-1, -1), (1, -1), (1, 1), (-1, 1)][ci]
        corner = [sx * w / 2, sy * h / 2, 0.0]
        tag_q = rng.normal(size=4)
        tag_q /= np.linalg.norm(tag_q)
        cam_q = rng.normal(size=4)
        cam_q /= np.linalg.norm(cam_q)
        if i % 3 == 1:   # un-normalised quaternions: the functor normalises internally
            tag_q *= 1.0000003
            cam_q *= 0.73
        tag_t = rng.uniform(-2, 2, size=3)
        # choose cam_t so that the point lands in front of the camera inside the image
        pw = np.array([float(v) for v in mp_rotate([mp.mpf(float(c)) for c in tag_q], [mp.mpf(c) for c in corner])]) + tag_t
        pc0 = np.array([float(v) for v in mp_rotate([mp.mpf(float(c)) for c in cam_q], [mp.mpf(float(c)) for c in pw])])
        target = np.array([rng.uniform(-0.3, 0.3), rng.uniform(-0.2, 0.2), rng.uniform(1.0, 4.0)])
        target[:2] *= target[2]
        cam_t = target - pc0
        obs = rng.uniform(0, 4000, size=2)
        m = lambda v: [mp.mpf(float(c)) for c in v]
        res = mp_functor(INTR, m(cam_t), m(cam_q), m(tag_t), m(tag_q), m(corner), m(obs))
        J = mp_local_jacobian(INTR, m(cam_t), m(cam_q), m(tag_t), m(tag_q), m(corner), m(obs))
        # cv2: object point = R_c X_w (pre-rotated), rvec = 0, tvec = cam_t
        Xrot = pc0.reshape(1, 1, 3)
        proj, jac = cv2.projectPoints(Xrot, np.zeros(3), cam_t.reshape(3, 1), K, dist)
        cases.append({
            "corner_index": ci, "tag_wh": [w, h],
            "cam_t": cam_t.tolist(), "cam_q": cam_q.tolist(), "tag_t": tag_t.tolist(), "tag_q": tag_q.tolist(),
            "obs": obs.tolist(),
            "mp_residual": [mp.nstr(r, 25) for r in res],
            "mp_jac_local": [[mp.nstr(v, 25) for v in row] for row in J],
            "cv2_projection": proj.reshape(2).tolist(),
            "cv2_dp_dtvec": jac[:, 3:6].tolist(),
            "cv2_dp_drvec_at_0": jac[:, 0:3].tolist(),
        })
    # the survey's probe: X=(0.3,-0.2,2.5) -> (3982.03977278, 1351.89242311)
    proj, _ = cv2.projectPoints(np.array([[[0.3, -0.2, 2.5]]]), np.zeros(3), np.zeros(3), K, dist)
    out = {"intrinsics": INTR, "cases": cases, "probe_point": [0.3, -0.2, 2.5], "probe_cv2": proj.reshape(2).tolist(),
           "generator": "tests/golden/make_golden.py", "cv2": cv2.__version__, "mpmath_dps": mp.mp.dps}
    with open(os.path.join(HERE, "corner_kats.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(cases), "cases; probe", out["probe_cv2"])


if __name__ == "__main__":
    main()
