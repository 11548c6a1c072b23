"""Golden PnP cases from OpenCV (python cv2 4.13, present in the build container only):
cv2.solvePnP(flags=SOLVEPNP_ITERATIVE) — what solvePnPEigen of the reference calls (src/EigenCVConversions.cpp:38-67) —
on tag-corner configurations as the reference feeds them: one tag's four corners (planar), and the corners of several
reconstructed tags seen in one image (planar sheet, and tags at different heights/tilts).

    python tests/golden/make_pnp_golden.py      # rewrites tests/golden/pnp_cases.json
"""
import json, os
import numpy as np
import cv2

K = np.array([[8075.2937867635346, 0, 3016.3896805084278], [0, 8083.1676114192869, 1996.2896554785455], [0, 0, 1]])
DIST = np.array([-0.1861818326266976, 0.37018092365577054, -2.9390604003594177e-4, 4.1533180829908799e-4, 5.7043887874185996e-2])
W, H = 0.1285, 0.1295
LOCAL = np.array([[-W / 2, -H / 2, 0], [W / 2, -H / 2, 0], [W / 2, H / 2, 0], [-W / 2, H / 2, 0]])


def rot(rv):
    return cv2.Rodrigues(np.asarray(rv, dtype=np.float64))[0]


def main():
    rng = np.random.default_rng(20261104)
    cases = []
    for kind in ("single_tag",) * 6 + ("sheet",) * 4 + ("relief",) * 4:
        # camera pose (object -> camera): looking roughly down the object's -z, 0.6 .. 3 m away, tilted up to 35 degrees
        R_true = rot(rng.normal(0, 0.3, 3)) @ rot([np.pi, 0, 0])
        t_true = np.array([rng.uniform(-0.2, 0.2), rng.uniform(-0.15, 0.15), rng.uniform(0.6, 3.0)])
        if kind == "single_tag":
            X = LOCAL.copy()
        else:
            n_tags = int(rng.integers(3, 9))
            X = []
            for _ in range(n_tags):
                c = np.array([rng.uniform(-0.5, 0.5), rng.uniform(-0.35, 0.35), 0.0 if kind == "sheet" else rng.uniform(-0.15, 0.15)])
                Rt = rot([0, 0, rng.uniform(-np.pi, np.pi)]) if kind == "sheet" else rot(rng.normal(0, 0.25, 3))
                X.append(LOCAL @ Rt.T + c)
            X = np.concatenate(X)
        uv, _ = cv2.projectPoints(X, cv2.Rodrigues(R_true)[0], t_true, K, DIST)
        uv = uv.reshape(-1, 2) + rng.normal(0, 0.3, (len(X), 2))
        ok, rvec, tvec = cv2.solvePnP(X, uv, K, DIST, flags=cv2.SOLVEPNP_ITERATIVE)
        assert ok
        proj, _ = cv2.projectPoints(X, rvec, tvec, K, DIST)
        rms = float(np.sqrt(((proj.reshape(-1, 2) - uv) ** 2).sum(1).mean()))
        cases.append({"kind": kind, "object_points": X.tolist(), "image_points": uv.tolist(), "cv2_R": cv2.Rodrigues(rvec)[0].tolist(),
                      "cv2_t": tvec.reshape(3).tolist(), "cv2_rms_px": rms, "true_R": R_true.tolist(), "true_t": t_true.tolist()})
    out = {"generator": "tests/golden/make_pnp_golden.py", "cv2": cv2.__version__, "K": K.tolist(), "dist": DIST.tolist(), "cases": cases}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pnp_cases.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, len(cases), "cases; cv2 rms", [round(c["cv2_rms_px"], 3) for c in cases])


if __name__ == "__main__":
    main()
