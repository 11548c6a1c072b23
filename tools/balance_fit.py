"""Fits the load-balance model of the streamed product: per-CTA time of the product phase against the segments, ring slots and
ring-resident tails the CTA was given (the partition of upload_seglist, restated here).  python tools/balance_fit.py [c3]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

REG_OBS = 224   # kSegRegObs of the default build
s = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c3")
fixed, tail = int(os.environ.get("VMM_BA_SEG_FIXED", "3")), int(os.environ.get("VMM_BA_SEG_TAIL", "0"))
G = 148
nobs = np.bincount(np.asarray(s.to_tag), minlength=s.n_markers)
act = (nobs > 0) & (np.arange(s.n_markers) != s.fixed_tag)
n = nobs[act]
slots = (((n + 7) // 8) + 3) // 4
has_tail = n > REG_OBS
cost = np.concatenate([[0], np.cumsum(slots + fixed + has_tail * tail)])
first = np.array([np.searchsorted(cost[:-1], cost[-1] * c // G, side="left") for c in range(G)] + [len(n)])
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0)
    ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
    ba.profile_enable(True)
    summ = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0)
    t = ba.pcg_cta_product_ns(G) / 1e3 / max(int(summ.pcg_iterations), 1)
rows = []
for c in range(G):
    a, b = first[c], first[c + 1]
    rows.append([int(b - a), int(slots[a:b].sum()), int(has_tail[a:b].sum()), int((slots[a:b] * has_tail[a:b]).sum())])
X = np.array(rows, dtype=float)
for cols, names in (((0, 1), ("segment", "slot")), ((0, 1, 2), ("segment", "slot", "tail segment")), ((0, 1, 2, 3), ("segment", "slot", "tail segment", "slot of a tail segment"))):
    A = X[:, cols]
    coef, *_ = np.linalg.lstsq(A, t, rcond=None)
    res = t - A @ coef
    print(json.dumps({"model": names, "us": [round(float(v), 4) for v in coef], "rms_residual_us": round(float(np.sqrt((res ** 2).mean())), 2),
                      "in_slots": [round(float(v / coef[1]), 2) for v in coef]}))
print(json.dumps({"cta_us_min_mean_max": [round(float(t.min()), 1), round(float(t.mean()), 1), round(float(t.max()), 1)], "fixed": fixed, "tail": tail,
                  "cta0": rows[0], "cta_mid": rows[74], "segments": int(len(n))}))
