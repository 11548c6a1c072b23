"""Product mode against parity mode on one scene: costs, iteration counts, pose differences, time per solve.
    python tools/diag_convergence.py c3 [huber|none|cauchy] [scale]
One JSON line per run on stdout (kept under profiles/ as evidence for DESIGN.md §2)."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
loss = {"huber": pkg.LOSS_HUBER, "none": pkg.LOSS_NONE, "cauchy": pkg.LOSS_CAUCHY}[sys.argv[2] if len(sys.argv) > 2 else "huber"]
a = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
s = synth.make_config(name)
init = (s.cam_q, s.cam_t, s.tag_q, s.tag_t)
runs = [("parity ftol=1e-12", dict(parity_mode=1, function_tolerance=1e-12)),
        ("parity default", dict(parity_mode=1)),
        ("product default", dict()),
        ("product confirm=1e-4", dict(pcg_confirm_tolerance=1e-4)),
        ("product no-confirm", dict(pcg_confirm_tolerance=0.0)),
        ("polish 2e-2 below 1e-2", dict(pcg_polish_tolerance=2e-2)),
        ("polish 5e-2 below 1e-2", dict(pcg_polish_tolerance=5e-2)),
        ("polish 5e-2 below 1e-3", dict(pcg_polish_tolerance=5e-2, pcg_polish_threshold=1e-3)),
        ("polish 1e-1 below 1e-2", dict(pcg_polish_tolerance=1e-1)),
        ("early 3e-3, polish 5e-2", dict(pcg_tolerance=3e-3, pcg_polish_tolerance=5e-2)),
        ("product pcg=5e-2", dict(pcg_tolerance=5e-2)),
        ("product ftol=1e-12", dict(function_tolerance=1e-12))]
ref = None
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    for label, kw in runs:
        ba.set_poses(*init)
        t0 = time.time()
        summ = ba.solve(max_num_iterations=200, loss_kind=loss, loss_scale=a, **kw)
        dt = time.time() - t0
        p = ba.get_poses()
        if ref is None:
            ref = (summ.final_cost, p)
        et = max(np.abs(p[1] - ref[1][1]).max(), np.abs(p[3] - ref[1][3]).max())
        er = max(synth.quat_angle(p[0], ref[1][0]).max(), synth.quat_angle(p[2], ref[1][2]).max())
        print(json.dumps({"scene": name, "loss": int(loss), "run": label, "final_cost": summ.final_cost, "rel_to_tight": (summ.final_cost - ref[0]) / ref[0],
                          "lm_iterations": summ.num_iterations, "pcg_iterations": int(summ.pcg_iterations), "confirmations": summ.num_confirmations,
                          "seconds": round(dt, 4), "message": summ.message.decode(), "max_trans_diff_m": float(et), "max_rot_diff_rad": float(er)}), flush=True)
