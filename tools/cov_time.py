"""Wall time of vmm_ba_marker_translation_covariance on a solved scene (site scale: the block-tridiagonal route).
    python tools/cov_time.py [c3]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
s = synth.make_config(name)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
    for rep in range(2):
        t0 = time.perf_counter()
        cov = ba.marker_translation_covariance(loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        dt = time.perf_counter() - t0
        cov = np.asarray(cov[0] if isinstance(cov, tuple) else cov)
        print(json.dumps({"scene": name, "markers": s.n_markers, "rep": rep, "seconds": round(dt, 3), "final_cost": summ.final_cost,
                          "median_sigma_m": float(np.sqrt(np.median(np.abs(cov.reshape(-1, 9)[:, [0, 4, 8]]))))}), flush=True)
