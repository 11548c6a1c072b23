"""How early may the inexact solve switch to its looser forcing term?  Final cost against the converged minimum (parity mode,
function_tolerance 1e-12) and PCG work for pcg_polish_threshold in a range, on the scenes of profiles/r02_convergence.md.
    python tools/polish_sweep.py c3/8 [c3/4 ...]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

for name in sys.argv[1:] or ["c3/8"]:
    s = synth.make_config(name)
    init = (s.cam_q, s.cam_t, s.tag_q, s.tag_t)
    runs = [("parity ftol=1e-12", dict(parity_mode=1, function_tolerance=1e-12)), ("parity default", dict(parity_mode=1))]
    for thr in (1e-2, 5e-2, 0.2, 0.5, 0.9):
        runs.append((f"polish 5e-2 below {thr:g}", dict(pcg_polish_threshold=thr)))
    runs.append(("polish 2e-2 below 0.5", dict(pcg_polish_threshold=0.5, pcg_polish_tolerance=2e-2)))
    runs.append(("polish 3e-2 below 0.9", dict(pcg_polish_threshold=0.9, pcg_polish_tolerance=3e-2)))
    ref = None
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        for label, kw in runs:
            ba.set_poses(*init)
            t0 = time.time()
            summ = ba.solve(max_num_iterations=200, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, **kw)
            dt = time.time() - t0
            ref = ref if ref is not None else summ.final_cost
            print(json.dumps({"scene": name, "run": label, "rel_to_tight": (summ.final_cost - ref) / ref, "lm": summ.num_iterations,
                              "pcg": int(summ.pcg_iterations), "confirmations": summ.num_confirmations, "ms": round(dt * 1e3, 1),
                              "msg": summ.message.decode()}), flush=True)
