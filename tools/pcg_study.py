"""GPU study: PCG iterations / LM iterations / time of a full solve on a scene for several PCG settings."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
s = synth.make_config(name)
print(name, s.n_markers, s.n_cameras, s.n_tag_obs, flush=True)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    init = (s.cam_q, s.cam_t, s.tag_q, s.tag_t)
    for pre in (1, 0):
        for tol in (1e-1, 1e-2, 1e-3):
            ba.set_poses(*init)
            t0 = time.perf_counter()
            sm = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0, pcg_tolerance=tol, preconditioner=pre,
                          pcg_max_iterations=2000)
            dt = time.perf_counter() - t0
            print(f"precond {pre} tol {tol:g}: LM {sm.num_iterations} (ok {sm.num_successful_steps}) pcg {sm.pcg_iterations} "
                  f"cost {sm.initial_cost:.6e} -> {sm.final_cost:.9e} {sm.message.decode()} {dt*1e3:.1f} ms "
                  f"(lin {sm.time_linearize_s*1e3:.1f} solve {sm.time_linear_solver_s*1e3:.1f})", flush=True)
    ba.set_poses(*init)
    ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0, verbose=1, pcg_max_iterations=2000)
