"""Verbose LM trace of one solve (stderr shows cost / PCG iterations per LM iteration) + kernel table.
usage: [VMM_BA_LIBRARY=...] python tools/solve_trace.py c3 [max_iters] [key=value solver options ...]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 12
kw = {}
for a in sys.argv[3:]:
    k, v = a.split("=")
    kw[k] = float(v) if ("." in v or "e" in v) else int(v)
s = synth.make_config(name)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ba.solve(max_num_iterations=2, loss_kind=1, loss_scale=1.0)
    ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
    ba.profile_enable(True)
    t0 = time.perf_counter()
    sm = ba.solve(max_num_iterations=iters, loss_kind=1, loss_scale=1.0, verbose=1, **kw)
    dt = time.perf_counter() - t0
    prof = ba.profile()
    print(f"LM {sm.num_iterations} ok {sm.num_successful_steps} pcg {sm.pcg_iterations} cost {sm.initial_cost:.6e} -> {sm.final_cost:.9e} "
          f"{sm.message.decode()} {dt*1e3:.1f} ms")
    for k, (n, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1])[:8]:
        print(f"  {k:34s} n={n:5d} avg={ms/n:.4f} ms total={ms:.2f}")
