"""One solve of the 'split' scene (markers longer than a segment: boundary path of the persistent kernel) — for compute-sanitizer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth
s = synth.make_scene(seed=77, n_markers=12, n_cameras=900, markers_per_camera=12)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    summ = ba.solve(max_num_iterations=5, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
    print("ok", summ.final_cost, summ.num_iterations, int(summ.pcg_iterations), summ.message.decode())
