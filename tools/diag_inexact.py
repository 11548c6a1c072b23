"""Diagnostic: inexact (product-mode) solve on C2 run to tight tolerances, verbose trace."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth
from oracle import oracle

s = synth.make_config("c2")
kw = dict(max_num_iterations=100, loss_kind=1, loss_scale=1.0, function_tolerance=1e-14, parameter_tolerance=1e-12)
o = oracle.solve(s, **kw)
print("oracle", o.summary.final_cost, o.summary.num_iterations, o.summary.message)
for tol in (1e-2, 1e-4, 1e-8):
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        summ = ba.solve(verbose=0, pcg_tolerance=tol, **kw)
        cq, ct, tq, tt = ba.get_poses()
    print("tol", tol, summ.final_cost, summ.num_iterations, summ.pcg_iterations, summ.message.decode(),
          "dct %.3e dtt %.3e" % (np.abs(ct - o.cam_t).max(), np.abs(tt - o.tag_t).max()),
          "rel cost %.3e" % (abs(summ.final_cost - o.summary.final_cost) / o.summary.final_cost))
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    summ = ba.solve(verbose=1, **kw)
