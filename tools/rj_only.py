"""Residual+Jacobian materialising pass only (for ncu): C3-sized scene, a few launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

s = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c3")
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ms = ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 6)
    print("corners", s.n_corners, "ms/pass", ms, "GB/s", s.n_corners * 226 / ms / 1e6)
