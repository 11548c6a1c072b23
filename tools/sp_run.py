"""A few Schur products on a scene (for ncu captures): python tools/sp_run.py c3 [path] [reps]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth
name = sys.argv[1] if len(sys.argv) > 1 else "c3"
path = int(sys.argv[2]) if len(sys.argv) > 2 else 0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
s = synth.make_config(name)
x = np.random.default_rng(3).normal(size=6 * s.n_cameras)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    for _ in range(reps):
        y = ba.schur_multiply(x, 1, 1.0, 1e4, path=path)
    print("ok", float(np.abs(y).max()))
