"""Time the one-pass Schur product alone (CUDA events via the engine's kernel profile)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth
s = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c3")
x = np.random.default_rng(3).normal(size=6 * s.n_cameras)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ba.schur_multiply(x, 1, 1.0, 1e4, path=0)
    ba.profile_enable(True)
    for _ in range(5):
        ba.schur_multiply(x, 1, 1.0, 1e4, path=0)
    for k, (n, ms) in ba.profile().items():
        if "schur" in k or "scale" in k:
            print(f"{k}: {ms/n:.4f} ms")
