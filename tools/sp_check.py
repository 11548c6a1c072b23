"""Repeatability / path-agreement check of the streamed Schur product at scale (GPU)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
s = synth.make_config(name)
rng = np.random.default_rng(3)
x = rng.normal(size=6 * s.n_cameras)
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ys = [ba.schur_multiply(x, 1, 1.0, 1e4, path=0) for _ in range(4)]
    y1 = [ba.schur_multiply(x, 1, 1.0, 1e4, path=1) for _ in range(2)]
    den = np.abs(ys[0]).max()
    print("path0 repeat diffs", [float(np.abs(y - ys[0]).max() / den) for y in ys[1:]])
    print("path1 repeat diff ", float(np.abs(y1[1] - y1[0]).max() / den))
    print("path0 vs path1    ", float(np.abs(y1[0] - ys[0]).max() / den))
    d = np.abs(y1[0] - ys[0]).reshape(-1, 6).max(axis=1) / den
    bad = np.nonzero(d > 1e-9)[0]
    print("cameras off by > 1e-9:", bad.size, bad[:20])
    # which markers / which observation positions inside the marker do the bad cameras belong to?
    if bad.size:
        dy = (ys[0] - y1[0]).reshape(-1, 6)
        badset = np.zeros(s.n_cameras, bool); badset[bad] = True
        order = np.argsort(s.to_tag, kind="stable")
        tag_sorted, cam_sorted = s.to_tag[order], s.to_cam[order]
        start = np.searchsorted(tag_sorted, np.arange(s.n_markers))
        pos = np.arange(order.size) - start[tag_sorted]
        counts = np.bincount(s.to_tag, minlength=s.n_markers)
        hit = badset[cam_sorted]
        per_marker = np.bincount(tag_sorted[hit], minlength=s.n_markers)
        top = np.argsort(-per_marker)[:12]
        for j in top:
            m = (tag_sorted == j) & hit
            print(f"marker {j}: nobs {counts[j]} bad {per_marker[j]} positions {np.sort(pos[m])[:6]}..{np.sort(pos[m])[-3:]}")
        print("markers with >= 8 bad cameras:", int((per_marker >= 8).sum()), " bad cameras:", bad.size)
