"""Load balance of the streamed product against the cost model of upload_seglist (VMM_BA_SEG_FIXED / VMM_BA_SEG_TAIL, read once per process)."""
import json, os, pickle, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CHILD = r'''
import json, os, pickle, sys, time
sys.path.insert(0, %r)
import visual_marker_mapping_b200 as pkg
s = pickle.load(open(sys.argv[1], "rb"))
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0)
    ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
    ba.profile_enable(True)
    t0 = time.time(); summ = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0); dt = time.time() - t0
    n = max(int(summ.pcg_iterations), 1)
    ph = {k: round(v / 1e3 / n, 1) for k, v in ba.pcg_phase_ns().items() if v}
    us = [ms * 1e3 / n for k, (c, ms) in ba.profile().items() if k.startswith("k_pcg_persistent")]
    print(json.dumps({"fixed": os.environ.get("VMM_BA_SEG_FIXED"), "tail": os.environ.get("VMM_BA_SEG_TAIL"), "pcg": n, "us_per_iteration": round(us[0], 1), "phases": ph}))
''' % ROOT
from visual_marker_mapping_b200 import synth
scene = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c3")
with tempfile.NamedTemporaryFile(suffix=".pkl", delete=False) as f:
    pickle.dump(scene, f); pk = f.name
with tempfile.NamedTemporaryFile(suffix=".py", delete=False, mode="w") as f:
    f.write(CHILD); child = f.name
for fixed, tail in ((3, 0), (0, 0), (6, 0), (10, 0), (20, 0), (1000, 0), (3, 2), (6, 3), (10, 4)):
    r = subprocess.run([sys.executable, child, pk], env=dict(os.environ, VMM_BA_SEG_FIXED=str(fixed), VMM_BA_SEG_TAIL=str(tail)), capture_output=True, text=True, timeout=200)
    print(r.stdout.strip() if r.returncode == 0 else json.dumps({"fixed": fixed, "error": r.stderr[-300:]}), flush=True)
