"""Times the streamed Schur product (and one bench-style solve) for every variant library in tools/variants/.
    python tools/sp_sweep.py [scene]        -> one JSON line per variant
The scene is generated once and handed to the child processes as a pickle."""
import glob, json, os, pickle, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CHILD = r'''
import json, os, pickle, sys, time
sys.path.insert(0, %r)
import numpy as np
import visual_marker_mapping_b200 as pkg
s = pickle.load(open(sys.argv[1], "rb"))
x = np.random.default_rng(3).normal(size=6 * s.n_cameras)
out = {"variant": sys.argv[2]}
with pkg.BundleAdjuster(0) as ba:
    ba.set_scene(s)
    y0 = ba.schur_multiply(x, 1, 1.0, 1e4, path=0)
    ba.profile_enable(True)
    for _ in range(8):
        ba.schur_multiply(x, 1, 1.0, 1e4, path=0)
    for k, (n, ms) in ba.profile().items():
        if "SP_FULL" in k:
            out["sp_full_ms"] = ms / n
            out["frac_hbm"] = 292.0 * s.n_tag_obs / (ms / n * 1e-3) / 1e9 / 6450.9
    ba.profile_enable(False)
    y1 = ba.schur_multiply(x, 1, 1.0, 1e4, path=1)
    out["paths_agree"] = float(np.abs(y0 - y1).max() / np.abs(y1).max())
    if not os.environ.get("SP_SWEEP_NO_SOLVE"):
        ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
        t0 = time.time(); summ = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0); dt = time.time() - t0
        ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
        t0 = time.time(); summ = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0); dt = time.time() - t0
        out.update({"solve_s": dt, "lm": summ.num_iterations, "pcg": int(summ.pcg_iterations), "cost": summ.final_cost})
        ba.set_poses(s.cam_q, s.cam_t, s.tag_q, s.tag_t)
        ba.profile_enable(True)
        summ = ba.solve(max_num_iterations=50, loss_kind=1, loss_scale=1.0)
        ph = ba.pcg_phase_ns()
        n = max(int(summ.pcg_iterations), 1)
        out["phase_us_per_pcg_iteration"] = {k: round(v / 1e3 / n, 1) for k, v in ph.items() if v}
        for k, (cnt, ms) in ba.profile().items():
            if k.startswith("k_pcg_persistent"):
                out["pcg_kernel_us_per_iteration"] = round(ms * 1e3 / n, 1)
        ba.profile_enable(False)
print(json.dumps(out))
''' % ROOT
from visual_marker_mapping_b200 import synth
scene = synth.make_config(sys.argv[1] if len(sys.argv) > 1 else "c3")
with tempfile.NamedTemporaryFile(suffix=".pkl", delete=False) as f:
    pickle.dump(scene, f)
    pk = f.name
with tempfile.NamedTemporaryFile(suffix=".py", delete=False, mode="w") as f:
    f.write(CHILD)
    child = f.name
for lib in sorted(glob.glob(os.path.join(ROOT, "tools", "variants", "libvmm_ba_*.so"))):
    name = os.path.basename(lib)[len("libvmm_ba_"):-3]
    if name.startswith("inv"):
        continue
    for legacy in (("0", "1") if name == "base" else ("0",)) * int(os.environ.get("SP_SWEEP_REPEAT", "1")):
        env = dict(os.environ, VMM_BA_LIBRARY=lib)
        if legacy == "1":
            env["VMM_BA_PCG_LEGACY"] = "1"
        r = subprocess.run([sys.executable, child, pk, name + ("/legacy" if legacy == "1" else "/persistent")], env=env, capture_output=True, text=True, timeout=300)
        print(r.stdout.strip() if r.returncode == 0 else json.dumps({"variant": name, "error": r.stderr[-400:]}), flush=True)
        for line in r.stderr.splitlines():
            if "[sp timing]" in line:
                print(json.dumps({"variant": name, "timing": line}), flush=True)
                break
