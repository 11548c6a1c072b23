"""BASELINE.json configurations beyond the bench line, on one GPU (or one rank per GPU under torchrun for C5):

  python tools/run_configs.py c5 [max_corners]   residual+Jacobian-only sweep 1M..256M corners (HBM roofline)
  python tools/run_configs.py c4                 outlier stress: 50k markers / 200k cameras, 20 % corrupted tag
                                                 observations, solved with Huber(1.0) and with Cauchy(3.0)
  python tools/run_configs.py c2                 synthetic room, parity mode against the default mode
One JSON line per measurement on stdout.
"""
import json, os, sys, time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]


def c5(max_corners):
    base = None
    with pkg.BundleAdjuster(0) as ba:
        n = 1_000_000
        while n <= max_corners:
            t0 = time.time()
            if n <= 16_000_000:
                scene = synth.make_rj_sweep_scene(n)
            else:
                base = base or synth.make_rj_sweep_scene(16_000_000)
                scene = synth.tile_scene(base, n // 16_000_000)
            t_gen = time.time() - t0
            t0 = time.time()
            ba.set_scene(scene)
            t_set = time.time() - t0
            ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 2)
            ms = ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 5)
            gbs = scene.n_corners * 226.0 / (ms * 1e-3) / 1e9
            print(json.dumps({"config": "c5", "corners": scene.n_corners, "ms_per_pass": round(ms, 4), "corners_per_s": scene.n_corners / (ms * 1e-3),
                              "gbs": round(gbs, 1), "frac_hbm": round(gbs / PEAK, 4), "set_problem_s": round(t_set, 3), "generate_s": round(t_gen, 1)}), flush=True)
            del scene
            n *= 4


def solve_line(ba, scene, tag, **kw):
    ba.set_poses(scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t)
    t0 = time.time()
    if "PCGTOL" in os.environ:
        kw = dict(kw, pcg_tolerance=float(os.environ["PCGTOL"]))
    s = ba.solve(max_num_iterations=100, **kw)
    dt = time.time() - t0
    cq, ct, tq, tt = ba.get_poses()
    row = {"config": tag, "markers": scene.n_markers, "cameras": scene.n_cameras, "corners": scene.n_corners, "solve_s": round(dt, 4),
           "lm_iterations": s.num_iterations, "successful": s.num_successful_steps, "pcg_iterations": int(s.pcg_iterations),
           "initial_cost": s.initial_cost, "final_cost": s.final_cost, "termination": s.message.decode(),
           "lm_iterations_per_s": s.num_iterations / dt}
    if scene.tag_t_gt is not None:
        em = np.sqrt(((tt - scene.tag_t_gt) ** 2).sum(1))
        ec = np.sqrt(((ct - scene.cam_t_gt) ** 2).sum(1))
        for name, e in (("marker", em), ("camera", ec)):
            row[name + "_translation_error_m"] = {"median": float(np.median(e)), "p95": float(np.percentile(e, 95)),
                                                  "p99.9": float(np.percentile(e, 99.9)), "max": float(e.max())}
    print(json.dumps(row), flush=True)


def c4():
    t0 = time.time()
    scene = synth.make_config("c4")
    print(json.dumps({"config": "c4", "generate_s": round(time.time() - t0, 1), "tag_obs": scene.n_tag_obs,
                      "corrupted": int(scene.outlier_mask.sum()) if scene.outlier_mask is not None else None}), flush=True)
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene)
        ba.solve(max_num_iterations=1, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)   # warm-up
        solve_line(ba, scene, "c4 huber(1.0)", loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        solve_line(ba, scene, "c4 cauchy(3.0)", loss_kind=pkg.LOSS_CAUCHY, loss_scale=3.0)


def c2():
    scene = synth.make_config("c2")
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(scene)
        ba.solve(max_num_iterations=1, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        solve_line(ba, scene, "c2 huber default", loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
        solve_line(ba, scene, "c2 huber parity_mode", loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "c5"
    if which == "c5":
        c5(int(sys.argv[2]) if len(sys.argv) > 2 else 256_000_000)
    elif which == "c4":
        c4()
    else:
        c2()
