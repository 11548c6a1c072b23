"""Generates tests/golden/<name>_oracle_solve.json: the CPU oracle's full Levenberg-Marquardt solve (exact sparse
Cholesky steps, Ceres default tolerances, Huber(1.0) — the options of src/TagReconstructor.cpp:721,725-735) of a
synthetic site-scale scene, for the GPU parity tests at the sizes the oracle needs minutes for.

    python tools/make_site_golden.py c3        # ~10 min on 8 cores, ~12 GB

The scene generator is deterministic (visual_marker_mapping_b200/synth.py, SplitMix64-seeded PCG32), so the fixture
is valid wherever the same generator version runs.  Test infrastructure: imports oracle/.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402
from visual_marker_mapping_b200 import synth  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    loss = {"huber": oracle.LOSS_HUBER, "none": oracle.LOSS_NONE, "cauchy": oracle.LOSS_CAUCHY}[sys.argv[2] if len(sys.argv) > 2 else "huber"]
    scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    oracle.set_num_threads(len(os.sched_getaffinity(0)))
    s = synth.make_config(name)
    t0 = time.time()
    r = oracle.solve(s, max_num_iterations=50, loss_kind=loss, loss_scale=scale)
    dt = time.time() - t0
    cams = np.arange(0, s.n_cameras, max(1, s.n_cameras // 128))
    tags = np.arange(0, s.n_markers, max(1, s.n_markers // 128))
    out = {
        "scene": name, "generator": synth.GENERATOR_VERSION, "markers": s.n_markers, "cameras": s.n_cameras, "tag_obs": s.n_tag_obs,
        "loss_kind": int(loss), "loss_scale": scale, "cores": oracle.num_threads(), "seconds": dt,
        "summary": {"initial_cost": r.summary.initial_cost, "final_cost": r.summary.final_cost, "num_iterations": r.summary.num_iterations,
                    "num_successful_steps": r.summary.num_successful_steps, "num_unsuccessful_steps": r.summary.num_unsuccessful_steps,
                    "termination_type": r.summary.termination_type, "message": r.summary.message.decode(),
                    "time_linear_solver_s": r.summary.time_linear_solver_s, "time_evaluate_s": r.summary.time_evaluate_s},
        "trace_columns": ["cost", "gradient_max_norm", "step_norm", "relative_decrease", "trust_region_radius", "successful"],
        "trace": r.trace.tolist(),
        "sample_cameras": cams.tolist(), "cam_q": r.cam_q[cams].tolist(), "cam_t": r.cam_t[cams].tolist(),
        "sample_markers": tags.tolist(), "tag_q": r.tag_q[tags].tolist(), "tag_t": r.tag_t[tags].tolist(),
        "input_checksum": {"to_uv_sum": float(np.sum(s.to_uv)), "cam_t_sum": float(np.sum(s.cam_t)), "to_tag_sum": int(np.sum(s.to_tag.astype(np.int64)))},
    }
    path = os.path.join(ROOT, "tests", "golden", name.replace("/", "_") + ("" if loss == oracle.LOSS_HUBER else f"_loss{int(loss)}") + "_oracle_solve.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print(path, "final cost", r.summary.final_cost, "iterations", r.summary.num_iterations, f"{dt:.1f} s")


if __name__ == "__main__":
    main()
