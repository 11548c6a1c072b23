"""GPU study: cost of the coarse-level set-up against the PCG iterations it saves.

usage: python tools/coarse_sweep.py [workload] [cams_per_cluster ...]
Prints, per cluster size: solve time, LM / PCG iterations, and the per-LM-iteration time of the set-up
pieces (k_coarse_w, cusolverDnDpotrf, cuBLAS trsm / gemm / syrk) from the per-kernel CUDA-event profile.
The k-d leaves halve, so the cluster count moves in powers of two (C3: 192..390 -> 256 clusters, the library's cap;
512 -> 256, 1024 -> 128, ...).  Environment: REFRESH (coarse_refresh_iterations), PCGTOL, PRECOND, VERBOSE.
"""
import os, sys, time, json

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3"
    sizes = [int(a) for a in sys.argv[2:]] or [192, 512, 1024, 2048]
    scene = synth.make_config(name)
    out = []
    with pkg.BundleAdjuster(device=0) as ba:
        for cs in sizes:
            ba.set_coarse_cluster_size(cs)
            t0 = time.time()
            ba.set_scene(scene)
            t_set = time.time() - t0
            ba.solve(max_num_iterations=2, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)   # warm-up
            ba.set_poses(scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t)
            t0 = time.time()
            s = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, verbose=int(os.environ.get("VERBOSE", "0")),
                         coarse_refresh_iterations=int(os.environ.get("REFRESH", "8")), pcg_tolerance=float(os.environ.get("PCGTOL", "5e-2")), preconditioner=int(os.environ.get("PRECOND", "2")))
            t_solve = time.time() - t0
            ba.set_poses(scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t)
            ba.profile_enable(True)
            ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, coarse_refresh_iterations=int(os.environ.get("REFRESH", "8")), pcg_tolerance=float(os.environ.get("PCGTOL", "5e-2")), preconditioner=int(os.environ.get("PRECOND", "2")))
            prof = ba.profile()
            ba.profile_enable(False)
            row = {"cams_per_cluster": cs, "set_problem_s": round(t_set, 3), "solve_s": round(t_solve, 4), "lm": s.num_iterations,
                   "pcg": s.pcg_iterations, "final_cost": s.final_cost,
                   "t_linear_solver_s": round(s.time_linear_solver_s, 4)}
            for k in ("k_coarse_w", "cublasDtrsm", "cublasDgemm", "cublasDsyrk", "cusolverDnDpotrf", "k_schur_product<SP_FULL>", "k_coarse_apply"):
                if k in prof:
                    row[k + "_ms_avg"] = round(prof[k][1] / max(prof[k][0], 1), 4)
            print(json.dumps(row), flush=True)
            out.append(row)


if __name__ == "__main__":
    main()
