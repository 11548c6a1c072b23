#!/usr/bin/env python
"""bench.py — LM iterations/s (and residual+Jacobian evaluations/s) of the bundle-adjustment path.

Contract:
  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA engine
  python bench.py --impl reference --gpus N --steps K ...  # the reference's algorithm on the host cores
A *step* is one Levenberg-Marquardt iteration (linear solve of the damped normal equations, cost at
the candidate point, re-linearisation when accepted) of the joint camera/marker refinement on the
BASELINE.json "large site" scene C3 (20k markers, 100k cameras, ~16M corners).  Solves run with the
reference's own options (Huber(1.0), Ceres default tolerances, src/TagReconstructor.cpp:721,725-735);
when a solve converges before K iterations are done the poses are reset to the initial guess and the
next solve continues the count.  N>1: the same scene sharded by camera over N ranks (strong scaling).
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "lm_iterations_per_s"
UNIT = "LM iterations/s"
RJ_BYTES_PER_CORNER = 226.0          # 16 uv + 2 idx + 16 r + 192 J
# algorithmic HBM bytes per tag observation of the solver kernels
KERNEL_BYTES_PER_TAG_OBS = {
    "k_linearize_cameras": 64 + 8,
    "k_linearize_markers": 64 + 4 + 288,
    "k_scale_e": 2 * 288,
    "k_schur_product<SP_FULL>": 288 + 4,
    "k_schur_product<SP_GATHER>": 288 + 4,
    "k_schur_product<SP_SCATTER>": 288 + 4,
    "k_schur_diag": 288 + 4,
    "k_cost": 64 + 8,
    "k_coarse_w": 288 + 4 + 1,
    "k_residual_jacobian": 4 * RJ_BYTES_PER_CORNER,
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` captures of this
# workload on one GPU (profiles/r01_ncu_solver.md, profiles/r01_ncu_rj.md); None for other workloads / shard sizes
NCU_TRAFFIC_BYTES_C3 = {
    "k_schur_product<SP_FULL>": 1.222487e9 + 0.009618e9,
    "k_residual_jacobian": 0.299655e9 + 3.268961e9,
    "k_coarse_w": 1.202321e9 + 0.011970e9,
}
# k_pcg_persistent (profiles/r02_ncu_pcg.md): the captured launch ran 47 PCG iterations and moved 61.956590 GB + 1.059650 GB, i.e.
# 1.341 GB per PCG iteration against 1.168 GB of scaled E blocks (the rest: B, M^-1, the camera-space vectors and A_c^-1 of
# phases B, C, D); scaled to the iterations of the measured launches below
NCU_TRAFFIC_BYTES_PER_PCG_ITERATION_C3 = (61.956590e9 + 1.059650e9) / 47


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", help="c3 (default), c2, c1 or c3/<divisor>")
    ap.add_argument("--cpu-iters", type=int, default=2, help="LM iterations of the CPU baseline leg of the GPU arm (on the workload itself)")
    ap.add_argument("--cpu-time-cap", type=float, default=150.0, help="--impl reference: stop after the LM iteration that crosses this many seconds")
    ap.add_argument("--no-cost-check", action="store_true")
    ap.add_argument("--no-c2-pair", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--rj-iters", type=int, default=10)
    ap.add_argument("--pcg-tolerance", type=float, default=None, help="A/B runs only: override the forcing term of the inexact solve (the default is the library's)")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([c.strip() for c in out.splitlines()[0].split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 5 + i and r[5 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def workload_description(scene, name):
    return {"workload": f"{name}: {scene.n_markers} markers, {scene.n_cameras} cameras, {scene.n_tag_obs} tag observations, "
                        f"{scene.n_corners} corners, radial-tangential distortion, Huber(1.0), fp64",
            "markers": scene.n_markers, "cameras": scene.n_cameras, "corners": scene.n_corners,
            "l2": "inputs larger than L2 (observations + E blocks stream > 126 MB per pass)" if scene.n_corners > 4_000_000
            else "inputs smaller than L2 (small parity-size workload)"}


# --------------------------------------------------------------------------------------------------
# CPU legs (the ONLY place outside tests/ and smoke() that executes oracle/): the reference's
# algorithm — dual-number Jacobians, Ceres-semantics LM with an exact linear solve — restated in
# oracle/vmm_oracle.cpp ("port": the reference itself needs Ceres/Eigen/OpenCV/Boost and cannot be
# built on this image), on all host cores, on a bounded sample of the workload.
# --------------------------------------------------------------------------------------------------
def host_cores() -> int:
    """Cores this process may use.  torchrun exports OMP_NUM_THREADS=1 to its ranks: the CPU legs ask for every core
    explicitly (oracle.set_num_threads) instead of inheriting that."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


CPU_SOLVER_NOTE = ("oracle/vmm_oracle.cpp: dual-number Jacobians, Ceres-semantics LM, exact sparse solve (cameras eliminated, block-envelope "
                   "Cholesky of the marker system in reverse Cuthill-McKee order; own 96x96 blocked AVX2 kernel + OpenMP, not BLAS / SuiteSparse)")


def cpu_lm_run(scene, name: str, max_iters: int, time_cap_s: float):
    """The reference algorithm on the host cores ON THE GIVEN SCENE ITSELF (no sub-sampling, no extrapolation): LM
    iterations until max_iters, convergence or the time cap (checked once per iteration, like Ceres' max_solver_time)."""
    from oracle import oracle

    oracle.set_num_threads(host_cores())
    t0 = time.perf_counter()
    r = oracle.solve(scene, max_num_iterations=max_iters, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0, max_solver_time_s=time_cap_s)
    dt = time.perf_counter() - t0
    its = max(1, r.summary.num_iterations - 1)
    return {
        "value": its / dt, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port", "same_config": True,
        "sample": f"{name} full scene ({scene.n_markers} markers, {scene.n_cameras} cameras, {scene.n_corners} corners): {its} LM iterations in {dt:.2f} s "
                  f"(iteration cap {max_iters}, time cap {time_cap_s:.0f} s; {r.summary.message.decode()})",
        "lm_iterations": its, "seconds": dt, "t_linear_solver_s": r.summary.time_linear_solver_s, "t_evaluate_s": r.summary.time_evaluate_s,
        "initial_cost": r.summary.initial_cost, "cost_after": r.summary.final_cost, "cost_trace": [float(c) for c in r.trace[:, 0]],
        "solver": CPU_SOLVER_NOTE,
    }


def cpu_rj_sample(scene, reps: int = 3):
    from oracle import oracle

    oracle.set_num_threads(host_cores())
    oracle.evaluate(scene, oracle.LOSS_HUBER, 1.0, gradient=False)
    t0 = time.perf_counter()
    for _ in range(reps):
        oracle.evaluate(scene, oracle.LOSS_HUBER, 1.0, gradient=False)
    dt = (time.perf_counter() - t0) / reps
    return {"value": scene.n_corners / dt, "unit": "corner evaluations/s", "cores": oracle.num_threads(), "kind": "port", "same_config": True,
            "sample": f"all {scene.n_corners} corners of the workload, dual-number residual+Jacobian, {reps} passes"}


def job_config(scene, name, world):
    return {**workload_description(scene, name), "parallelism": f"camera shards x{world}, markers replicated"}


def run_reference(args):
    """--impl reference: the reference algorithm on the host cores, on the workload itself (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    from visual_marker_mapping_b200 import synth

    scene = synth.make_config(args.workload)
    oracle.set_num_threads(host_cores())
    oracle.solve(synth.make_config("c1"), max_num_iterations=3, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)   # warm-up: thread pool
    t0 = time.perf_counter()
    base = cpu_lm_run(scene, args.workload, max(1, args.steps), args.cpu_time_cap)
    wall = time.perf_counter() - t0
    rj = cpu_rj_sample(scene, reps=2)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": base["lm_iterations"],
        "steps_requested": args.steps, "warmup": 0, "ms_per_step": 1e3 / base["value"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": job_config(scene, args.workload, args.gpus),
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "same_config", "solver")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "rj": {"value": rj["value"], "unit": rj["unit"], "sample": rj["sample"], "cores": rj["cores"]},
        "solve": {"initial_cost": base["initial_cost"], "cost_after": base["cost_after"], "cost_trace": base["cost_trace"],
                  "t_linear_solver_s": base["t_linear_solver_s"], "t_evaluate_s": base["t_evaluate_s"]},
        "gpu_launches": 0, "wall_s": wall,
        "note": "steps = LM iterations actually run and timed on the workload itself (the first solve of the scene, stopped by the iteration "
                "cap --steps, convergence or the time cap --cpu-time-cap); no warm-up steps are discarded (iteration 0's evaluation is inside the time)",
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import visual_marker_mapping_b200 as pkg
    from visual_marker_mapping_b200 import synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        box = [pkg.engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]

    scene = synth.make_config(args.workload)
    ba = pkg.BundleAdjuster(device=local_rank, rank=rank, world_size=world, nccl_id=nccl_id)
    stream = torch.cuda.Stream()               # torch.cuda.Event only sees the stream it is recorded on:
    ba.set_stream(stream.cuda_stream)          # run the engine on a torch stream and record on that one
    ba.set_scene(scene)
    opts = dict(loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)
    if args.pcg_tolerance is not None:
        opts["pcg_tolerance"] = args.pcg_tolerance
    init = (scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t)

    def pinned(a, dtype):
        """page-locked host copy of a numpy array in the ABI's dtype (the e2e leg uploads from pinned host memory)"""
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=dtype)).pin_memory()
        return t.numpy()

    host = None   # pinned copies of everything vmm_ba_set_problem reads, made on first use

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def lm_steps(k: int, e2e: bool = False):
        """k LM iterations as consecutive solves from the initial guess; returns the summaries."""
        done, sums = 0, []
        while done < k:
            if e2e:
                # H2D of the caller's observation arrays + initial poses from (pinned) host buffers; both device orders
                # are built on the GPU inside vmm_ba_set_problem
                ba.set_problem(host["cam_q"], host["cam_t"], host["tag_q"], host["tag_t"], host["tag_wh"], scene.fixed_tag,
                               host["to_cam"], host["to_tag"], host["to_uv"])
            else:
                ba.set_poses(*init)
            s = ba.solve(max_num_iterations=k - done, **opts)
            if e2e:
                ba.get_poses()                    # D2H of the result
            if s.num_iterations <= 0:
                raise SystemExit(f"solve made no progress: {s.message.decode()}")
            done += s.num_iterations
            sums.append(s)
        return sums

    def timed(fn):
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        out = fn()
        b.record(stream)
        barrier()
        return max_over_ranks(a.elapsed_time(b)), out

    # ---- warm-up, then the timed K steps (inputs resident in HBM) ----
    lm_steps(max(args.warmup, 3))
    with ClockSampler(local_rank) as clocks:
        ms_total, sums = timed(lambda: lm_steps(args.steps))
    ms_per_step = ms_total / args.steps
    value = 1e3 / ms_per_step
    launches = int(sum(s.kernel_launches for s in sums))

    # ---- the same K steps with per-kernel CUDA events (roofline of the dominant kernel) ----
    ba.profile_enable(True)
    ms_prof, sums_prof = timed(lambda: lm_steps(args.steps))
    prof = ba.profile()
    pcg_phases = ba.pcg_phase_ns()
    ba.profile_enable(False)
    pcg_its_prof = int(sum(s.pcg_iterations for s in sums_prof))

    # ---- end to end through the C ABI from host buffers ----
    e2e = None
    if not args.no_e2e:
        host = {k: pinned(getattr(scene, k), np.int32 if k in ("to_cam", "to_tag") else np.float64)
                for k in ("cam_q", "cam_t", "tag_q", "tag_t", "tag_wh", "to_cam", "to_tag", "to_uv")}
        lm_steps(1, e2e=True)
        b0 = ba.h2d_bytes()
        ms_e2e, sums_e = timed(lambda: lm_steps(args.steps, e2e=True))
        ncalls = len(sums_e)
        a0, a1, ntol = ba.shard_info()
        h2d = (ba.h2d_bytes() - b0) / args.steps        # counted by the library from the copies it issues
        d2h = ncalls * ((a1 - a0) * 56 + scene.n_markers * 56) / args.steps
        e2e = {"value": 1e3 / (ms_e2e / args.steps), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "solves": ncalls, "note": "per solve: vmm_ba_set_problem from pinned host buffers (the structure of the problem equals the previous "
                       "call's, so the library keeps its plans and copies this rank's observations, marker sizes and poses) + vmm_ba_solve + "
                       "vmm_ba_get_poses; VMM_BA_NO_PLAN_CACHE=1 rebuilds everything per call"}

    # ---- residual+Jacobian materialising pass (HBM roofline kernel) ----
    ba.set_poses(*init)
    ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 3)
    ms_rj = max_over_ranks(ba.evaluate_device(pkg.LOSS_HUBER, 1.0, args.rj_iters))
    peak, peak_src = load_peaks()
    a0, a1, ntol = ba.shard_info()
    ntol_t = torch.tensor([ntol], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ntol_t, op=dist.ReduceOp.MAX)
    ntol_max = float(ntol_t.item())
    rj_gbs = 4 * ntol_max * RJ_BYTES_PER_CORNER / (ms_rj * 1e-3) / 1e9
    rj = {"value": scene.n_corners / (ms_rj * 1e-3), "unit": "corner evaluations/s", "ms_per_pass": ms_rj,
          "roofline": {"bound": "hbm", "achieved": rj_gbs, "peak": peak, "unit": "GB/s", "frac": rj_gbs / peak,
                       "traffic": NCU_TRAFFIC_BYTES_C3["k_residual_jacobian"] if (args.workload == "c3" and world == 1) else None,
                       "algorithmic_bytes": 4 * ntol_max * RJ_BYTES_PER_CORNER,
                       "kernel": "k_residual_jacobian", "bytes_per_corner": RJ_BYTES_PER_CORNER, "peak_source": peak_src}}

    if rank == 0:
        top = max(prof.items(), key=lambda kv: kv[1][1]) if prof else None
        roofline = None
        kernels = {}
        for name, (n, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
            ent = {"launches": n, "ms_total": round(ms, 4), "ms_avg": ms / n, "share": ms / ms_prof}
            if name in KERNEL_BYTES_PER_TAG_OBS:
                gbs = KERNEL_BYTES_PER_TAG_OBS[name] * ntol / (ms / n * 1e-3) / 1e9
                ent.update({"gbs": gbs, "frac_hbm": gbs / peak})
            elif name in ("k_pcg_persistent", "k_pcg_persistent_dist"):
                # one launch = one whole PCG solve: every iteration streams the scaled E blocks once (292 B per tag observation);
                # the grid barriers, the vector updates and the coarse mat-vec of each iteration are inside the measured time
                per_launch = KERNEL_BYTES_PER_TAG_OBS["k_schur_product<SP_FULL>"] * ntol * pcg_its_prof / max(n, 1)
                gbs = per_launch / (ms / n * 1e-3) / 1e9
                ent.update({"gbs": gbs, "frac_hbm": gbs / peak, "pcg_iterations": pcg_its_prof, "ms_per_pcg_iteration": ms / max(pcg_its_prof, 1),
                            "algorithmic_bytes_per_launch": per_launch,
                            "phase_us_per_iteration_cta0": {k: v / 1e3 / max(pcg_its_prof, 1) for k, v in pcg_phases.items()}})
            kernels[name] = ent
        if top and "gbs" in kernels[top[0]]:
            k = kernels[top[0]]
            per_obs = KERNEL_BYTES_PER_TAG_OBS.get(top[0], KERNEL_BYTES_PER_TAG_OBS["k_schur_product<SP_FULL>"])
            roofline = {"bound": "hbm", "achieved": k["gbs"], "peak": peak, "unit": "GB/s", "frac": k["frac_hbm"],
                        "traffic": NCU_TRAFFIC_BYTES_C3.get(top[0]) if (args.workload == "c3" and world == 1) else None,
                        "algorithmic_bytes": k.get("algorithmic_bytes_per_launch", per_obs * ntol),
                        "kernel": top[0], "bytes_per_tag_obs": per_obs, "peak_source": peak_src, "share_of_step": k["share"],
                        "launch_ms": k["ms_avg"]}
            if top[0] == "k_pcg_persistent":
                if args.workload == "c3" and world == 1:
                    roofline["traffic"] = NCU_TRAFFIC_BYTES_PER_PCG_ITERATION_C3 * pcg_its_prof / max(k["launches"], 1)
                roofline["note"] = ("one launch = one PCG solve of %.1f iterations on average; per iteration 292 B per tag observation are streamed once; "
                                    "traffic = DRAM bytes per PCG iteration of the ncu capture (47 iterations, 63.0 GB: profiles/r02_ncu_pcg.md) x the iterations of an average launch here" % (pcg_its_prof / max(k["launches"], 1)))
            roofline["rj"] = rj
        # ---- cost check (north_star: the solve reaches the reference's cost): default-tolerance solve, both modes at
        #      function_tolerance 1e-12, and the oracle's exact-Cholesky solve of the same scene (committed fixture) ----
        cost_check = None
        if not args.no_cost_check and world == 1:
            cost_check = {}
            for label, kw in (("default", {}), ("parity_ftol_1e-12", dict(parity_mode=1, function_tolerance=1e-12)),
                              ("product_ftol_1e-12", dict(function_tolerance=1e-12))):
                ba.set_poses(*init)
                sc = ba.solve(max_num_iterations=200, **opts, **kw)
                cost_check[label] = {"final_cost": sc.final_cost, "lm_iterations": sc.num_iterations, "pcg_iterations": int(sc.pcg_iterations),
                                     "confirmations": sc.num_confirmations, "termination": sc.message.decode()}
            gpath = os.path.join(ROOT, "tests", "golden", args.workload.replace("/", "_") + "_oracle_solve.json")
            if os.path.exists(gpath):
                with open(gpath) as f:
                    g = json.load(f)
                cost_check["oracle_exact_solve"] = {"final_cost": g["summary"]["final_cost"], "lm_iterations": g["summary"]["num_iterations"] - 1,
                                                    "source": "tests/golden/" + os.path.basename(gpath)}
                cost_check["default_vs_oracle_rel"] = abs(cost_check["default"]["final_cost"] - g["summary"]["final_cost"]) / g["summary"]["final_cost"]
            tight = cost_check["parity_ftol_1e-12"]["final_cost"]
            cost_check["default_vs_tight_rel"] = (cost_check["default"]["final_cost"] - tight) / tight
            cost_check["product_tight_vs_parity_tight_rel"] = (cost_check["product_ftol_1e-12"]["final_cost"] - tight) / tight
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            # the reference algorithm on this box's host cores, ON THE WORKLOAD ITSELF: a capped number of LM iterations
            cpu = cpu_lm_run(scene, args.workload, args.cpu_iters, 1e9)
            cpu["rj"] = cpu_rj_sample(scene)
            cpu["gpu_over_cpu"] = {"value": value / cpu["value"], "e2e": (e2e["value"] / cpu["value"]) if e2e else None,
                                   "rj": rj["value"] / cpu["rj"]["value"], "same_config": True}
            if not args.no_c2_pair and args.workload != "c2":
                # a second pair where BOTH arms run the whole solve to convergence: C2 (BASELINE.json configs[1])
                from oracle import oracle
                c2 = synth.make_config("c2")
                t0 = time.perf_counter()
                oc = oracle.solve(c2, max_num_iterations=50, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
                t_cpu = time.perf_counter() - t0
                hc2 = {k: pinned(getattr(c2, k), np.int32 if k in ("to_cam", "to_tag") else np.float64)
                       for k in ("cam_q", "cam_t", "tag_q", "tag_t", "tag_wh", "to_cam", "to_tag", "to_uv")}

                def c2_solve():
                    ba.set_problem(hc2["cam_q"], hc2["cam_t"], hc2["tag_q"], hc2["tag_t"], hc2["tag_wh"], c2.fixed_tag, hc2["to_cam"], hc2["to_tag"], hc2["to_uv"])
                    sg = ba.solve(max_num_iterations=50, **opts)
                    ba.get_poses()
                    return sg
                c2_solve()
                t0 = time.perf_counter()
                torch.cuda.synchronize()
                sg = c2_solve()
                torch.cuda.synchronize()
                t_gpu = time.perf_counter() - t0
                cpu["c2_full_solve_pair"] = {
                    "workload": workload_description(c2, "c2")["workload"], "same_config": True,
                    "cpu": {"lm_iterations": oc.summary.num_iterations - 1, "seconds": t_cpu, "value": (oc.summary.num_iterations - 1) / t_cpu,
                            "final_cost": oc.summary.final_cost, "cores": oracle.num_threads()},
                    "gpu_e2e": {"lm_iterations": sg.num_iterations, "seconds": t_gpu, "value": sg.num_iterations / t_gpu, "final_cost": sg.final_cost,
                                "note": "host buffers -> vmm_ba_set_problem -> vmm_ba_solve -> vmm_ba_get_poses, wall clock"},
                    "unit": UNIT, "final_cost_rel_diff": abs(sg.final_cost - oc.summary.final_cost) / oc.summary.final_cost,
                    "time_to_solution_ratio": t_cpu / t_gpu}
        s0 = sums[0]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(job_config(scene, args.workload, world), **({"pcg_tolerance_override": args.pcg_tolerance} if args.pcg_tolerance is not None else {})),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks.summary(),
            "rj": rj, "kernels": kernels,
            "solve": {"solves_in_timed_region": len(sums), "first_solve": {"iterations": s0.num_iterations, "initial_cost": s0.initial_cost,
                      "final_cost": s0.final_cost, "pcg_iterations": int(s0.pcg_iterations), "termination": s0.message.decode(),
                      "t_linearize_s": s0.time_linearize_s, "t_linear_solver_s": s0.time_linear_solver_s, "t_cost_s": s0.time_cost_eval_s}},
            "profiled_pass_ms_per_step": ms_prof / args.steps,
            "fp64_peak_note": "FP64 fractions anywhere in this repo are against a builder-measured DFMA rate (vmm_ba_measure_fp64_peak, 34.3 TFLOP/s); "
                              "MEASURED_PEAKS.json has no FP64 entry",
        }
        line["solve"]["cost_check"] = cost_check
        print(json.dumps(line), flush=True)
    ba.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
