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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c3", help="c3 (default), c2, c1 or c3/<divisor>")
    ap.add_argument("--cpu-sample", default="c3/40", help="bounded CPU sample of the workload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--rj-iters", type=int, default=10)
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
def cpu_lm_sample(sample_name: str, full_corners: int, max_iters: int):
    from oracle import oracle
    from visual_marker_mapping_b200 import synth

    s = synth.make_config(sample_name)
    t0 = time.perf_counter()
    r = oracle.solve(s, max_num_iterations=max_iters, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    dt = time.perf_counter() - t0
    its = max(1, r.summary.num_iterations - 1)
    its_per_s_sample = its / dt
    scale = s.n_corners / float(full_corners)          # linear-in-corners extrapolation (favours the CPU:
    return {                                            # its exact factorisation grows super-linearly)
        "value": its_per_s_sample * scale,
        "unit": UNIT,
        "cores": oracle.num_threads(),
        "kind": "port",
        "sample": f"{sample_name} ({s.n_markers} markers, {s.n_cameras} cameras, {s.n_corners} corners): {its} LM iterations in "
                  f"{dt:.2f} s = {its_per_s_sample:.3f} it/s on the sample, scaled by corners {s.n_corners}/{full_corners}",
        "final_cost": r.summary.final_cost,
    }, s


def cpu_rj_sample(scene, reps: int = 3):
    from oracle import oracle

    oracle.evaluate(scene, oracle.LOSS_HUBER, 1.0, gradient=False)
    t0 = time.perf_counter()
    for _ in range(reps):
        oracle.evaluate(scene, oracle.LOSS_HUBER, 1.0, gradient=False)
    dt = (time.perf_counter() - t0) / reps
    return {"value": scene.n_corners / dt, "unit": "corner evaluations/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": f"{scene.n_corners} corners, dual-number residual+Jacobian, {reps} passes"}


def run_reference(args):
    """--impl reference: the reference algorithm on the host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from visual_marker_mapping_b200 import synth

    if args.workload == "c3":
        full_corners = 4 * 100000 * 40    # nominal C3 size; generating it is not needed on this arm
        wl = {"workload": "c3: 20000 markers, 100000 cameras, ~16M corners, radial-tangential distortion, Huber(1.0), fp64"}
    else:
        sc = synth.make_config(args.workload)
        full_corners = sc.n_corners
        wl = workload_description(sc, args.workload)
    sample = args.cpu_sample if args.workload == "c3" else args.workload
    # warm-up: thread pool + page faults
    cpu_lm_sample("c1", 1656, 3)
    t0 = time.perf_counter()
    base, s = cpu_lm_sample(sample, full_corners, max(1, args.steps))
    wall = time.perf_counter() - t0
    rj = cpu_rj_sample(s)
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 / base["value"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": wl,
        "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "rj": {"value": rj["value"], "unit": rj["unit"], "sample": rj["sample"]},
        "gpu_launches": 0, "wall_s": wall,
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
    ms_prof, _ = timed(lambda: lm_steps(args.steps))
    prof = ba.profile()
    ba.profile_enable(False)

    # ---- end to end through the C ABI from host buffers ----
    e2e = None
    if not args.no_e2e:
        host = {k: pinned(getattr(scene, k), np.int32 if k in ("to_cam", "to_tag") else np.float64)
                for k in ("cam_q", "cam_t", "tag_q", "tag_t", "tag_wh", "to_cam", "to_tag", "to_uv")}
        lm_steps(1, e2e=True)
        ms_e2e, sums_e = timed(lambda: lm_steps(args.steps, e2e=True))
        ncalls = len(sums_e)
        a0, a1, ntol = ba.shard_info()
        h2d = ncalls * (scene.n_tag_obs * (64 + 8) + (a1 - a0) * 56 + scene.n_markers * (56 + 16)) / args.steps
        d2h = ncalls * ((a1 - a0) * 56 + scene.n_markers * 56) / args.steps
        e2e = {"value": 1e3 / (ms_e2e / args.steps), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "solves": ncalls, "note": "vmm_ba_set_problem (H2D of the caller's arrays from pinned host memory, device-side sorts, host-side "
                       "cluster/segment set-up) + vmm_ba_solve + vmm_ba_get_poses per solve"}

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
            kernels[name] = ent
        if top and top[0] in KERNEL_BYTES_PER_TAG_OBS:
            k = kernels[top[0]]
            roofline = {"bound": "hbm", "achieved": k["gbs"], "peak": peak, "unit": "GB/s", "frac": k["frac_hbm"],
                        "traffic": NCU_TRAFFIC_BYTES_C3.get(top[0]) if (args.workload == "c3" and world == 1) else None,
                        "algorithmic_bytes": KERNEL_BYTES_PER_TAG_OBS[top[0]] * ntol,
                        "kernel": top[0], "bytes_per_tag_obs": KERNEL_BYTES_PER_TAG_OBS[top[0]], "peak_source": peak_src,
                        "share_of_step": k["share"]}
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            sample = args.cpu_sample if args.workload == "c3" else args.workload
            cpu, s_cpu = cpu_lm_sample(sample, scene.n_corners, 8)
            cpu["rj"] = cpu_rj_sample(s_cpu)
        s0 = sums[0]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": {**workload_description(scene, args.workload), "parallelism": f"camera shards x{world}, markers replicated"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks.summary(),
            "rj": rj, "kernels": kernels,
            "solve": {"solves_in_timed_region": len(sums), "first_solve": {"iterations": s0.num_iterations, "initial_cost": s0.initial_cost,
                      "final_cost": s0.final_cost, "pcg_iterations": int(s0.pcg_iterations), "termination": s0.message.decode(),
                      "t_linearize_s": s0.time_linearize_s, "t_linear_solver_s": s0.time_linear_solver_s, "t_cost_s": s0.time_cost_eval_s}},
            "profiled_pass_ms_per_step": ms_prof / args.steps,
        }
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
