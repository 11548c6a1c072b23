"""BASELINE.json configurations C4 and C5 on N GPUs of one box (one rank per GPU, camera shards):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/run_configs_dist.py c5 [max_corners]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/run_configs_dist.py c4 [divisor]
  ... run_configs_dist.py c5c4 max_corners [divisor]     both in one process (C4_HUBER_ONLY=1: skip the Cauchy solve)
also runs as a plain script (N = 1).  Rank 0 prints one JSON line per measurement; times are CUDA events on the engine's
stream, max over ranks (the whole job is as slow as its slowest shard)."""
import json, os, sys, time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
nccl_id = None
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    box = [pkg.engine.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    nccl_id = box[0]


def max_over_ranks(v: float) -> float:
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def out(row):
    if rank == 0:
        print(json.dumps(dict(row, n_gpus=world)), flush=True)


what = sys.argv[1] if len(sys.argv) > 1 else "c5"
with pkg.BundleAdjuster(device=local, rank=rank, world_size=world, nccl_id=nccl_id) as ba:
    if what in ("c5", "c5c4"):
        max_corners = int(sys.argv[2]) if len(sys.argv) > 2 else 256_000_000
        n, base = 1_000_000, None
        while n <= max_corners:
            if n <= 16_000_000:
                scene = synth.make_rj_sweep_scene(n)
            else:
                base = base or synth.make_rj_sweep_scene(16_000_000)
                scene = synth.tile_scene(base, n // 16_000_000)
            ba.set_scene(scene)
            ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 2)
            ms = max_over_ranks(ba.evaluate_device(pkg.LOSS_HUBER, 1.0, 5))   # this rank's shard; the job = the slowest shard
            gbs = scene.n_corners * 226.0 / (ms * 1e-3) / 1e9
            out({"config": "c5", "corners": scene.n_corners, "ms_per_pass": round(ms, 4), "corners_per_s": scene.n_corners / (ms * 1e-3),
                 "gbs_whole_job": round(gbs, 1), "frac_hbm_of_n_gpus": round(gbs / (PEAK * world), 4)})
            del scene
            n *= 4
    if what in ("c4", "c5c4"):
        name = "c4" if len(sys.argv) <= 3 - (what == "c4") or sys.argv[-1] in ("1", "0") else f"c4/{sys.argv[-1]}"
        scene = synth.make_config(name)
        ba.set_scene(scene)
        losses = (("huber(1.0)", dict(loss_kind=pkg.LOSS_HUBER, loss_scale=1.0)), ("cauchy(3.0)", dict(loss_kind=pkg.LOSS_CAUCHY, loss_scale=3.0)))
        for label, kw in losses[:1 if os.environ.get("C4_HUBER_ONLY") else 2]:
            ba.set_poses(scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.time()
            s = ba.solve(max_num_iterations=100, **kw)
            torch.cuda.synchronize()
            dt = max_over_ranks(time.time() - t0)
            cq, ct, tq, tt = ba.get_poses()
            row = {"config": name, "loss": label, "markers": scene.n_markers, "cameras": scene.n_cameras, "corners": scene.n_corners,
                   "solve_s": round(dt, 4), "lm_iterations": s.num_iterations, "successful": s.num_successful_steps,
                   "pcg_iterations": int(s.pcg_iterations), "initial_cost": s.initial_cost, "final_cost": s.final_cost,
                   "termination": s.message.decode(), "lm_iterations_per_s": s.num_iterations / dt}
            if scene.tag_t_gt is not None:
                em = np.sqrt(((tt - scene.tag_t_gt) ** 2).sum(1))
                row["marker_translation_error_m"] = {"median": float(np.median(em)), "p95": float(np.percentile(em, 95)), "max": float(em.max())}
            out(row)
if world > 1:
    dist.destroy_process_group()
