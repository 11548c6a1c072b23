"""Multi-rank logic.

CPU (gloo, world_size 2, 127.0.0.1): the sharding plan of the engine — cameras split into contiguous
ranges balanced by observation count (vmm_ba_partition_cameras), markers replicated, marker-side
sums / cost all-reduced — reproduces the single-rank cost and gradient when the per-shard numbers
come from the oracle.

GPU (nccl, needs >= 2 devices): two engines created with vmm_ba_create_distributed solve the same
problem as one engine, to 1e-12 relative on the cost.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene():
    return synth.make_scene(n_markers=30, n_cameras=61, markers_per_camera=9, seed=8, outlier_frac=0.1)


def _gloo_worker(rank, world, port, out):
    from oracle import oracle

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = _scene()
    bounds = pkg.partition_cameras(s.n_cameras, s.to_cam, world)
    c0, c1 = int(bounds[rank]), int(bounds[rank + 1])
    keep = (s.to_cam >= c0) & (s.to_cam < c1)
    shard = s.copy()
    shard.to_cam, shard.to_tag, shard.to_uv = s.to_cam[keep].copy(), s.to_tag[keep].copy(), s.to_uv[keep].copy()
    ev = oracle.evaluate(shard, oracle.LOSS_HUBER, 1.0)
    g = ev.gradient.copy()
    # camera blocks outside the shard carry no residuals on this rank
    cam_part = g[:6 * s.n_cameras].reshape(-1, 6)
    assert not cam_part[:c0].any() and not cam_part[c1:].any()
    # what travels: marker gradient + cost (one all-reduce); camera gradients stay local and are gathered
    packed = torch.from_numpy(np.concatenate([g[6 * s.n_cameras:], [ev.cost, float(keep.sum())]]))
    dist.all_reduce(packed)
    cams = torch.zeros(6 * s.n_cameras, dtype=torch.float64)
    cams[6 * c0:6 * c1] = torch.from_numpy(g[6 * c0:6 * c1])
    dist.all_reduce(cams)          # the engine's gather_cameras(): zero-padded sum
    if rank == 0:
        out.put((bounds.tolist(), packed.numpy().copy(), cams.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_camera_sharding_plan_gloo_world2(oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    bounds, packed, cams = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    s = _scene()
    ref = oracle.evaluate(s, oracle.LOSS_HUBER, 1.0)
    assert bounds[0] == 0 and bounds[-1] == s.n_cameras
    assert int(packed[-1]) == s.n_tag_obs
    assert abs(packed[-2] - ref.cost) <= 1e-12 * ref.cost
    gm = ref.gradient[6 * s.n_cameras:]
    assert np.abs(packed[:-2] - gm).max() <= 1e-12 * np.abs(gm).max()
    assert np.abs(cams - ref.gradient[:6 * s.n_cameras]).max() <= 1e-12 * np.abs(ref.gradient).max()


# ------------------------------------------------------------------------------------------------
def _nccl_worker(rank, world, nccl_id, out):
    s = synth.make_config("c2")
    with pkg.BundleAdjuster(device=rank, rank=rank, world_size=world, nccl_id=nccl_id) as ba:
        ba.set_scene(s)
        a0, a1, ntol = ba.shard_info()
        ev = ba.evaluate(pkg.LOSS_HUBER, 1.0, residuals=False, jacobians=False)
        summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)
        cq, ct, tq, tt = ba.get_poses()
        out.put((rank, a0, a1, ntol, ev.cost, ev.gradient, summ.final_cost, summ.num_iterations, cq, ct, tq, tt))


@pytest.mark.gpu
def test_two_gpu_solve_matches_one_gpu():
    if pkg.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    s = synth.make_config("c2")
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        ev1 = ba.evaluate(pkg.LOSS_HUBER, 1.0, residuals=False, jacobians=False)
        s1 = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, parity_mode=1)
        p1 = ba.get_poses()
    nccl_id = pkg.engine.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, nccl_id, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=int(os.environ.get("VMM_TEST_2GPU_TIMEOUT_S", "240"))) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == s.n_cameras
    assert res[0][3] + res[1][3] == s.n_tag_obs
    for r in res:
        assert abs(r[4] - ev1.cost) <= 1e-12 * ev1.cost
        assert np.abs(r[5] - ev1.gradient).max() <= 1e-11 * np.abs(ev1.gradient).max()
        assert abs(r[6] - s1.final_cost) <= 1e-9 * s1.final_cost
        assert r[7] == s1.num_iterations
        assert np.abs(r[9] - p1[1]).max() < 1e-7 and np.abs(r[11] - p1[3]).max() < 1e-7
    # both ranks hold the same full result
    assert np.array_equal(res[0][9], res[1][9]) and np.array_equal(res[0][11], res[1][11])


def _nccl_worker_inexact(rank, world, nccl_id, out):
    s = synth.make_config("c3/40")
    with pkg.BundleAdjuster(device=rank, rank=rank, world_size=world, nccl_id=nccl_id) as ba:
        ba.set_scene(s)
        summ = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, function_tolerance=1e-10)
        sums = ba.reprojection_error_sums()
        cq, ct, tq, tt = ba.get_poses()
        a0, a1, ntol = ba.shard_info()
        out.put((rank, summ.final_cost, summ.num_iterations, int(summ.pcg_iterations), ct, tt, ntol, sums))


@pytest.mark.gpu
def test_two_gpu_inexact_solve_reaches_one_gpu_minimum():
    """Default (inexact Newton, two-level preconditioner) mode on a scene with interior and boundary markers:
    the sharded solve ends in the same minimum as the single-GPU one."""
    if pkg.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")
    s = synth.make_config("c3/40")
    _, boundary = pkg.shard_plan(s.cam_q, s.cam_t, s.n_markers, s.to_cam, s.to_tag, 2)
    assert 0 < (boundary >= 0).sum() < s.n_markers          # both paths of the sharded product are exercised
    with pkg.BundleAdjuster(0) as ba:
        ba.set_scene(s)
        s1 = ba.solve(max_num_iterations=50, loss_kind=pkg.LOSS_HUBER, loss_scale=1.0, function_tolerance=1e-10)
        p1 = ba.get_poses()
        sums1 = ba.reprojection_error_sums()
    nccl_id = pkg.engine.nccl_unique_id()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_nccl_worker_inexact, args=(r, 2, nccl_id, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=int(os.environ.get("VMM_TEST_2GPU_TIMEOUT_S", "240"))) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res[0][6] + res[1][6] == s.n_tag_obs
    for r in res:
        assert abs(r[1] - s1.final_cost) <= 1e-6 * s1.final_cost
        assert np.abs(r[4] - p1[1]).max() < 1e-4 and np.abs(r[5] - p1[3]).max() < 1e-4
    assert np.array_equal(res[0][4], res[1][4]) and np.array_equal(res[0][5], res[1][5])
    # per-observation outputs come back in the caller's order from both shards (each fills its own entries)
    owned0, owned1 = res[0][7] != 0, res[1][7] != 0
    assert not (owned0 & owned1).any() and (owned0 | owned1).all()
    merged = res[0][7] + res[1][7]
    assert np.abs(merged - sums1).max() < 1e-3


# ------------------------------------------------------------------------------------------------
# The sharded Schur product: spatial camera shards (vmm_ba_shard_plan), markers seen from one rank
# handled there in one pass, the others through a compact all-reduced boundary vector.
def _plan_scene():
    return synth.make_scene(n_markers=200, n_cameras=400, markers_per_camera=8, seed=11)


def _schur_pieces(s, ev, lam=1e-3):
    """per tag observation: E (6x6 camera x marker), and per marker C + lam I (undamped B is not needed here)"""
    J = ev.jac_local.reshape(-1, 8, 12)
    Jc, Jm = J[:, :, :6], J[:, :, 6:]
    E = np.einsum("kra,krb->kab", Jc, Jm)
    C = np.zeros((s.n_markers, 6, 6))
    np.add.at(C, s.to_tag, np.einsum("kra,krb->kab", Jm, Jm))
    C += lam * np.eye(6)[None]
    return E, C


def _plan_worker(rank, world, port, out):
    from oracle import oracle

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = _plan_scene()
    cam_rank, boundary = pkg.shard_plan(s.cam_q, s.cam_t, s.n_markers, s.to_cam, s.to_tag, world, cameras_per_cluster=16)
    ev = oracle.evaluate(s, oracle.LOSS_NONE, 1.0, gradient=False)
    E, C = _schur_pieces(s, ev)
    Cinv = np.linalg.inv(C)
    x = np.random.default_rng(5).standard_normal((s.n_cameras, 6))
    mine = cam_rank[s.to_cam] == rank                      # this rank's tag observations
    # w_j = sum_k E_k^T x[cam(k)] over the local observations
    w = np.zeros((s.n_markers, 6))
    np.add.at(w, s.to_tag[mine], np.einsum("kab,ka->kb", E[mine], x[s.to_cam[mine]]))
    # interior markers are complete locally; only the boundary ones travel, as one compact vector
    nb = int(boundary.max()) + 1 if (boundary >= 0).any() else 0
    wb = torch.zeros((max(nb, 1), 6), dtype=torch.float64)
    bsel = boundary >= 0
    wb[boundary[bsel]] = torch.from_numpy(w[bsel])
    dist.all_reduce(wb)
    w[bsel] = wb.numpy()[boundary[bsel]]
    z = np.einsum("jab,jb->ja", Cinv, w)
    y = np.zeros((s.n_cameras, 6))
    np.add.at(y, s.to_cam[mine], np.einsum("kab,kb->ka", E[mine], z[s.to_tag[mine]]))
    yt = torch.from_numpy(y)
    dist.all_reduce(yt)                                    # test-only: assemble the camera-sharded result
    if rank == 0:
        out.put((cam_rank, boundary, yt.numpy().copy(), nb))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_schur_product_plan_gloo_world2(oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_plan_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    cam_rank, boundary, y, nb = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    s = _plan_scene()
    # the plan: both ranks used, balanced by tag observations, interior markers really are interior
    counts = np.bincount(cam_rank[s.to_cam], minlength=2)
    assert counts.min() > 0.4 * s.n_tag_obs
    ranks_seen = [set(cam_rank[s.to_cam[s.to_tag == j]].tolist()) for j in range(s.n_markers)]
    for j in range(s.n_markers):
        if boundary[j] < 0:
            assert len(ranks_seen[j]) <= 1
        else:
            assert len(ranks_seen[j]) == 2 or (s.to_tag == j).sum() > 400
    ids = np.sort(boundary[boundary >= 0])
    assert np.array_equal(ids, np.arange(nb))
    # the spatial cut keeps most markers on one rank (a random cut of this scene leaves almost none)
    assert nb < 0.5 * s.n_markers
    # the product equals the single-rank one
    ev = oracle.evaluate(s, oracle.LOSS_NONE, 1.0, gradient=False)
    E, C = _schur_pieces(s, ev)
    x = np.random.default_rng(5).standard_normal((s.n_cameras, 6))
    w = np.zeros((s.n_markers, 6))
    np.add.at(w, s.to_tag, np.einsum("kab,ka->kb", E, x[s.to_cam]))
    z = np.einsum("jab,jb->ja", np.linalg.inv(C), w)
    ref = np.zeros((s.n_cameras, 6))
    np.add.at(ref, s.to_cam, np.einsum("kab,kb->ka", E, z[s.to_tag]))
    assert np.abs(y - ref).max() <= 1e-11 * np.abs(ref).max()
