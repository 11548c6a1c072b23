"""Synthetic marker-mapping scenes of the BASELINE.json shapes.

The reference ships no dataset (README.md:109 points at an external download), so every test and
benchmark of the bundle-adjustment path runs on scenes generated here.  A scene is exactly the input
of ``TagReconstructor::doBundleAdjustment`` (src/TagReconstructor.cpp:646-724) flattened to dense
0-based indices: one shared camera model, camera poses (world->camera), marker poses
(marker->world), and tag observations (4 corners each, order LL, LR, UR, UL,
include/visual_marker_mapping/TagReconstructor.h:44-52).

Geometry ("marker sheet"): markers lie on a gently undulating ground sheet with tilted normals,
cameras hover above it looking down with random yaw/tilt, and each camera records its
``markers_per_camera`` nearest fully visible markers.  Intrinsics are the README example
(README.md:132-145), marker size 0.1285 x 0.1295 m (README.md:116).  Marker 0 is the fixed origin
tag at the identity pose (src/TagReconstructor.cpp:134-141).  Deterministic: numpy PCG64 with a
fixed seed per configuration.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np

GENERATOR_VERSION = "vmm-synth-4"

# README.md:132-145 — fx, fy, cx, cy, k1, k2, p1, p2, k3
README_INTRINSICS = np.array(
    [
        8075.2937867635346,
        8083.1676114192869,
        3016.3896805084278,
        1996.2896554785455,
        -0.1861818326266976,
        0.37018092365577054,
        -2.9390604003594177e-4,
        4.1533180829908799e-4,
        5.7043887874185996e-2,
    ],
    dtype=np.float64,
)
IMAGE_W, IMAGE_H = 6000, 4000
MARKER_W, MARKER_H = 0.1285, 0.1295
CORNER_SIGNS = np.array([[-1.0, -1.0], [1.0, -1.0], [1.0, 1.0], [-1.0, 1.0]])


# ----------------------------------------------------------------------------------------------
# quaternion helpers, (w, x, y, z) order as in Camera.h:13 / TagReconstructor.h:19
# ----------------------------------------------------------------------------------------------
def quat_to_rot(q: np.ndarray) -> np.ndarray:
    q = np.asarray(q, dtype=np.float64)
    q = q / np.linalg.norm(q, axis=-1, keepdims=True)
    w, x, y, z = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
    R = np.empty(q.shape[:-1] + (3, 3))
    R[..., 0, 0] = 1 - 2 * (y * y + z * z)
    R[..., 0, 1] = 2 * (x * y - w * z)
    R[..., 0, 2] = 2 * (x * z + w * y)
    R[..., 1, 0] = 2 * (x * y + w * z)
    R[..., 1, 1] = 1 - 2 * (x * x + z * z)
    R[..., 1, 2] = 2 * (y * z - w * x)
    R[..., 2, 0] = 2 * (x * z - w * y)
    R[..., 2, 1] = 2 * (y * z + w * x)
    R[..., 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def quat_to_rot_raw(q: np.ndarray) -> np.ndarray:
    """Eigen::Quaterniond::toRotationMatrix: no normalisation (computeMarkerCorners3D, TagReconstructor.h:33-43)."""
    q = np.asarray(q, dtype=np.float64)
    n2 = np.sum(q * q, axis=-1, keepdims=True)
    # R_raw = I + (R_unit - I) * |q|^2  for the 2*q_i*q_j form used by Eigen
    R = quat_to_rot(q)
    return np.eye(3) + (R - np.eye(3)) * n2[..., None]


def rot_to_quat(R: np.ndarray) -> np.ndarray:
    """Batch rotation matrix -> unit quaternion (w,x,y,z), w >= 0."""
    R = np.asarray(R, dtype=np.float64)
    batch = R.shape[:-2]
    Rf = R.reshape(-1, 3, 3)
    n = Rf.shape[0]
    q = np.empty((n, 4))
    tr = Rf[:, 0, 0] + Rf[:, 1, 1] + Rf[:, 2, 2]
    # four candidate branches, pick the best conditioned per element
    cand = np.stack(
        [
            1 + tr,
            1 + Rf[:, 0, 0] - Rf[:, 1, 1] - Rf[:, 2, 2],
            1 - Rf[:, 0, 0] + Rf[:, 1, 1] - Rf[:, 2, 2],
            1 - Rf[:, 0, 0] - Rf[:, 1, 1] + Rf[:, 2, 2],
        ],
        axis=1,
    )
    best = np.argmax(cand, axis=1)
    for b in range(4):
        m = best == b
        if not np.any(m):
            continue
        Rm = Rf[m]
        s = 2.0 * np.sqrt(cand[m, b])
        if b == 0:
            q[m, 0] = 0.25 * s
            q[m, 1] = (Rm[:, 2, 1] - Rm[:, 1, 2]) / s
            q[m, 2] = (Rm[:, 0, 2] - Rm[:, 2, 0]) / s
            q[m, 3] = (Rm[:, 1, 0] - Rm[:, 0, 1]) / s
        elif b == 1:
            q[m, 0] = (Rm[:, 2, 1] - Rm[:, 1, 2]) / s
            q[m, 1] = 0.25 * s
            q[m, 2] = (Rm[:, 0, 1] + Rm[:, 1, 0]) / s
            q[m, 3] = (Rm[:, 0, 2] + Rm[:, 2, 0]) / s
        elif b == 2:
            q[m, 0] = (Rm[:, 0, 2] - Rm[:, 2, 0]) / s
            q[m, 1] = (Rm[:, 0, 1] + Rm[:, 1, 0]) / s
            q[m, 2] = 0.25 * s
            q[m, 3] = (Rm[:, 1, 2] + Rm[:, 2, 1]) / s
        else:
            q[m, 0] = (Rm[:, 1, 0] - Rm[:, 0, 1]) / s
            q[m, 1] = (Rm[:, 0, 2] + Rm[:, 2, 0]) / s
            q[m, 2] = (Rm[:, 1, 2] + Rm[:, 2, 1]) / s
            q[m, 3] = 0.25 * s
    q *= np.where(q[:, :1] < 0, -1.0, 1.0)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.reshape(batch + (4,))


def rotvec_to_rot(rv: np.ndarray) -> np.ndarray:
    rv = np.asarray(rv, dtype=np.float64)
    th = np.linalg.norm(rv, axis=-1)
    safe = np.where(th > 1e-300, th, 1.0)
    k = rv / safe[..., None]
    K = np.zeros(rv.shape[:-1] + (3, 3))
    K[..., 0, 1] = -k[..., 2]
    K[..., 0, 2] = k[..., 1]
    K[..., 1, 0] = k[..., 2]
    K[..., 1, 2] = -k[..., 0]
    K[..., 2, 0] = -k[..., 1]
    K[..., 2, 1] = k[..., 0]
    s = np.sin(th)[..., None, None]
    c = np.cos(th)[..., None, None]
    return np.eye(3) + s * K + (1 - c) * (K @ K)


def quat_angle(q1: np.ndarray, q2: np.ndarray) -> np.ndarray:
    """Rotation angle (rad) between two quaternions, sign- and scale-insensitive."""
    q1 = q1 / np.linalg.norm(q1, axis=-1, keepdims=True)
    q2 = q2 / np.linalg.norm(q2, axis=-1, keepdims=True)
    d = np.abs(np.sum(q1 * q2, axis=-1))
    # angle of q1 * q2^-1: 2*acos(|<q1,q2>|); use the sine form for accuracy near zero
    diff = np.minimum(np.linalg.norm(q1 - q2, axis=-1), np.linalg.norm(q1 + q2, axis=-1))
    return np.where(d > 0.9, 4.0 * np.arcsin(np.clip(diff / 2.0, 0, 1)), 2.0 * np.arccos(np.clip(d, 0, 1)))


def project(intr: np.ndarray, Xc: np.ndarray) -> np.ndarray:
    """OpenCV pinhole + radial-tangential model, the functor's maths
    (TagReconstructionCostFunction.h:125-152) in plain numpy."""
    fx, fy, cx, cy, k1, k2, p1, p2, k3 = intr
    x = Xc[..., 0] / Xc[..., 2]
    y = Xc[..., 1] / Xc[..., 2]
    r2 = x * x + y * y
    c = 1 + r2 * (k1 + r2 * (k2 + r2 * k3))
    xd = x * c + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
    yd = y * c + 2 * p2 * x * y + p1 * (r2 + 2 * y * y)
    return np.stack([fx * xd + cx, fy * yd + cy], axis=-1)


@dataclass
class Scene:
    """Flattened BA input (dense indices). Poses are the INITIAL GUESS; *_gt hold ground truth."""

    intr: np.ndarray
    cam_q: np.ndarray
    cam_t: np.ndarray
    tag_q: np.ndarray
    tag_t: np.ndarray
    tag_wh: np.ndarray
    fixed_tag: int
    to_cam: np.ndarray
    to_tag: np.ndarray
    to_uv: np.ndarray
    cam_q_gt: Optional[np.ndarray] = None
    cam_t_gt: Optional[np.ndarray] = None
    tag_q_gt: Optional[np.ndarray] = None
    tag_t_gt: Optional[np.ndarray] = None
    outlier_mask: Optional[np.ndarray] = None
    meta: dict = field(default_factory=dict)

    @property
    def n_cameras(self) -> int:
        return int(self.cam_q.shape[0])

    @property
    def n_markers(self) -> int:
        return int(self.tag_q.shape[0])

    @property
    def n_tag_obs(self) -> int:
        return int(self.to_cam.shape[0])

    @property
    def n_corners(self) -> int:
        return 4 * self.n_tag_obs

    def copy(self) -> "Scene":
        kw = {}
        for k, v in self.__dict__.items():
            kw[k] = v.copy() if isinstance(v, np.ndarray) else (dict(v) if isinstance(v, dict) else v)
        return Scene(**kw)

    def at_ground_truth(self) -> "Scene":
        s = self.copy()
        s.cam_q, s.cam_t, s.tag_q, s.tag_t = self.cam_q_gt.copy(), self.cam_t_gt.copy(), self.tag_q_gt.copy(), self.tag_t_gt.copy()
        return s


def _corner_world(tag_q, tag_t, tag_wh):
    """world corners [Nm,4,3]"""
    R = quat_to_rot(tag_q)
    local = np.zeros((tag_q.shape[0], 4, 3))
    local[:, :, 0] = CORNER_SIGNS[:, 0][None, :] * tag_wh[:, 0:1] / 2.0
    local[:, :, 1] = CORNER_SIGNS[:, 1][None, :] * tag_wh[:, 1:2] / 2.0
    return np.einsum("mij,mkj->mki", R, local) + tag_t[:, None, :]


def make_scene(
    n_markers: int,
    n_cameras: int,
    markers_per_camera: int,
    seed: int,
    noise_px: float = 0.3,
    rot_sigma_deg: float = 1.0,
    trans_sigma: float = 0.03,
    outlier_frac: float = 0.0,
    relabel: str = "visible",
    height_range=(2.5, 3.5),
    max_tilt_deg: float = 12.0,
    marker_tilt_deg: float = 20.0,
    relief: float = 0.3,
    float32_round: bool = True,
    intr: Optional[np.ndarray] = None,
) -> Scene:
    from scipy.spatial import cKDTree

    rng = np.random.Generator(np.random.PCG64(seed))
    intr = README_INTRINSICS.copy() if intr is None else np.asarray(intr, dtype=np.float64)
    fx, fy, cx, cy = intr[:4]
    K = int(markers_per_camera)
    h_mid = 0.5 * (height_range[0] + height_range[1])
    # footprint of the image on the sheet at the lowest flying height, shrunk for tilt
    fp_w = IMAGE_W / fx * height_range[0] * 0.80
    fp_h = IMAGE_H / fy * height_range[0] * 0.80
    density = 1.35 * K / (fp_w * fp_h)
    # markers must not overlap: cap density, fly higher instead if needed
    max_density = 1.0 / (0.21 * 0.21)
    if density > max_density:
        scale_h = np.sqrt(density / max_density)
        height_range = (height_range[0] * scale_h, height_range[1] * scale_h)
        fp_w *= scale_h
        fp_h *= scale_h
        h_mid *= scale_h
        density = max_density
    area = n_markers / density
    L = float(np.sqrt(area))
    g = int(np.ceil(np.sqrt(n_markers)))
    cell = L / g
    # jittered grid: choose n_markers of the g*g cells
    cells = rng.permutation(g * g)[:n_markers]
    cells.sort()
    gx, gy = cells % g, cells // g
    jit = max(0.0, 0.5 * (cell - 0.20))
    mx = (gx + 0.5) * cell + rng.uniform(-jit, jit, n_markers)
    my = (gy + 0.5) * cell + rng.uniform(-jit, jit, n_markers)
    # the origin tag (index 0, fixed at the identity pose) must be one the cameras actually see — the
    # reference picks it from the start image (src/TagReconstructor.cpp:115-141) — otherwise the gauge
    # is free: take the marker nearest the centre of the sheet.
    centre = int(np.argmin((mx - 0.5 * L) ** 2 + (my - 0.5 * L) ** 2))
    for arr in (mx, my):
        arr[0], arr[centre] = arr[centre], arr[0]
    mz = relief * np.sin(mx * 1.3) * np.cos(my * 0.9) + rng.uniform(-0.05, 0.05, n_markers)
    tag_t = np.stack([mx, my, mz], axis=1)
    tilt = np.deg2rad(marker_tilt_deg) * np.sqrt(rng.uniform(0, 1, n_markers))
    tdir = rng.uniform(0, 2 * np.pi, n_markers)
    spin = rng.uniform(0, 2 * np.pi, n_markers)
    R_tilt = rotvec_to_rot(np.stack([np.cos(tdir) * tilt, np.sin(tdir) * tilt, np.zeros(n_markers)], axis=1))
    R_spin = rotvec_to_rot(np.stack([np.zeros(n_markers), np.zeros(n_markers), spin], axis=1))
    R_m = R_tilt @ R_spin  # marker z (normal) points up, tilted
    tag_wh = np.tile(np.array([[MARKER_W, MARKER_H]]), (n_markers, 1))

    # cameras: centre C, looking down (camera z = -world z), yaw + small tilt
    margin_x = min(0.5 * fp_w, 0.45 * L)
    margin_y = min(0.5 * fp_h, 0.45 * L)
    margin = max(margin_x, margin_y)
    margin = min(margin, 0.45 * L)
    Cx = rng.uniform(margin, L - margin, n_cameras)
    Cy = rng.uniform(margin, L - margin, n_cameras)
    Cz = rng.uniform(height_range[0], height_range[1], n_cameras)
    C = np.stack([Cx, Cy, Cz], axis=1)
    yaw = rng.uniform(0, 2 * np.pi, n_cameras)
    ctilt = np.deg2rad(max_tilt_deg) * np.sqrt(rng.uniform(0, 1, n_cameras))
    cdir = rng.uniform(0, 2 * np.pi, n_cameras)
    # camera-to-world rotation: flip (x right, y down, z forward=-Z world) then yaw, then tilt
    flip = np.array([[1.0, 0, 0], [0, -1.0, 0], [0, 0, -1.0]])
    R_yaw = rotvec_to_rot(np.stack([np.zeros(n_cameras), np.zeros(n_cameras), yaw], axis=1))
    R_ct = rotvec_to_rot(np.stack([np.cos(cdir) * ctilt, np.sin(cdir) * ctilt, np.zeros(n_cameras)], axis=1))
    R_c2w = R_ct @ R_yaw @ flip
    R_c = np.transpose(R_c2w, (0, 2, 1))  # world->camera
    t_c = -np.einsum("nij,nj->ni", R_c, C)

    # put marker 0 at the identity pose (origin tag): world' = T0^-1 world
    R0, t0 = R_m[0].copy(), tag_t[0].copy()
    R_m = np.einsum("ji,njk->nik", R0, R_m)
    tag_t = (tag_t - t0) @ R0
    # camera: Xc = R_c Xw + t_c, Xw = R0 Xw' + t0  ->  R_c' = R_c R0, t_c' = R_c t0 + t_c
    t_c = np.einsum("nij,j->ni", R_c, t0) + t_c
    R_c = R_c @ R0
    tag_q_gt = rot_to_quat(R_m)
    tag_q_gt[0] = np.array([1.0, 0, 0, 0])
    tag_t[0] = 0.0
    cam_q_gt = rot_to_quat(R_c)
    cam_t_gt = t_c
    tag_t_gt = tag_t

    # visibility: nearest candidates in the sheet plane around the optical-axis footprint
    corners_w = _corner_world(tag_q_gt, tag_t_gt, tag_wh)  # [Nm,4,3]
    normals = quat_to_rot(tag_q_gt)[:, :, 2]
    C_w = -np.einsum("nji,nj->ni", quat_to_rot(cam_q_gt), cam_t_gt)  # camera centres in the new world
    # intersection of the optical axis with the sheet (z'=0 plane of the original world is lost; use marker cloud)
    axis_w = quat_to_rot(cam_q_gt)[:, 2, :]  # third row of R_c = camera z axis in world coords
    # distance along the axis to the mean sheet: approximate with height h_mid
    foot = C_w + axis_w * (C[:, 2:3] / np.maximum(1e-6, np.cos(ctilt))[:, None])
    tree = cKDTree(tag_t_gt)
    n_query = int(min(n_markers, max(2 * K + 8, int(3.0 * K))))
    to_cam_l, to_tag_l, to_uv_l = [], [], []
    chunk = 4096
    Rc_all = quat_to_rot(cam_q_gt)
    for s in range(0, n_cameras, chunk):
        e = min(n_cameras, s + chunk)
        _, idx = tree.query(foot[s:e], k=n_query)
        if n_query == 1:
            idx = idx[:, None]
        cw = corners_w[idx]  # [B,Q,4,3]
        Xc = np.einsum("bij,bqkj->bqki", Rc_all[s:e], cw) + cam_t_gt[s:e, None, None, :]
        uv = project(intr, Xc)
        inside = (
            (Xc[..., 2] > 0.3)
            & (uv[..., 0] > 20)
            & (uv[..., 0] < IMAGE_W - 20)
            & (uv[..., 1] > 20)
            & (uv[..., 1] < IMAGE_H - 20)
        ).all(axis=2)
        # incidence angle < 70 deg
        view = C_w[s:e, None, :] - tag_t_gt[idx]
        view /= np.linalg.norm(view, axis=2, keepdims=True)
        cosang = np.einsum("bqi,bqi->bq", view, normals[idx])
        ok = inside & (cosang > np.cos(np.deg2rad(70.0)))
        rank = np.cumsum(ok, axis=1)
        keep = ok & (rank <= K)
        b, qd = np.nonzero(keep)
        to_cam_l.append((b + s).astype(np.int32))
        to_tag_l.append(idx[b, qd].astype(np.int32))
        to_uv_l.append(uv[b, qd].reshape(-1, 8))
    to_cam = np.concatenate(to_cam_l)
    to_tag = np.concatenate(to_tag_l)
    to_uv = np.concatenate(to_uv_l)
    # detector noise, then float32 precision
    to_uv = to_uv + rng.normal(0.0, noise_px, to_uv.shape) if noise_px > 0 else to_uv
    outlier_mask = np.zeros(to_cam.shape[0], dtype=bool)
    if outlier_frac > 0:
        n_out = int(round(outlier_frac * to_cam.shape[0]))
        sel = rng.permutation(to_cam.shape[0])[:n_out]
        outlier_mask[sel] = True
        half = n_out // 2
        shift_sel, relabel_sel = sel[:half], sel[half:]
        # half: common shift U(-40,40) px + N(0,5^2) per corner
        sh = rng.uniform(-40, 40, (shift_sel.size, 1, 2)) + rng.normal(0, 5.0, (shift_sel.size, 4, 2))
        to_uv[shift_sel] += sh.reshape(-1, 8)
        # half: re-labelled to another marker id.  "visible" (default): the id of another marker seen in the same
        # image — a decoding mix-up between detections, residuals of hundreds to thousands of pixels.  "any": a
        # random id of the whole map; on a site-sized sheet that marker is usually hundreds of metres away, its
        # projection (no Z check in the functor, TagReconstructionCostFunction.h:125-126) runs through the r^6
        # distortion term and the cost starts at 1e50: a stress case no loss function rescues, kept for tests.
        if relabel == "visible":
            order = np.argsort(to_cam, kind="stable")
            counts = np.bincount(to_cam, minlength=n_cameras)
            starts = np.concatenate([[0], np.cumsum(counts)[:-1]])
            c = to_cam[relabel_sel]
            r = rng.integers(0, np.maximum(counts[c], 1))
            cand = order[starts[c] + r]
            same = to_tag[cand] == to_tag[relabel_sel]
            cand = np.where(same, order[starts[c] + (r + 1) % np.maximum(counts[c], 1)], cand)
            new_ids = to_tag[cand].astype(np.int32)     # gathered from the unmodified labels
        else:
            new_ids = rng.integers(1, n_markers, relabel_sel.size).astype(np.int32)
            new_ids = np.where(new_ids == to_tag[relabel_sel], (new_ids % (n_markers - 1)) + 1, new_ids).astype(np.int32)
        to_tag[relabel_sel] = new_ids
    if float32_round:
        to_uv = to_uv.astype(np.float32).astype(np.float64)

    # initial guess = ground truth perturbed (stands in for the PnP initialisation)
    def perturb(q, t, fixed=None, about_centre=False):
        n = q.shape[0]
        rv = rng.normal(0.0, np.deg2rad(rot_sigma_deg), (n, 3))
        dt = rng.normal(0.0, trans_sigma, (n, 3))
        if fixed is not None:
            rv[fixed] = 0.0
            dt[fixed] = 0.0
        R = quat_to_rot(q)
        Rn = rotvec_to_rot(rv) @ R
        qn = rot_to_quat(Rn)
        if fixed is not None:
            qn[fixed] = q[fixed]
        if about_centre:
            # world->camera pose: rotate about the camera centre C = -R^T t and move the centre by dt,
            # which is what a PnP initialisation gets wrong (not a rotation about the far world origin)
            C = -np.einsum("nji,nj->ni", R, t)
            return qn, -np.einsum("nij,nj->ni", Rn, C + dt)
        return qn, t + dt

    if rot_sigma_deg > 0 or trans_sigma > 0:
        cam_q, cam_t = perturb(cam_q_gt, cam_t_gt, about_centre=True)
        tag_q, tag_t0 = perturb(tag_q_gt, tag_t_gt, fixed=0)
    else:
        cam_q, cam_t, tag_q, tag_t0 = cam_q_gt.copy(), cam_t_gt.copy(), tag_q_gt.copy(), tag_t_gt.copy()

    return Scene(
        intr=intr,
        cam_q=np.ascontiguousarray(cam_q),
        cam_t=np.ascontiguousarray(cam_t),
        tag_q=np.ascontiguousarray(tag_q),
        tag_t=np.ascontiguousarray(tag_t0),
        tag_wh=np.ascontiguousarray(tag_wh),
        fixed_tag=0,
        to_cam=np.ascontiguousarray(to_cam, dtype=np.int32),
        to_tag=np.ascontiguousarray(to_tag, dtype=np.int32),
        to_uv=np.ascontiguousarray(to_uv, dtype=np.float64),
        cam_q_gt=cam_q_gt,
        cam_t_gt=cam_t_gt,
        tag_q_gt=tag_q_gt,
        tag_t_gt=tag_t_gt,
        outlier_mask=outlier_mask,
        meta={
            "generator": GENERATOR_VERSION,
            "seed": int(seed),
            "n_markers": int(n_markers),
            "n_cameras": int(n_cameras),
            "markers_per_camera": K,
            "noise_px": float(noise_px),
            "outlier_frac": float(outlier_frac),
            "site_side_m": L,
        },
    )


# BASELINE.json configs. Seeds 0x564D4D00 + config number.
CONFIGS = {
    "c1": dict(n_markers=12, n_cameras=40, markers_per_camera=12, seed=0x564D4D01),
    "c2": dict(n_markers=500, n_cameras=2000, markers_per_camera=25, seed=0x564D4D02),
    "c3": dict(n_markers=20000, n_cameras=100000, markers_per_camera=40, seed=0x564D4D03),
    "c4": dict(n_markers=50000, n_cameras=200000, markers_per_camera=40, seed=0x564D4D04, outlier_frac=0.2),
}


def make_config(name: str, **overrides) -> Scene:
    """Scene for one of the BASELINE.json configurations: c1, c2, c3, c4 or 'c3/<divisor>'
    (the C3 generator scaled down by an integer divisor — used for bounded CPU samples)."""
    if "/" in name:
        base, div = name.split("/")
        kw = dict(CONFIGS[base])
        d = int(div)
        kw["n_markers"] = max(2, kw["n_markers"] // d)
        kw["n_cameras"] = max(2, kw["n_cameras"] // d)
    else:
        kw = dict(CONFIGS[name])
    kw.update(overrides)
    return make_scene(**kw)


def tile_scene(scene: Scene, copies: int) -> Scene:
    """`copies` disjoint replicas of a scene in one problem (camera / marker indices offset per replica, poses and
    observations repeated): evaluation-only workloads beyond what the generator produces in reasonable time (C5)."""
    if copies <= 1:
        return scene
    nc, nm = scene.n_cameras, scene.n_markers
    off_c = np.repeat(np.arange(copies, dtype=np.int64) * nc, scene.n_tag_obs)
    off_m = np.repeat(np.arange(copies, dtype=np.int64) * nm, scene.n_tag_obs)
    return Scene(
        intr=scene.intr, cam_q=np.tile(scene.cam_q, (copies, 1)), cam_t=np.tile(scene.cam_t, (copies, 1)),
        tag_q=np.tile(scene.tag_q, (copies, 1)), tag_t=np.tile(scene.tag_t, (copies, 1)), tag_wh=np.tile(scene.tag_wh, (copies, 1)),
        fixed_tag=scene.fixed_tag,
        to_cam=(np.tile(scene.to_cam.astype(np.int64), copies) + off_c).astype(np.int32),
        to_tag=(np.tile(scene.to_tag.astype(np.int64), copies) + off_m).astype(np.int32),
        to_uv=np.tile(scene.to_uv, (copies,) + (1,) * (scene.to_uv.ndim - 1)),
        meta=dict(scene.meta, tiled=copies))


def make_rj_sweep_scene(n_corners: int, seed: int = 0x564D4D05) -> Scene:
    """C5: evaluation-only scene with ~n_corners corner observations at C3 density
    (40 markers per camera, 5 cameras per marker-equivalent)."""
    n_to = max(1, n_corners // 4)
    n_cam = max(1, n_to // 40)
    n_mark = max(48, n_cam // 5)
    return make_scene(n_markers=n_mark, n_cameras=n_cam, markers_per_camera=40, seed=seed)
