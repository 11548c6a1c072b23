"""Project directories in the reference's on-disk formats: camera_intrinsics.json,
marker_detections.json, reconstruction.json — written in the boost::property_tree write_json dialect
(every leaf a quoted string, doubles with 17 significant digits, 4-space indentation, empty array = "").

Python plumbing for tests and benchmarks; the product reader/writer is the C++ one in csrc/host/.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Tuple

import numpy as np

from . import synth

TAG_TYPE = "apriltag_36h11"


def _esc(s: str) -> str:
    return s.replace("\\", "\\\\").replace('"', '\\"').replace("/", "\\/")


def _num(v) -> str:
    if isinstance(v, (int, np.integer)):
        return str(int(v))
    return "%.17g" % float(v)


def _write(node, indent: int, out: list):
    """node: str leaf | dict (object, insertion order) | list (array); boost write_json_helper layout."""
    if isinstance(node, str):
        out.append('"' + _esc(node) + '"')
        return
    is_obj = isinstance(node, dict)
    items = list(node.items()) if is_obj else [(None, v) for v in node]
    if indent > 0 and not items:       # an empty ptree node has no children and empty data
        out.append('""')
        return
    pad = " " * (4 * (indent + 1))
    out.append("{\n" if is_obj else "[\n")
    for i, (k, v) in enumerate(items):
        out.append(pad)
        if is_obj:
            out.append('"' + _esc(k) + '": ')
        _write(v, indent + 1, out)
        out.append(",\n" if i + 1 < len(items) else "\n")
    out.append(" " * (4 * indent) + ("}" if is_obj else "]"))


def dumps_ptree(root: dict) -> str:
    out: list = []
    _write(root, 0, out)
    return "".join(out) + "\n"


def _vec(a):
    return [_num(x) for x in a]


def camera_model_tree(intr) -> dict:
    return {"fx": _num(intr[0]), "fy": _num(intr[1]), "cx": _num(intr[2]), "cy": _num(intr[3]),
            "distortion_coefficients": _vec(intr[4:9]), "vertical_resolution": _num(synth.IMAGE_H),
            "horizontal_resolution": _num(synth.IMAGE_W)}


def reconstruction_tree(intr, cam_ids, cam_q, cam_t, tag_ids, tag_q, tag_t, tag_wh) -> dict:
    tags, corners, cams = [], [], []
    R = synth.quat_to_rot_raw(tag_q)
    for j, tid in enumerate(tag_ids):
        tags.append({"id": _num(tid), "type": TAG_TYPE, "width": _num(tag_wh[j, 0]), "height": _num(tag_wh[j, 1]),
                     "rotation": _vec(tag_q[j]), "translation": _vec(tag_t[j])})
        for c in range(4):
            p = np.array([synth.CORNER_SIGNS[c, 0] * tag_wh[j, 0] / 2.0, synth.CORNER_SIGNS[c, 1] * tag_wh[j, 1] / 2.0, 0.0])
            corners.append({"marker_id": _num(tid), "corner_index": _num(c), "coords": _vec(R[j] @ p + tag_t[j])})
    for i, cid in enumerate(cam_ids):
        cams.append({"id": _num(cid), "rotation": _vec(cam_q[i]), "translation": _vec(cam_t[i])})
    return {"reconstructed_tags": tags, "reconstructed_marker_corners": corners, "reconstructed_cameras": cams,
            "camera_model": camera_model_tree(intr)}


def write_project(scene: synth.Scene, path: str, image_id_offset: int = 0, tag_id_offset: int = 0) -> None:
    """camera_intrinsics.json + marker_detections.json + initial_reconstruction.json (the scene's initial guess)."""
    os.makedirs(path, exist_ok=True)
    with open(os.path.join(path, "camera_intrinsics.json"), "w") as f:
        f.write(dumps_ptree(camera_model_tree(scene.intr)))
    cam_ids = np.arange(scene.n_cameras) + image_id_offset
    tag_ids = np.arange(scene.n_markers) + tag_id_offset
    det = {
        "images": [{"filename": "images/img_%06d.jpg" % i, "id": _num(i)} for i in cam_ids],
        "tags": [{"id": _num(t), "tag_type": TAG_TYPE, "width": _num(scene.tag_wh[j, 0]), "height": _num(scene.tag_wh[j, 1])}
                 for j, t in enumerate(tag_ids)],
        "tag_observations": [
            {"image_id": _num(cam_ids[scene.to_cam[k]]), "tag_id": _num(tag_ids[scene.to_tag[k]]),
             "observations": [[_num(scene.to_uv[k, 2 * c]), _num(scene.to_uv[k, 2 * c + 1])] for c in range(4)]}
            for k in range(scene.n_tag_obs)],
    }
    with open(os.path.join(path, "marker_detections.json"), "w") as f:
        f.write(dumps_ptree(det))
    with open(os.path.join(path, "initial_reconstruction.json"), "w") as f:
        f.write(dumps_ptree(reconstruction_tree(scene.intr, cam_ids, scene.cam_q, scene.cam_t, tag_ids, scene.tag_q, scene.tag_t,
                                                scene.tag_wh)))


def read_reconstruction(path: str) -> Tuple[Dict[int, dict], Dict[int, dict]]:
    """{tag id: {q, t, width, height}}, {camera id: {q, t}} from a reconstruction.json (quoted or bare numbers)."""
    with open(path) as f:
        root = json.load(f)

    def arr(v):
        return np.array([float(x) for x in v]) if v != "" else np.zeros(0)

    tags = {int(t["id"]): {"q": arr(t["rotation"]), "t": arr(t["translation"]), "width": float(t["width"]),
                           "height": float(t["height"])} for t in (root["reconstructed_tags"] or [])}
    cams = {int(c["id"]): {"q": arr(c["rotation"]), "t": arr(c["translation"])} for c in (root["reconstructed_cameras"] or [])}
    return tags, cams
