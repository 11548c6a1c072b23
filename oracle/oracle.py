"""ctypes binding of the CPU oracle (oracle/vmm_oracle.cpp).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs. The product package (visual_marker_mapping_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libvmm_oracle.so")

LOSS_NONE, LOSS_HUBER, LOSS_CAUCHY = 0, 1, 2


class Options(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int),
        ("loss_kind", C.c_int),
        ("loss_scale", C.c_double),
        ("function_tolerance", C.c_double),
        ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double),
        ("initial_trust_region_radius", C.c_double),
        ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double),
        ("min_relative_decrease", C.c_double),
        ("min_lm_diagonal", C.c_double),
        ("max_lm_diagonal", C.c_double),
        ("max_num_consecutive_invalid_steps", C.c_int),
        ("jacobi_scaling", C.c_int),
        ("linear_solver", C.c_int),          # 0 block-envelope Cholesky (RCM marker order), 1 dense (cross-check)
        ("max_solver_time_s", C.c_double),   # Ceres' max_solver_time_in_seconds
    ]


SOLVER_ENVELOPE, SOLVER_DENSE = 0, 1


class Summary(C.Structure):
    _fields_ = [
        ("termination_type", C.c_int),
        ("initial_cost", C.c_double),
        ("final_cost", C.c_double),
        ("num_successful_steps", C.c_int),
        ("num_unsuccessful_steps", C.c_int),
        ("num_iterations", C.c_int),
        ("time_total_s", C.c_double),
        ("time_evaluate_s", C.c_double),
        ("time_linear_solver_s", C.c_double),
        ("num_jacobian_evals", C.c_int64),
        ("num_residual_evals", C.c_int64),
        ("message", C.c_char * 256),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only, no reference build system)."""
    src = os.path.join(_HERE, "vmm_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.vmo_num_threads.restype = C.c_int
    return _lib


def _p(a, typ=C.c_double):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(typ))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def default_options(**kw) -> Options:
    o = Options()
    lib().vmo_options_init(C.byref(o))
    for k, v in kw.items():
        setattr(o, k, v)
    return o


def num_threads() -> int:
    return int(lib().vmo_num_threads())


def set_num_threads(n: int) -> None:
    lib().vmo_set_num_threads(C.c_int(n))


@dataclass
class Evaluation:
    cost: float
    residuals: Optional[np.ndarray]  # [Nobs, 2]
    jac_local: Optional[np.ndarray]  # [Nobs, 2, 12]
    jac_global: Optional[np.ndarray]  # [Nobs, 2, 14]
    gradient: Optional[np.ndarray]  # [6Nc + 6Nm]


def _scene_args(s):
    cam_q, cam_t = _f64(s.cam_q), _f64(s.cam_t)
    tag_q, tag_t, tag_wh = _f64(s.tag_q), _f64(s.tag_t), _f64(s.tag_wh)
    to_cam, to_tag, to_uv = _i32(s.to_cam), _i32(s.to_tag), _f64(s.to_uv)
    intr = _f64(s.intr)
    keep = (intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv)
    return keep


def evaluate(s, loss_kind=LOSS_NONE, loss_scale=1.0, residuals=True, jac_local=True, jac_global=False, gradient=True) -> Evaluation:
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    Nc, Nm, Nto = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
    cost = C.c_double(0)
    r = np.zeros((4 * Nto, 2)) if residuals else None
    jl = np.zeros((4 * Nto, 2, 12)) if jac_local else None
    jg = np.zeros((4 * Nto, 2, 14)) if jac_global else None
    g = np.zeros(6 * (Nc + Nm)) if gradient else None
    rc = lib().vmo_evaluate(
        _p(intr), C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(tag_wh),
        C.c_int64(s.fixed_tag), C.c_int64(Nto), _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(to_uv),
        C.c_int(loss_kind), C.c_double(loss_scale), C.byref(cost), _p(r), _p(jl), _p(jg), _p(g),
    )
    assert rc == 0
    return Evaluation(cost.value, r, jl, jg, g)


def plus(s, delta):
    """Apply Ceres' Plus to every block; returns (cam_q, cam_t, tag_q, tag_t)."""
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    Nc, Nm = cam_q.shape[0], tag_q.shape[0]
    d = _f64(delta)
    assert d.shape[0] == 6 * (Nc + Nm)
    oq, ot, mq, mt = np.empty_like(cam_q), np.empty_like(cam_t), np.empty_like(tag_q), np.empty_like(tag_t)
    lib().vmo_plus(C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(d), _p(oq), _p(ot), _p(mq), _p(mt))
    return oq, ot, mq, mt


def reprojection_error_sums(s) -> np.ndarray:
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    Nc, Nm, Nto = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
    out = np.zeros(Nto)
    lib().vmo_reprojection_error_sums(
        _p(intr), C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(tag_wh),
        C.c_int64(Nto), _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(to_uv), _p(out),
    )
    return out


@dataclass
class SolveResult:
    cam_q: np.ndarray
    cam_t: np.ndarray
    tag_q: np.ndarray
    tag_t: np.ndarray
    summary: Summary
    trace: np.ndarray  # [num_iterations, 6]: cost, gradient_max_norm, step_norm, relative_decrease, radius, successful


def solve(s, options: Optional[Options] = None, **kw) -> SolveResult:
    """Ceres-semantics LM with an exact linear solve; does not modify ``s``."""
    opt = options if options is not None else default_options(**kw)
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    cam_q, cam_t, tag_q, tag_t = cam_q.copy(), cam_t.copy(), tag_q.copy(), tag_t.copy()
    Nc, Nm, Nto = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
    summ = Summary()
    trace = np.zeros((opt.max_num_iterations + 2, 6))
    rc = lib().vmo_solve(
        _p(intr), C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(tag_wh),
        C.c_int64(s.fixed_tag), C.c_int64(Nto), _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(to_uv),
        C.byref(opt), C.byref(summ), _p(trace),
    )
    assert rc == 0
    return SolveResult(cam_q, cam_t, tag_q, tag_t, summ, trace[: summ.num_iterations])


def lm_step_residual(s, delta, radius=1e4, loss_kind=LOSS_NONE, loss_scale=1.0, jacobi_scaling=True,
                     min_lm_diagonal=1e-6, max_lm_diagonal=1e32) -> dict:
    """Matrix-free check of one LM step ``delta`` (tangent, [6Nc+6Nm], cameras first) computed elsewhere for the FIRST
    iteration at the scene's poses: residual of the damped, Jacobi-scaled normal equations from the oracle's own
    dual-number Jacobian (no factorisation; linear in the corner count, so it runs at C3/C4 scale)."""
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    Nc, Nm, Nto = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
    d = _f64(delta).reshape(-1)
    assert d.shape[0] == 6 * (Nc + Nm)
    out = np.zeros(4)
    rc = lib().vmo_lm_step_residual(
        _p(intr), C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(tag_wh),
        C.c_int64(s.fixed_tag), C.c_int64(Nto), _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(to_uv),
        C.c_int(loss_kind), C.c_double(loss_scale), C.c_int(1 if jacobi_scaling else 0), C.c_double(min_lm_diagonal),
        C.c_double(max_lm_diagonal), C.c_double(radius), _p(d), _p(out))
    assert rc == 0
    return {"residual_norm": out[0], "gradient_norm": out[1], "relative": out[0] / out[1] if out[1] > 0 else out[0],
            "model_cost_change": out[2], "cost": out[3]}


def marker_translation_covariance_sparse(s, markers, loss_kind=LOSS_NONE, loss_scale=1.0) -> np.ndarray:
    """3x3 translation blocks of (J^T J)^-1 for the listed markers through the envelope factorisation of the
    marker-reduced matrix (site scale; src/TagReconstructor.cpp:744-783).  [len(markers), 3, 3]."""
    intr, cam_q, cam_t, tag_q, tag_t, tag_wh, to_cam, to_tag, to_uv = _scene_args(s)
    Nc, Nm, Nto = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
    sel = np.ascontiguousarray(markers, dtype=np.int64)
    out = np.zeros((sel.shape[0], 3, 3))
    rc = lib().vmo_marker_translation_covariance_sparse(
        _p(intr), C.c_int64(Nc), _p(cam_q), _p(cam_t), C.c_int64(Nm), _p(tag_q), _p(tag_t), _p(tag_wh),
        C.c_int64(s.fixed_tag), C.c_int64(Nto), _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(to_uv),
        C.c_int(loss_kind), C.c_double(loss_scale), C.c_int64(sel.shape[0]), _p(sel, C.c_int64), _p(out))
    if rc != 0:
        raise RuntimeError("the normal equations are rank deficient")
    return out


def sparse_structure(s) -> dict:
    """Size of the envelope factor the sparse path would build for this scene (no numbers are computed)."""
    to_cam, to_tag = _i32(s.to_cam), _i32(s.to_tag)
    out = np.zeros(5)
    lib().vmo_sparse_structure(C.c_int64(s.n_cameras), C.c_int64(s.n_markers), C.c_int64(s.fixed_tag), C.c_int64(to_cam.shape[0]),
                               _p(to_cam, C.c_int32), _p(to_tag, C.c_int32), _p(out))
    return {"free_markers": int(out[0]), "order": int(out[1]), "blocks": int(out[2]), "factor_bytes": out[3], "factor_flops": out[4]}


def marker_translation_covariance(s, loss_kind=LOSS_NONE, loss_scale=1.0) -> np.ndarray:
    """3x3 blocks of (J^T J)^-1 for every marker's translation, as ceres::Covariance delivers them at
    src/TagReconstructor.cpp:744-783: the FULL tangent-space normal matrix (cameras and markers, 6 dofs per pose,
    loss-corrected Jacobians) is formed densely and inverted; the constant origin tag has no columns and gets a zero
    block, like every constant parameter block in Ceres.  Dense — small scenes only.  numpy restatement, no
    elimination trick: an independent check of the engine's Schur-reduced computation."""
    ev = evaluate(s, loss_kind, loss_scale, residuals=False, jac_local=True, gradient=False)
    Nc, Nm = s.n_cameras, s.n_markers
    J = ev.jac_local.reshape(-1, 4, 2, 12)                       # per tag observation: 4 corners x 2 rows x 12 columns
    n = 6 * (Nc + Nm)
    H = np.zeros((n, n))
    cam = np.asarray(s.to_cam, dtype=np.int64)
    tag = np.asarray(s.to_tag, dtype=np.int64)
    for k in range(J.shape[0]):
        Jk = J[k].reshape(8, 12)
        cols = np.concatenate([6 * cam[k] + np.arange(6), 6 * (Nc + tag[k]) + np.arange(6)])
        H[np.ix_(cols, cols)] += Jk.T @ Jk
    free = np.ones(n, dtype=bool)
    observed_m = np.zeros(Nm, dtype=bool)
    observed_m[tag] = True
    observed_c = np.zeros(Nc, dtype=bool)
    observed_c[cam] = True
    for j in range(Nm):
        if j == s.fixed_tag or not observed_m[j]:
            free[6 * (Nc + j):6 * (Nc + j) + 6] = False
    for i in range(Nc):
        if not observed_c[i]:
            free[6 * i:6 * i + 6] = False
    idx = np.flatnonzero(free)
    Hinv = np.linalg.inv(H[np.ix_(idx, idx)])
    pos = -np.ones(n, dtype=np.int64)
    pos[idx] = np.arange(idx.size)
    cov = np.zeros((Nm, 3, 3))
    for j in range(Nm):
        p = pos[6 * (Nc + j):6 * (Nc + j) + 3]
        if (p >= 0).all():
            cov[j] = Hinv[np.ix_(p, p)]
    return cov
