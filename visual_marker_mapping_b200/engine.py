"""ctypes binding of libvmm_ba.so — the thin Python view of the C ABI in include/vmm_ba.h.

Python is plumbing here (tests, bench.py, torchrun ranks); the product is the CUDA library and the
C++ facade in csrc/host/.  There is no fallback: if the library is missing or no sm_100a device is
present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import _build

LOSS_NONE, LOSS_HUBER, LOSS_CAUCHY = 0, 1, 2
CONVERGENCE, NO_CONVERGENCE, FAILURE = 0, 1, 2
PRECOND_JACOBI, PRECOND_SCHUR_JACOBI, PRECOND_TWO_LEVEL = 0, 1, 2
COV_AUTO, COV_DENSE, COV_BLOCK_TRIDIAGONAL = 0, 1, 2

# every symbol include/vmm_ba.h declares
ABI_SYMBOLS = [
    "vmm_ba_options_init", "vmm_ba_abi_version", "vmm_ba_device_count", "vmm_ba_create",
    "vmm_ba_create_distributed", "vmm_ba_nccl_unique_id", "vmm_ba_destroy", "vmm_ba_last_error",
    "vmm_ba_set_stream", "vmm_ba_set_intrinsics", "vmm_ba_set_problem", "vmm_ba_set_poses",
    "vmm_ba_get_poses", "vmm_ba_get_step", "vmm_ba_solve", "vmm_ba_evaluate", "vmm_ba_evaluate_device",
    "vmm_ba_reprojection_error_sums", "vmm_ba_marker_translation_covariance", "vmm_ba_marker_translation_covariance_ex", "vmm_ba_schur_multiply", "vmm_ba_debug_coarse_matrix", "vmm_ba_set_coarse_cluster_size", "vmm_ba_shard_info", "vmm_ba_partition_cameras", "vmm_ba_shard_plan",
    "vmm_ba_kernel_launches", "vmm_ba_h2d_bytes", "vmm_ba_measure_fp64_peak", "vmm_ba_measure_copy_bandwidth",
    "vmm_ba_pcg_phase_ns", "vmm_ba_pcg_cta_product_ns", "vmm_ba_profile_enable", "vmm_ba_profile_num_kernels", "vmm_ba_profile_kernel_name", "vmm_ba_profile_get",
    "vmm_ba_cost_no_alloc",
]
NO_ALLOC_MAX_CAMERAS, NO_ALLOC_MAX_MARKERS, NO_ALLOC_MAX_TAG_OBS = 4096, 1024, 131072


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
        ("preconditioner", C.c_int),
        ("pcg_tolerance", C.c_double),
        ("pcg_max_iterations", C.c_int),
        ("pcg_min_iterations", C.c_int),
        ("parity_mode", C.c_int),
        ("pcg_parity_tolerance", C.c_double),
        ("pcg_parity_max_iterations", C.c_int),
        ("verbose", C.c_int),
        ("coarse_refresh_iterations", C.c_int),
        ("pcg_confirm_tolerance", C.c_double),
        ("pcg_polish_tolerance", C.c_double),
        ("pcg_polish_threshold", C.c_double),
    ]


class Summary(C.Structure):
    _fields_ = [
        ("termination_type", C.c_int),
        ("initial_cost", C.c_double),
        ("final_cost", C.c_double),
        ("num_successful_steps", C.c_int),
        ("num_unsuccessful_steps", C.c_int),
        ("num_iterations", C.c_int),
        ("num_linearizations", C.c_int64),
        ("num_cost_evaluations", C.c_int64),
        ("pcg_iterations", C.c_int64),
        ("kernel_launches", C.c_int64),
        ("time_total_s", C.c_double),
        ("time_linearize_s", C.c_double),
        ("time_linear_solver_s", C.c_double),
        ("time_cost_eval_s", C.c_double),
        ("gradient_max_norm", C.c_double),
        ("trust_region_radius", C.c_double),
        ("num_confirmations", C.c_int),
        ("reserved", C.c_int),
        ("message", C.c_char * 256),
    ]

    def as_dict(self) -> dict:
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["message"] = self.message.decode()
        return d


class VmmBaError(RuntimeError):
    pass


_lib = None


def library_path() -> str:
    return _build.LIB_PATH


def load_library():
    """dlopen libvmm_ba.so; raises (loudly) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("VMM_BA_LIBRARY", _build.LIB_PATH)   # override: kernel-variant experiments (tools/)
    if not os.path.exists(path):
        raise VmmBaError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback."
        )
    lib = C.CDLL(path)
    lib.vmm_ba_last_error.restype = C.c_char_p
    lib.vmm_ba_last_error.argtypes = [C.c_void_p]
    lib.vmm_ba_kernel_launches.restype = C.c_int64
    lib.vmm_ba_kernel_launches.argtypes = [C.c_void_p]
    lib.vmm_ba_h2d_bytes.restype = C.c_int64
    lib.vmm_ba_h2d_bytes.argtypes = [C.c_void_p]
    lib.vmm_ba_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    lib.vmm_ba_create_distributed.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
    lib.vmm_ba_destroy.argtypes = [C.c_void_p]
    lib.vmm_ba_set_stream.argtypes = [C.c_void_p, C.c_void_p]
    lib.vmm_ba_set_intrinsics.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_double)]
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)
    lib.vmm_ba_set_problem.argtypes = [C.c_void_p, C.c_int64, dp, dp, C.c_int64, dp, dp, dp, C.c_int64, C.c_int64, ip, ip, dp]
    lib.vmm_ba_set_poses.argtypes = [C.c_void_p, dp, dp, dp, dp]
    lib.vmm_ba_get_poses.argtypes = [C.c_void_p, dp, dp, dp, dp]
    lib.vmm_ba_get_step.argtypes = [C.c_void_p, dp, dp]
    lib.vmm_ba_solve.argtypes = [C.c_void_p, C.POINTER(Options), C.POINTER(Summary)]
    lib.vmm_ba_evaluate.argtypes = [C.c_void_p, C.c_int, C.c_double, dp, dp, dp, dp]
    lib.vmm_ba_evaluate_device.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, dp]
    lib.vmm_ba_reprojection_error_sums.argtypes = [C.c_void_p, dp]
    lib.vmm_ba_marker_translation_covariance.argtypes = [C.c_void_p, C.c_int, C.c_double, dp]
    lib.vmm_ba_marker_translation_covariance_ex.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_int, dp]
    lib.vmm_ba_set_coarse_cluster_size.argtypes = [C.c_void_p, C.c_int]
    lib.vmm_ba_schur_multiply.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, dp, dp]
    lib.vmm_ba_debug_coarse_matrix.argtypes = [C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int64, dp, dp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.vmm_ba_shard_info.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.vmm_ba_partition_cameras.argtypes = [C.c_int64, C.c_int64, ip, C.c_int, C.POINTER(C.c_int64)]
    lib.vmm_ba_shard_plan.argtypes = [C.c_int64, dp, dp, C.c_int64, C.c_int64, ip, ip, C.c_int, C.c_int, ip, ip]
    lib.vmm_ba_shard_plan.restype = C.c_int
    lib.vmm_ba_measure_fp64_peak.argtypes = [C.c_void_p, dp]
    lib.vmm_ba_measure_copy_bandwidth.argtypes = [C.c_void_p, C.c_size_t, dp]
    lib.vmm_ba_nccl_unique_id.argtypes = [C.c_void_p, C.c_size_t]
    lib.vmm_ba_pcg_phase_ns.argtypes = [C.c_void_p, dp]
    lib.vmm_ba_pcg_cta_product_ns.argtypes = [C.c_void_p, dp, C.c_int]
    lib.vmm_ba_profile_enable.argtypes = [C.c_void_p, C.c_int]
    lib.vmm_ba_profile_kernel_name.restype = C.c_char_p
    lib.vmm_ba_profile_kernel_name.argtypes = [C.c_int]
    lib.vmm_ba_profile_get.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64), dp]
    vp = C.c_void_p
    lib.vmm_ba_cost_no_alloc.argtypes = [C.c_int, vp, C.c_double, C.c_double, C.c_double, C.c_double, dp, C.c_int64, vp, vp, C.c_int64, vp, vp, vp,
                                         C.c_int64, vp, vp, vp, C.c_int, C.c_double, vp]
    _lib = lib
    return lib


def default_options(**kw) -> Options:
    o = Options()
    load_library().vmm_ba_options_init(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def cost_no_alloc(device: int, stream: int, intrinsics, num_cameras: int, d_cam_q: int, d_cam_t: int, num_markers: int, d_tag_q: int,
                  d_tag_t: int, d_tag_wh: int, num_tag_obs: int, d_to_cam: int, d_to_tag: int, d_to_uv: int, loss_kind: int, loss_scale: float,
                  d_cost: int) -> None:
    """vmm_ba_cost_no_alloc: the robust cost on the caller's DEVICE buffers (raw device addresses as ints), scratch in the
    library's static device array; allocates nothing, does not synchronise.  Raises VmmBaError beyond the static limits."""
    lib = load_library()
    intr = np.ascontiguousarray(intrinsics, dtype=np.float64)
    dist = np.ascontiguousarray(intr[4:9])
    rc = lib.vmm_ba_cost_no_alloc(int(device), C.c_void_p(stream or None), float(intr[0]), float(intr[1]), float(intr[2]), float(intr[3]), _dp(dist),
                                  int(num_cameras), C.c_void_p(d_cam_q or None), C.c_void_p(d_cam_t or None), int(num_markers), C.c_void_p(d_tag_q or None),
                                  C.c_void_p(d_tag_t or None), C.c_void_p(d_tag_wh or None), int(num_tag_obs), C.c_void_p(d_to_cam or None),
                                  C.c_void_p(d_to_tag or None), C.c_void_p(d_to_uv or None), int(loss_kind), float(loss_scale), C.c_void_p(d_cost or None))
    if rc != 0:
        raise VmmBaError(f"status {rc}: {lib.vmm_ba_last_error(None).decode()}")


def device_count() -> int:
    return int(load_library().vmm_ba_device_count())


def partition_cameras(num_cameras: int, to_cam: np.ndarray, world_size: int) -> np.ndarray:
    """Host-side shard map (no GPU needed): world_size+1 camera boundaries."""
    to_cam = np.ascontiguousarray(to_cam, dtype=np.int32)
    out = np.zeros(world_size + 1, dtype=np.int64)
    rc = load_library().vmm_ba_partition_cameras(
        num_cameras, to_cam.shape[0], to_cam.ctypes.data_as(C.POINTER(C.c_int32)), world_size,
        out.ctypes.data_as(C.POINTER(C.c_int64)))
    if rc != 0:
        raise VmmBaError(f"vmm_ba_partition_cameras failed with status {rc}")
    return out


def shard_plan(cam_q, cam_t, num_markers: int, to_cam, to_tag, world_size: int, cameras_per_cluster: int = 0):
    """The sharding plan of a distributed engine (no GPU needed): (owner rank per camera, per marker -1 when all its
    blocks live on one rank else its position in the all-reduced boundary vector)."""
    cam_q, cam_t = _f64(cam_q), _f64(cam_t)
    to_cam = np.ascontiguousarray(to_cam, dtype=np.int32)
    to_tag = np.ascontiguousarray(to_tag, dtype=np.int32)
    cam_rank = np.zeros(cam_q.shape[0], dtype=np.int32)
    boundary = np.zeros(num_markers, dtype=np.int32)
    ip = C.POINTER(C.c_int32)
    rc = load_library().vmm_ba_shard_plan(cam_q.shape[0], _dp(cam_q), _dp(cam_t), num_markers, to_cam.shape[0], to_cam.ctypes.data_as(ip),
                                          to_tag.ctypes.data_as(ip), world_size, cameras_per_cluster, cam_rank.ctypes.data_as(ip),
                                          boundary.ctypes.data_as(ip))
    if rc != 0:
        raise VmmBaError(f"vmm_ba_shard_plan failed with status {rc}")
    return cam_rank, boundary


def nccl_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    rc = load_library().vmm_ba_nccl_unique_id(buf, 128)
    if rc != 0:
        raise VmmBaError(load_library().vmm_ba_last_error(None).decode())
    return buf.raw


def _dp(a):
    return None if a is None else a.ctypes.data_as(C.POINTER(C.c_double))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


@dataclass
class Evaluation:
    cost: float
    residuals: Optional[np.ndarray]
    jacobians: Optional[np.ndarray]
    gradient: Optional[np.ndarray]


class BundleAdjuster:
    """One engine on one GPU (optionally one rank of a distributed solve)."""

    def __init__(self, device: int = 0, rank: int = 0, world_size: int = 1, nccl_id: Optional[bytes] = None):
        self._lib = load_library()
        self._h = C.c_void_p()
        if world_size > 1:
            rc = self._lib.vmm_ba_create_distributed(device, rank, world_size, nccl_id, len(nccl_id), C.byref(self._h))
        else:
            rc = self._lib.vmm_ba_create(device, C.byref(self._h))
        if rc != 0:
            raise VmmBaError(self._lib.vmm_ba_last_error(None).decode())
        self.rank, self.world_size = rank, world_size
        self.n_cameras = self.n_markers = self.n_tag_obs = 0

    def _check(self, rc):
        if rc != 0:
            raise VmmBaError(f"status {rc}: " + self._lib.vmm_ba_last_error(self._h).decode())

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.vmm_ba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_stream(self, cuda_stream_ptr: int):
        self._check(self._lib.vmm_ba_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def set_intrinsics(self, intr):
        intr = _f64(intr)
        dist = _f64(intr[4:9])
        self._check(self._lib.vmm_ba_set_intrinsics(self._h, intr[0], intr[1], intr[2], intr[3], _dp(dist)))

    def set_problem(self, cam_q, cam_t, tag_q, tag_t, tag_wh, fixed_tag, to_cam, to_tag, to_uv):
        cam_q, cam_t, tag_q, tag_t, tag_wh, to_uv = map(_f64, (cam_q, cam_t, tag_q, tag_t, tag_wh, to_uv))
        to_cam = np.ascontiguousarray(to_cam, dtype=np.int32)
        to_tag = np.ascontiguousarray(to_tag, dtype=np.int32)
        self.n_cameras, self.n_markers, self.n_tag_obs = cam_q.shape[0], tag_q.shape[0], to_cam.shape[0]
        ip = C.POINTER(C.c_int32)
        self._check(self._lib.vmm_ba_set_problem(
            self._h, self.n_cameras, _dp(cam_q), _dp(cam_t), self.n_markers, _dp(tag_q), _dp(tag_t), _dp(tag_wh),
            int(fixed_tag), self.n_tag_obs, to_cam.ctypes.data_as(ip), to_tag.ctypes.data_as(ip), _dp(to_uv)))

    def set_scene(self, scene):
        self.set_intrinsics(scene.intr)
        self.set_problem(scene.cam_q, scene.cam_t, scene.tag_q, scene.tag_t, scene.tag_wh, scene.fixed_tag,
                         scene.to_cam, scene.to_tag, scene.to_uv)

    def set_poses(self, cam_q, cam_t, tag_q, tag_t):
        cam_q, cam_t, tag_q, tag_t = map(_f64, (cam_q, cam_t, tag_q, tag_t))
        self._check(self._lib.vmm_ba_set_poses(self._h, _dp(cam_q), _dp(cam_t), _dp(tag_q), _dp(tag_t)))

    def get_poses(self):
        cq, ct = np.zeros((self.n_cameras, 4)), np.zeros((self.n_cameras, 3))
        tq, tt = np.zeros((self.n_markers, 4)), np.zeros((self.n_markers, 3))
        self._check(self._lib.vmm_ba_get_poses(self._h, _dp(cq), _dp(ct), _dp(tq), _dp(tt)))
        return cq, ct, tq, tt

    def get_step(self) -> np.ndarray:
        """Tangent step of the last LM iteration, [6 * (n_cameras + n_markers)], cameras first (the oracle's layout)."""
        d = np.zeros(6 * (self.n_cameras + self.n_markers))
        self._check(self._lib.vmm_ba_get_step(self._h, _dp(d[:6 * self.n_cameras]), _dp(d[6 * self.n_cameras:])))
        return d

    def solve(self, options: Optional[Options] = None, **kw) -> Summary:
        opt = options if options is not None else default_options(**kw)
        s = Summary()
        self._check(self._lib.vmm_ba_solve(self._h, C.byref(opt), C.byref(s)))
        return s

    def evaluate(self, loss_kind=LOSS_NONE, loss_scale=1.0, residuals=True, jacobians=True, gradient=True) -> Evaluation:
        nobs = 4 * self.n_tag_obs
        cost = C.c_double(0)
        r = np.zeros((nobs, 2)) if residuals else None
        j = np.zeros((nobs, 2, 12)) if jacobians else None
        g = np.zeros(6 * (self.n_cameras + self.n_markers)) if gradient else None
        self._check(self._lib.vmm_ba_evaluate(self._h, loss_kind, loss_scale, C.byref(cost), _dp(r), _dp(j), _dp(g)))
        return Evaluation(cost.value, r, j, g)

    def evaluate_device(self, loss_kind=LOSS_NONE, loss_scale=1.0, iterations=1) -> float:
        ms = C.c_double(0)
        self._check(self._lib.vmm_ba_evaluate_device(self._h, loss_kind, loss_scale, iterations, C.byref(ms)))
        return ms.value

    def reprojection_error_sums(self) -> np.ndarray:
        out = np.zeros(self.n_tag_obs)
        self._check(self._lib.vmm_ba_reprojection_error_sums(self._h, _dp(out)))
        return out

    def marker_translation_covariance(self, loss_kind=LOSS_NONE, loss_scale=1.0, method=COV_AUTO) -> np.ndarray:
        """3x3 blocks of (J^T J)^-1 of every marker's translation at the current poses: [n_markers, 3, 3]."""
        out = np.zeros((self.n_markers, 3, 3))
        self._check(self._lib.vmm_ba_marker_translation_covariance_ex(self._h, int(loss_kind), float(loss_scale), int(method), _dp(out)))
        return out

    def schur_multiply(self, x, loss_kind=LOSS_NONE, loss_scale=1.0, radius=1e4, path=0) -> np.ndarray:
        """y = S x with the reduced camera system at the current poses (test entry, one GPU)."""
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1)
        assert x.size == 6 * self.n_cameras
        y = np.zeros_like(x)
        self._check(self._lib.vmm_ba_schur_multiply(self._h, loss_kind, loss_scale, radius, path, _dp(x), _dp(y)))
        return y

    def debug_coarse_matrix(self, loss_kind=LOSS_NONE, loss_scale=1.0, radius=1e4):
        """(A_c, A_c^-1 as computed on the device, order) of the two-level preconditioner; (None, None, 0) without a coarse space."""
        n, ld = C.c_int(0), C.c_int(0)
        self._check(self._lib.vmm_ba_debug_coarse_matrix(self._h, loss_kind, loss_scale, radius, 0, None, None, C.byref(n), C.byref(ld)))
        if n.value == 0:
            self.coarse_unavailable_reason = self._lib.vmm_ba_last_error(self._h).decode()
            return None, None, 0
        a, inv = np.zeros((ld.value, ld.value)), np.zeros((ld.value, ld.value))
        self._check(self._lib.vmm_ba_debug_coarse_matrix(self._h, loss_kind, loss_scale, radius, a.size, _dp(a), _dp(inv), C.byref(n), C.byref(ld)))
        return a, inv, n.value

    def set_coarse_cluster_size(self, cameras_per_cluster: int):
        """Target cameras per cluster of the two-level preconditioner (before set_scene)."""
        self._check(self._lib.vmm_ba_set_coarse_cluster_size(self._h, int(cameras_per_cluster)))

    def shard_info(self):
        a, b, n = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        self._check(self._lib.vmm_ba_shard_info(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value

    def h2d_bytes(self) -> int:
        return int(self._lib.vmm_ba_h2d_bytes(self._h))

    def kernel_launches(self) -> int:
        return int(self._lib.vmm_ba_kernel_launches(self._h))

    def profile_enable(self, on: bool = True):
        """Bracket every kernel launch with CUDA events on the launching stream (clears the counters)."""
        self._check(self._lib.vmm_ba_profile_enable(self._h, 1 if on else 0))

    def profile(self) -> dict:
        """{kernel name: (launches, total device ms)} since profile_enable()."""
        out = {}
        for i in range(self._lib.vmm_ba_profile_num_kernels()):
            n, ms = C.c_int64(0), C.c_double(0)
            self._check(self._lib.vmm_ba_profile_get(self._h, i, C.byref(n), C.byref(ms)))
            if n.value:
                out[self._lib.vmm_ba_profile_kernel_name(i).decode()] = (n.value, ms.value)
        return out

    def pcg_phase_ns(self) -> dict:
        """In-kernel phase clock of the persistent PCG since profile_enable(True) (CTA 0)."""
        v = np.zeros(10)
        self._check(self._lib.vmm_ba_pcg_phase_ns(self._h, _dp(v)))
        return dict(zip(("product", "update_restrict", "coarse_matvec", "direction", "grid_barriers", "product_cta_min", "product_cta_mean",
                         "product_cta_max", "boundary_exchange", "boundary_scatter"), (float(x) for x in v)))

    def pcg_cta_product_ns(self, n: int = 148) -> np.ndarray:
        v = np.zeros(n)
        self._check(self._lib.vmm_ba_pcg_cta_product_ns(self._h, _dp(v), n))
        return v

    def measure_fp64_peak(self) -> float:
        v = C.c_double(0)
        self._check(self._lib.vmm_ba_measure_fp64_peak(self._h, C.byref(v)))
        return v.value

    def measure_copy_bandwidth(self, nbytes: int = 1 << 30) -> float:
        v = C.c_double(0)
        self._check(self._lib.vmm_ba_measure_copy_bandwidth(self._h, nbytes, C.byref(v)))
        return v.value
