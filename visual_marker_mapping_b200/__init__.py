"""visual_marker_mapping_b200 — B200-native bundle adjustment for visual marker mapping.

Drop-in for one path of schmidtniko/visual_marker_mapping: TagReconstructor::doBundleAdjustment
(src/TagReconstructor.cpp:646-784).  The product is csrc/ (CUDA kernels + C ABI, include/vmm_ba.h)
and csrc/host/ (C++ TagReconstructor facade); this package is the Python plumbing around it.
"""
from . import synth  # noqa: F401
from .engine import (  # noqa: F401
    BundleAdjuster, Options, Summary, VmmBaError, default_options, device_count, load_library,
    partition_cameras, shard_plan, LOSS_NONE, LOSS_HUBER, LOSS_CAUCHY, CONVERGENCE, NO_CONVERGENCE, FAILURE, COV_AUTO, COV_DENSE, COV_BLOCK_TRIDIAGONAL,
)

__all__ = [
    "BundleAdjuster", "Options", "Summary", "VmmBaError", "default_options", "device_count",
    "load_library", "partition_cameras", "shard_plan", "synth",
]
