"""In-tree build of libvmm_ba.so (nvcc, sm_100a only) and of the C++ host facade.

The built objects stay next to their sources (git-ignored) so that they travel to the GPU box with
the repository snapshot.
"""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
REPO = os.path.dirname(_HERE)
LIB_PATH = os.path.join(CSRC, "libvmm_ba.so")
HOST_CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

NVCC_FLAGS = [
    "-ccbin", HOST_CXX,
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]

CUDA_SOURCES = ["ba_engine.cu"]
CUDA_HEADERS = ["ba_kernels.cuh", "ba_schur.cuh", "ba_coarse.cuh", "ba_pcg.cuh", "ba_pcg_dist.cuh", "ba_coarse_inv.cuh", "ba_stage.cuh", "ba_cov.cuh", "ba_peer.cuh", "ba_cov_block.cuh", "ba_no_alloc.cuh", "ba_math.h", "nccl_dyn.h", os.path.join(REPO, "include", "vmm_ba.h")]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libvmm_ba.so cannot be built (there is no CPU fallback)")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES]
    deps = srcs + [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in CUDA_HEADERS]
    if force or _stale(LIB_PATH, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-o", LIB_PATH] + srcs + ["-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True, cwd=CSRC)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
        if verbose:
            print(res.stderr)
    return LIB_PATH


HOST_DIR = os.path.join(CSRC, "host")


def build_host(force: bool = False) -> dict:
    """Build the C++ facade (TagReconstructor over the C ABI), its CLI and its test driver."""
    out = {}
    mk = os.path.join(HOST_DIR, "Makefile")
    if os.path.exists(mk):
        args = ["make", "-C", HOST_DIR, "-s"] + (["-B"] if force else [])
        res = subprocess.run(args, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("host build failed:\n" + res.stdout + res.stderr)
        out["dir"] = HOST_DIR
    return out
