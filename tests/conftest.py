"""Shared fixtures.  `-m "not gpu"` runs everywhere (oracle vs golden vectors, host maths, ABI symbol
checks, gloo sharding logic); `-m gpu` tests are the parity tests proper and call through the C ABI."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (sm_100a); run with -m gpu on the GPU box")


def _have_gpu() -> bool:
    try:
        import visual_marker_mapping_b200 as pkg

        return pkg.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    # a `-m gpu` run on a machine without a device must fail loudly, not skip: the tests themselves
    # raise VmmBaError from vmm_ba_create (there is no CPU fallback to fall through to).
    return


@pytest.fixture(scope="session")
def lib():
    """libvmm_ba.so, built on demand (nvcc cross-compiles without a GPU)."""
    from visual_marker_mapping_b200 import _build, load_library

    _build.build_library()
    return load_library()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o

    o.build()
    return o


@pytest.fixture(scope="session")
def math_harness():
    """csrc/ba_math.h compiled for the host (test-only shared object)."""
    import ctypes as C

    d = os.path.join(ROOT, "tests", "host_harness")
    subprocess.run(["make", "-C", d, "-s"], check=True)
    h = C.CDLL(os.path.join(d, "libmath_harness.so"))
    h.hm_corner_full.restype = C.c_double
    h.hm_corner_cost.restype = C.c_double
    return h


@pytest.fixture(scope="session")
def scene_c1():
    from visual_marker_mapping_b200 import synth

    return synth.make_config("c1")


@pytest.fixture(scope="session")
def scene_c2():
    from visual_marker_mapping_b200 import synth

    return synth.make_config("c2")


@pytest.fixture(scope="session")
def have_gpu():
    return _have_gpu()
