"""The C-ABI boundary without a GPU: libvmm_ba.so loads, exports every symbol include/vmm_ba.h declares,
its structs match the ctypes mirror, host-only entry points work, and every computing entry point
fails loudly (VMM_BA_ERR_CUDA, no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    with open(os.path.join(ROOT, "include", "vmm_ba.h")) as f:
        text = f.read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vmm_ba_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 24
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/vmm_ba.h but not exported"
    assert sorted(engine.ABI_SYMBOLS) == syms, "engine.ABI_SYMBOLS out of sync with include/vmm_ba.h"


def test_abi_version_and_defaults(lib):
    assert lib.vmm_ba_abi_version() == 2
    o = pkg.default_options()
    # Ceres defaults in force at src/TagReconstructor.cpp:725-735
    assert o.function_tolerance == 1e-6 and o.gradient_tolerance == 1e-10 and o.parameter_tolerance == 1e-8
    assert o.initial_trust_region_radius == 1e4 and o.max_trust_region_radius == 1e16
    assert o.min_relative_decrease == 1e-3 and o.min_lm_diagonal == 1e-6 and o.max_lm_diagonal == 1e32
    assert o.max_num_consecutive_invalid_steps == 5 and o.jacobi_scaling == 1
    assert o.loss_kind == pkg.LOSS_NONE and o.loss_scale == 1.0 and o.parity_mode == 0 and o.pcg_confirm_tolerance == 1e-3 and o.pcg_tolerance == 1e-2


def test_struct_layout_matches_header():
    """Field order of the ctypes mirrors == field order in the header."""
    with open(os.path.join(ROOT, "include", "vmm_ba.h")) as f:
        text = f.read()

    def fields(struct_name):
        body = re.search(r"typedef struct \{([^{}]*)\} " + struct_name + ";", text).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        return [re.sub(r"\[.*\]", "", m) for m in re.findall(r"\b(\w+(?:\[\d+\])?)\s*;", body)]

    assert fields("vmm_ba_options") == [n for n, _ in engine.Options._fields_]
    assert fields("vmm_ba_summary") == [n for n, _ in engine.Summary._fields_]


def test_partition_cameras_is_balanced_and_contiguous():
    rng = np.random.default_rng(0)
    to_cam = np.sort(rng.integers(0, 1000, 40000)).astype(np.int32)
    for world in (1, 2, 3, 8):
        b = pkg.partition_cameras(1000, to_cam, world)
        assert b[0] == 0 and b[-1] == 1000 and np.all(np.diff(b) >= 0)
        counts = [np.count_nonzero((to_cam >= b[r]) & (to_cam < b[r + 1])) for r in range(world)]
        assert sum(counts) == 40000
        assert max(counts) - min(counts) <= 2 * 40000 / 1000 * 3 + 100
    # no observations at all: split cameras evenly
    b = pkg.partition_cameras(10, np.zeros(0, dtype=np.int32), 4)
    assert list(b) == [0, 2, 5, 7, 10]
    with pytest.raises(pkg.VmmBaError):
        pkg.partition_cameras(10, np.array([11], dtype=np.int32), 2)


def test_no_device_fails_loudly(lib, have_gpu):
    if have_gpu:
        pytest.skip("a CUDA device is present")
    assert pkg.device_count() == 0
    with pytest.raises(pkg.VmmBaError, match="no CPU fallback"):
        pkg.BundleAdjuster(0)
    h = C.c_void_p()
    assert lib.vmm_ba_create(0, C.byref(h)) == 2      # VMM_BA_ERR_CUDA
    assert not h.value
    assert b"no CUDA device" in lib.vmm_ba_last_error(None)


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under visual_marker_mapping_b200/ may reference it."""
    pkg_dir = os.path.join(ROOT, "visual_marker_mapping_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or fn == "Makefile":
                with open(os.path.join(dirpath, fn), errors="ignore") as f:
                    t = f.read()
                assert "vmm_oracle" not in t and "from oracle" not in t and "import oracle" not in t, fn


def test_no_alloc_entry_checks_its_limits_before_touching_the_device(lib):
    """vmm_ba_cost_no_alloc (static device scratch, no allocation): sizes beyond the scratch, bad losses and null pointers are
    refused with a clear text before any CUDA call — so this runs without a GPU."""
    from visual_marker_mapping_b200 import engine
    intr = np.arange(1.0, 10.0)
    ok = dict(num_cameras=4, num_markers=2, num_tag_obs=3)
    for bad, text in ((dict(num_cameras=engine.NO_ALLOC_MAX_CAMERAS + 1), "exceed the static scratch"),
                      (dict(num_markers=engine.NO_ALLOC_MAX_MARKERS + 1), "exceed the static scratch"),
                      (dict(num_tag_obs=engine.NO_ALLOC_MAX_TAG_OBS + 1), "exceed the static scratch"),
                      (dict(num_cameras=-1), "negative size")):
        n = dict(ok, **bad)
        with pytest.raises(pkg.VmmBaError, match=text):
            engine.cost_no_alloc(0, 0, intr, n["num_cameras"], 8, 8, n["num_markers"], 8, 8, 8, n["num_tag_obs"], 8, 8, 8, pkg.LOSS_HUBER, 1.0, 8)
    with pytest.raises(pkg.VmmBaError, match="bad loss"):
        engine.cost_no_alloc(0, 0, intr, 4, 8, 8, 2, 8, 8, 8, 3, 8, 8, 8, 7, 1.0, 8)
    with pytest.raises(pkg.VmmBaError, match="null"):
        engine.cost_no_alloc(0, 0, intr, 4, 0, 8, 2, 8, 8, 8, 3, 8, 8, 8, pkg.LOSS_HUBER, 1.0, 8)
    with pytest.raises(pkg.VmmBaError, match="null"):
        engine.cost_no_alloc(0, 0, intr, 4, 8, 8, 2, 8, 8, 8, 3, 8, 8, 8, pkg.LOSS_HUBER, 1.0, 0)


def test_header_is_plain_c_and_links_from_c99(lib, tmp_path):
    """The boundary is a C ABI: include/vmm_ba.h must compile with a C compiler (-std=c99 -pedantic, no warnings) and a C program
    must link against libvmm_ba.so and call it."""
    import shutil, subprocess
    cc = shutil.which("gcc")
    if not cc:
        pytest.skip("no gcc")
    src = tmp_path / "c_abi.c"
    src.write_text('#include "vmm_ba.h"\n#include <stdio.h>\n'
                   'int main(void) { vmm_ba_options o; vmm_ba_options_init(&o);\n'
                   '  printf("%d %g %d\\n", vmm_ba_abi_version(), o.pcg_tolerance, VMM_BA_NO_ALLOC_MAX_CAMERAS);\n'
                   '  return vmm_ba_cost_no_alloc(0, 0, 1, 1, 0, 0, 0, 1, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 1, 1.0, 0) == VMM_BA_ERR_INVALID_ARGUMENT ? 0 : 1; }\n')
    libdir = os.path.dirname(pkg.engine.library_path())
    r = subprocess.run([cc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", str(tmp_path / "c_abi"), str(src),
                        "-L", libdir, "-lvmm_ba", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(tmp_path / "c_abi")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    version, tol, cams = r.stdout.split()
    assert int(version) == 2 and float(tol) == 1e-2 and int(cams) == 4096
