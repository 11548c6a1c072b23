"""The C++ host side of the drop-in (visual_marker_mapping_b200/csrc/host/): TagReconstructor facade, the
readers/writers of the three project files in the reference's boost::property_tree JSON dialect, and the
visual_marker_mapping-compatible CLI.  CPU tests cover formats and error behaviour; the GPU test runs the
reference's closing sequence BA(Huber) -> cull 2 px -> BA (src/TagReconstructor.cpp:271-277) through the
CLI and compares reconstruction.json with the same sequence done by the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

from visual_marker_mapping_b200 import _build, project_io, synth

HOST = _build.HOST_DIR
CLI = os.path.join(HOST, "visual_marker_mapping")
DRIVER = os.path.join(HOST, "vmm_host_test")


@pytest.fixture(scope="module")
def host(lib):
    _build.build_host()
    assert os.path.exists(CLI) and os.path.exists(DRIVER)
    return HOST


def culled_scene():
    """C1-like scene in which marker 5 is badly detected in most images: the cull step must drop it."""
    s = synth.make_scene(n_markers=12, n_cameras=40, markers_per_camera=12, seed=0x564D4D01)
    rng = np.random.default_rng(4)
    idx = np.nonzero(s.to_tag == 5)[0]
    bad = idx[rng.random(idx.size) < 0.7]
    s.to_uv[bad] += rng.uniform(8, 20, (bad.size, 1)) * np.tile(rng.choice([-1.0, 1.0], (bad.size, 2)), (1, 4))
    return s


def test_roundtrip_is_byte_exact(host, tmp_path, scene_c1):
    src, dst = tmp_path / "src", tmp_path / "dst"
    dst.mkdir()
    project_io.write_project(scene_c1, str(src), image_id_offset=3, tag_id_offset=100)
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert f"images 40 tags 12 observations {scene_c1.n_tag_obs} rec_tags 12 rec_cams 40" in r.stdout
    assert (src / "marker_detections.json").read_bytes() == (dst / "marker_detections.json").read_bytes()
    a = (src / "initial_reconstruction.json").read_text().splitlines()
    b = (dst / "initial_reconstruction.json").read_text().splitlines()
    assert len(a) == len(b)
    for x, y in zip(a, b):
        if x != y:   # only the derived world corners may differ, in the last digit
            fx, fy = float(x.strip().strip('",')), float(y.strip().strip('",'))
            assert abs(fx - fy) <= 4e-16 * max(abs(fx), 1.0)


def test_number_text_roundtrip_extremes(host, tmp_path, scene_c1):
    """Doubles go out as boost's setprecision(17) prints them (= C "%.17g", here through std::to_chars) and come back through
    std::from_chars / strtod: denormals, the largest double, negative zero, integers beyond 2^53, exponent forms."""
    src, dst = tmp_path / "src", tmp_path / "dst"
    dst.mkdir()
    project_io.write_project(scene_c1, str(src))
    det = json.loads((src / "marker_detections.json").read_text())
    vals = [5e-324, 2.2250738585072014e-308, 1.7976931348623157e308, -0.0, 0.1, 1.0 / 3.0, 123456789012345678.0, 1e22, 1e-7,
            -4536.70849609375, 6.02214076e23, 2.0 ** 53 + 2, 1e15, 1e16, 1e17, 0.30000000000000004]
    rng = np.random.default_rng(3)
    vals += list(np.ldexp(rng.standard_normal(200), rng.integers(-1000, 1000, 200)))
    k = 0
    for o in det["tag_observations"]:
        for c in o["observations"]:
            for i in range(2):
                c[i] = "%.17g" % vals[k % len(vals)]
                k += 1
    assert k >= len(vals)
    text = project_io.dumps_ptree(det)
    (src / "marker_detections.json").write_text(text)
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (dst / "marker_detections.json").read_text() == text
    # the same values as bare JSON numbers in Python's shortest form, '+' exponents and capital E included
    for o in det["tag_observations"]:
        for c in o["observations"]:
            c[0], c[1] = float(c[0]), float(c[1])
    (src / "marker_detections.json").write_text(json.dumps(det).replace("e+", "E+"))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (dst / "marker_detections.json").read_text() == text


def test_dialect_literals(tmp_path, scene_c1):
    """boost write_json: quoted leaves, 17 significant digits, 4-space indent, '/' escaped, empty array = \"\"."""
    project_io.write_project(scene_c1, str(tmp_path))
    text = (tmp_path / "marker_detections.json").read_text()
    assert text.startswith('{\n    "images": [\n        {\n            "filename": "images\\/img_000000.jpg",\n            "id": "0"\n        },')
    assert '"tag_type": "apriltag_36h11"' in text and '"width": "0.1285"' in text and '"height": "0.1295"' in text
    assert text.endswith("}\n")
    intr = (tmp_path / "camera_intrinsics.json").read_text()
    assert '"fx": "8075.2937867635346"' in intr and '"horizontal_resolution": "6000"' in intr
    assert project_io.dumps_ptree({"tag_observations": [], "tags": {}}) == '{\n    "tag_observations": "",\n    "tags": ""\n}\n'
    json.loads(text)   # it is still JSON


def test_reader_accepts_bare_numbers_and_empty_arrays(host, tmp_path, scene_c1):
    src, dst = tmp_path / "src", tmp_path / "dst"
    dst.mkdir()
    project_io.write_project(scene_c1, str(src))
    det = json.loads((src / "marker_detections.json").read_text())
    det["tag_observations"] = ""
    for t in det["tags"]:
        t["id"], t["width"] = int(t["id"]), float(t["width"])
    (src / "marker_detections.json").write_text(json.dumps(det))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "tags 12 observations 0" in r.stdout
    assert '"tag_observations": ""' in (dst / "marker_detections.json").read_text()


def test_reader_error_messages(host, tmp_path, scene_c1):
    src, dst = tmp_path / "src", tmp_path / "dst"
    dst.mkdir()
    project_io.write_project(scene_c1, str(src))
    good = (src / "marker_detections.json").read_text()
    det = json.loads(good)
    det["tag_observations"][0]["observations"].pop()           # 3 corners (src/DetectionIO.cpp:50-51)
    (src / "marker_detections.json").write_text(json.dumps(det))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "Unexpected number of values" in r.stderr
    (src / "marker_detections.json").write_text(good)
    intr = json.loads((src / "camera_intrinsics.json").read_text())
    intr["distortion_coefficients"].pop()                       # PropertyTreeUtilities.h:118-121
    (src / "camera_intrinsics.json").write_text(json.dumps(intr))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "Not enough many parameters in vector. Expected 5, got 4!" in r.stderr


def test_streamed_io_is_bounded_in_memory(host, tmp_path):
    """site-sized files go through the chunked section reader and the streaming writer.  400 k tag
    observations (a 237 MB marker_detections.json) are written, read back and compared; the process may grow by
    the DetectionResult itself (~50 MB) but never by a whole-file tree (several times the file size)."""
    r = subprocess.run([DRIVER, "bigio", str(tmp_path), "400000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    f = dict(zip(r.stdout.split()[0::2], map(int, r.stdout.split()[1::2])))
    assert f["observations"] == 400000 and f["mismatches"] == 0
    assert f["file_bytes"] > 200e6
    assert (f["maxrss_kb"] - f["base_rss_kb"]) * 1024 < 0.35 * f["file_bytes"]
    text = (tmp_path / "marker_detections.json").read_text()
    assert text.startswith('{\n    "images": [\n        {\n            "filename": "images\\/img_0.jpg",\n            "id": "0"\n        },')
    assert text.endswith('                ]\n            ]\n        }\n    ]\n}\n')


def test_streamed_reader_edge_cases(host, tmp_path, scene_c1):
    """Section reader: key order free, unknown sections and members skipped, object-valued sections iterate their
    members (ptree makes no array/object distinction), missing sections and syntax errors reported."""
    src, dst = tmp_path / "src", tmp_path / "dst"
    dst.mkdir()
    project_io.write_project(scene_c1, str(src))
    good = (src / "marker_detections.json").read_text()
    det = json.loads(good)
    shuffled = {"comment": {"deep": [[1, {"x": []}], "y"]}, "tag_observations": det["tag_observations"],
                "tags": {f"t{i}": dict(t, extra=[1, 2]) for i, t in enumerate(det["tags"])}, "version": 3, "images": det["images"]}
    (src / "marker_detections.json").write_text(json.dumps(shuffled, indent=1))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert (dst / "marker_detections.json").read_text() == good
    del det["tags"]
    (src / "marker_detections.json").write_text(json.dumps(det))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "No such node (tags)" in r.stderr
    (src / "marker_detections.json").write_text(good[: len(good) // 2])
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "marker_detections.json(" in r.stderr
    (src / "marker_detections.json").write_text(good.rstrip() + " x")
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "garbage after data" in r.stderr
    (src / "marker_detections.json").write_text(good)
    rec = json.loads((src / "initial_reconstruction.json").read_text())
    del rec["reconstructed_cameras"]
    (src / "initial_reconstruction.json").write_text(json.dumps(rec))
    r = subprocess.run([DRIVER, "roundtrip", str(src), str(dst)], capture_output=True, text=True)
    assert r.returncode == 1 and "No such node (reconstructed_cameras)" in r.stderr


def _ptree_leaves(node):
    """What boost::property_tree keeps of a JSON document: every leaf as text."""
    if isinstance(node, dict):
        return {k: _ptree_leaves(v) for k, v in node.items()}
    if isinstance(node, list):
        return [_ptree_leaves(v) for v in node]
    if isinstance(node, str):
        return node
    return json.dumps(node)   # true / false / null / numbers: the literal's text


def test_tree_and_streamed_json_paths_agree(host, tmp_path):
    """Property test of the two JSON paths of the host library on random documents: read_json + write_json (tree) and
    read_json_sections + Writer (streamed) must print the same bytes, and those are the boost write_json dialect as
    project_io.dumps_ptree states it (quoted leaves, "" for empty containers, escaped '/')."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    text = st.text(alphabet=st.characters(min_codepoint=32, max_codepoint=126), max_size=12)
    leaf = st.one_of(text, st.integers(-10**9, 10**9), st.booleans(), st.none(),
                     st.floats(allow_nan=False, allow_infinity=False, width=64))
    keys = st.text(alphabet=st.characters(min_codepoint=33, max_codepoint=126), min_size=1, max_size=8)
    value = st.recursive(leaf, lambda c: st.one_of(st.lists(c, max_size=4), st.dictionaries(keys, c, max_size=4)), max_leaves=20)
    container = st.one_of(st.lists(value, max_size=5), st.dictionaries(keys, value, max_size=5))
    docs = st.dictionaries(keys, container, max_size=5)
    src, tree, stream = tmp_path / "in.json", tmp_path / "tree.json", tmp_path / "stream.json"

    @settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
    @given(docs, st.sampled_from([None, 1, 4]), st.booleans())
    def check(doc, indent, ascii_only):
        src.write_text(json.dumps(doc, indent=indent, ensure_ascii=ascii_only))
        r = subprocess.run([DRIVER, "json", str(src), str(tree), str(stream)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert tree.read_bytes() == stream.read_bytes()
        assert tree.read_text() == project_io.dumps_ptree(_ptree_leaves(doc))

    check()
    # control characters and non-ASCII text: both paths escape / pass them through alike
    src.write_text(json.dumps({"a": ["tab\there", "nl\n", "\u00e9\u4e2d", "q\"b\\/"], "b": {"k\u0001": ["\u001f"]}, "c": [], "d": {}}))
    r = subprocess.run([DRIVER, "json", str(src), str(tree), str(stream)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert tree.read_bytes() == stream.read_bytes()
    out = tree.read_text()
    assert '"tab\\there"' in out and '"\\u001F"' in out and '"q\\"b\\\\\\/"' in out and "\u00e9\u4e2d" in out
    assert out.endswith('    "c": "",\n    "d": ""\n}\n')


@pytest.mark.parametrize("shape", [(40, 12, 475), (2000, 500, 200000)])
def test_flatten_matches_map_restatement(host, shape):
    """TagReconstructor::flatProblem (the arrays doBundleAdjustment hands to vmm_ba_set_problem: id tables + a threaded
    two-pass fill) against the observation-by-observation std::map restatement of src/TagReconstructor.cpp:663-724, with
    sparse ids (table and binary-search look-ups), unreconstructed images / tags, and a second call on the reused buffers."""
    r = subprocess.run([DRIVER, "flatten", *map(str, shape)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    f = r.stdout.split()
    assert f[0] == "kept" and int(f[1]) > 0.7 * shape[2] and int(f[3]) == shape[2] and f[4:6] == ["mismatches", "0"]


@pytest.fixture(scope="module")
def abi_double(host):
    """Environment that puts tests/host_harness/libabi_test_double.so (identity solve + plain C++ reprojection sums; test
    infrastructure, see its header) in front of libvmm_ba.so: the facade's host logic then runs without a GPU."""
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host_harness")
    subprocess.run(["make", "-C", d, "-s"], check=True)
    return dict(os.environ, LD_PRELOAD=os.path.join(d, "libabi_test_double.so"))


def test_facade_statistics_host_logic(host, oracle, abi_double, tmp_path, scene_c1):
    """flatten + per-image / per-tag aggregation of computeReprojectionErrorPerImg/PerTag (src/TagReconstructor.cpp:340-428)
    with sparse ids, against numpy on the oracle's per-observation sums.  (The same check runs on the GPU in
    test_facade_reprojection_statistics; here the sums come from the ABI test double.)"""
    project_io.write_project(scene_c1, str(tmp_path), image_id_offset=11, tag_id_offset=300)
    r = subprocess.run([DRIVER, "stats", str(tmp_path)], capture_output=True, text=True, env=abi_double)
    assert r.returncode == 0, r.stderr
    got = json.loads(r.stdout)
    s = scene_c1
    sums = oracle.reprojection_error_sums(s)
    per_tag = np.bincount(s.to_tag, sums, s.n_markers) / (4 * np.bincount(s.to_tag, minlength=s.n_markers))
    per_img = np.bincount(s.to_cam, sums, s.n_cameras) / (4 * np.bincount(s.to_cam, minlength=s.n_cameras))
    assert sorted(map(int, got["per_tag"])) == [300 + j for j in range(s.n_markers)]
    assert sorted(map(int, got["per_img"])) == [11 + i for i in range(s.n_cameras)]
    for j in range(s.n_markers):
        assert abs(float(got["per_tag"][str(300 + j)]) - per_tag[j]) < 1e-9 * max(per_tag[j], 1.0)
    for i in range(s.n_cameras):
        assert abs(float(got["per_img"][str(11 + i)]) - per_img[i]) < 1e-9 * max(per_img[i], 1.0)
    assert abs(float(got["avg"]) - sums.sum() / (4 * s.n_tag_obs)) < 1e-9 * float(got["avg"])


def test_cli_closing_sequence_host_logic(host, oracle, abi_double, tmp_path):
    """The CLI's closing sequence BA -> removeBadMarkers(2) -> removeBadCameras(2) -> BA -> export
    (src/TagReconstructor.cpp:271-277, :786-816) with the identity solve of the ABI test double, started at the ground
    truth: which markers and cameras survive is host logic and must equal the numpy statement of the same rules —
    the badly detected marker goes, the origin tag stays whatever its error, a camera left without observations goes."""
    s = culled_scene()
    s.cam_q, s.cam_t, s.tag_q, s.tag_t = s.cam_q_gt.copy(), s.cam_t_gt.copy(), s.tag_q_gt.copy(), s.tag_t_gt.copy()
    lonely = 17                                   # camera 17 only sees marker 5: culled once the marker is gone (-1, :379-383, :809)
    keep = (s.to_cam != lonely) | (s.to_tag == 5)
    s.to_cam, s.to_tag, s.to_uv = s.to_cam[keep], s.to_tag[keep], s.to_uv[keep]
    s.to_uv[s.to_tag == s.fixed_tag] += 6.0       # the origin tag is never removed (:794)
    project_io.write_project(s, str(tmp_path), image_id_offset=7, tag_id_offset=20)
    r = subprocess.run([CLI, "--project_path", str(tmp_path), "--start_tag_id", str(20 + s.fixed_tag), "--overwrite",
                        "--initial_reconstruction", str(tmp_path / "initial_reconstruction.json")], capture_output=True,
                       text=True, env=abi_double)
    assert r.returncode == 0 and "Error" not in r.stderr, r.stderr
    sums = oracle.reprojection_error_sums(s)
    tag_err = np.bincount(s.to_tag, sums, s.n_markers) / np.maximum(4 * np.bincount(s.to_tag, minlength=s.n_markers), 1)
    keep_tag = ~((tag_err > 2.0) & (np.arange(s.n_markers) != s.fixed_tag))
    ko = keep_tag[s.to_tag]
    cam_n = np.bincount(s.to_cam[ko], minlength=s.n_cameras)
    cam_err = np.where(cam_n > 0, np.bincount(s.to_cam[ko], sums[ko], s.n_cameras) / np.maximum(4 * cam_n, 1), -1.0)
    keep_cam = ~((cam_err > 2.0) | (cam_err < 0))
    assert not keep_tag[5] and keep_tag[s.fixed_tag] and tag_err[s.fixed_tag] > 2.0 and not keep_cam[lonely]
    tags, cams = project_io.read_reconstruction(str(tmp_path / "reconstruction.json"))
    assert sorted(tags) == [20 + int(j) for j in np.nonzero(keep_tag)[0]]
    assert sorted(cams) == [7 + int(i) for i in np.nonzero(keep_cam)[0]]
    assert "Removing bad marker with id 25" in r.stdout and f"Removing bad camera with id {7 + lonely}" in r.stdout
    for j in np.nonzero(keep_tag)[0]:            # identity solve: the poses come back as they went in
        assert np.array_equal(tags[20 + int(j)]["t"], s.tag_t[j]) and np.array_equal(tags[20 + int(j)]["q"], s.tag_q[j])
    for i in np.nonzero(keep_cam)[0]:
        assert np.array_equal(cams[7 + int(i)]["t"], s.cam_t[i])


@pytest.mark.parametrize("wave", [False, True])
def test_cli_incremental_schedule_host_logic(host, abi_double, tmp_path, scene_c1, wave):
    """startReconstruction from the detections alone (src/TagReconstructor.cpp:146-267) with the identity solve of the ABI
    test double: every image and every tag gets initialised by the in-repo PnP in the reference's order, one bundle
    adjustment is requested per added image (per round with --wave) plus the two closing ones, and the un-refined PnP
    poses are already close to ground truth (0.3 px corner noise).  The refined result is checked on the GPU
    (test_cli_reconstructs_from_detections_alone)."""
    s = scene_c1
    project_io.write_project(s, str(tmp_path), image_id_offset=3, tag_id_offset=0)
    os.remove(tmp_path / "initial_reconstruction.json")
    r = subprocess.run([CLI, "--project_path", str(tmp_path), "--overwrite"] + (["--wave"] if wave else []), capture_output=True, text=True,
                       env=abi_double)
    assert r.returncode == 0 and "Error" not in r.stderr, r.stderr
    n_ba = r.stdout.count("Solution ")
    assert r.stdout.count("Initialized camera with id") == s.n_cameras
    assert r.stdout.count("Initialized tag with id") == s.n_markers - 1      # the origin tag is set, not initialised
    assert (n_ba < 10) if wave else (n_ba == s.n_cameras + 2)
    tags, cams = project_io.read_reconstruction(str(tmp_path / "reconstruction.json"))
    if not wave:   # (per round, poses that were never refined may miss the 2 px of the closing cull: survivors only)
        assert sorted(tags) == list(range(s.n_markers)) and sorted(cams) == [i + 3 for i in range(s.n_cameras)]
    assert set(tags) <= set(range(s.n_markers)) and set(cams) <= {i + 3 for i in range(s.n_cameras)} and len(cams) > 0
    assert np.array_equal(tags[s.fixed_tag]["t"], np.zeros(3))
    for j in tags:
        assert np.abs(tags[j]["t"] - s.tag_t_gt[j]).max() < 0.05
    for i in cams:
        assert np.abs(cams[i]["t"] - s.cam_t_gt[i - 3]).max() < 0.05


def test_cli_without_device_reports_and_returns_zero(host, tmp_path, scene_c1, have_gpu):
    """No CPU fallback; like the reference's main the exception is printed and the exit code stays 0."""
    if have_gpu:
        pytest.skip("a CUDA device is present")
    project_io.write_project(scene_c1, str(tmp_path))
    r = subprocess.run([CLI, "--project_path", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0
    assert "no CUDA device" in r.stderr and "no CPU fallback" in r.stderr
    assert not (tmp_path / "reconstruction.json").exists()
    r = subprocess.run([CLI], capture_output=True, text=True)
    assert "--project_path" in r.stderr


def oracle_closing_sequence(oracle, s):
    """BA(Huber, 1500) -> removeBadMarkers(2) -> removeBadCameras(2) -> BA(no loss, 1500) with the oracle."""
    o1 = oracle.solve(s, max_num_iterations=1500, loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
    s1 = s.copy()
    s1.cam_q, s1.cam_t, s1.tag_q, s1.tag_t = o1.cam_q, o1.cam_t, o1.tag_q, o1.tag_t
    sums = oracle.reprojection_error_sums(s1)
    tag_err = np.bincount(s1.to_tag, sums, s.n_markers) / np.maximum(4 * np.bincount(s1.to_tag, minlength=s.n_markers), 1)
    keep_tag = ~((tag_err > 2.0) & (np.arange(s.n_markers) != s.fixed_tag))
    ko = keep_tag[s1.to_tag]
    cam_n = np.bincount(s1.to_cam[ko], minlength=s.n_cameras)
    cam_err = np.where(cam_n > 0, np.bincount(s1.to_cam[ko], sums[ko], s.n_cameras) / np.maximum(4 * cam_n, 1), -1.0)
    keep_cam = ~((cam_err > 2.0) | (cam_err < 0))
    ko &= keep_cam[s1.to_cam]
    tmap = np.cumsum(keep_tag) - 1
    cmap = np.cumsum(keep_cam) - 1
    s2 = synth.Scene(intr=s.intr, cam_q=o1.cam_q[keep_cam], cam_t=o1.cam_t[keep_cam], tag_q=o1.tag_q[keep_tag], tag_t=o1.tag_t[keep_tag],
                     tag_wh=s.tag_wh[keep_tag], fixed_tag=int(tmap[s.fixed_tag]), to_cam=cmap[s1.to_cam[ko]].astype(np.int32),
                     to_tag=tmap[s1.to_tag[ko]].astype(np.int32), to_uv=s1.to_uv[ko])
    o2 = oracle.solve(s2, max_num_iterations=1500)
    s2.cam_q, s2.cam_t, s2.tag_q, s2.tag_t = o2.cam_q, o2.cam_t, o2.tag_q, o2.tag_t
    o2.final_scene = s2
    return np.nonzero(keep_tag)[0], np.nonzero(keep_cam)[0], o2


@pytest.mark.gpu
def test_cli_closing_sequence_matches_oracle(host, oracle, tmp_path):
    s = culled_scene()
    project_io.write_project(s, str(tmp_path), image_id_offset=7, tag_id_offset=20)
    r = subprocess.run([CLI, "--project_path", str(tmp_path), "--start_tag_id", "20", "--parity", "--overwrite",
                        "--initial_reconstruction", str(tmp_path / "initial_reconstruction.json")], capture_output=True, text=True)
    assert r.returncode == 0 and "Error" not in r.stderr, r.stderr
    assert "Solution 0" in r.stdout and "Removing bad marker with id 25" in r.stdout
    tags, cams = project_io.read_reconstruction(str(tmp_path / "reconstruction.json"))
    keep_tag, keep_cam, o2 = oracle_closing_sequence(oracle, s)
    assert 5 not in keep_tag
    assert sorted(tags) == [int(t) + 20 for t in keep_tag] and sorted(cams) == [int(c) + 7 for c in keep_cam]
    for j, t in enumerate(keep_tag):
        assert np.abs(tags[int(t) + 20]["t"] - o2.tag_t[j]).max() < 1e-6
        assert synth.quat_angle(tags[int(t) + 20]["q"], o2.tag_q[j]) < 1e-6
    for i, c in enumerate(keep_cam):
        assert np.abs(cams[int(c) + 7]["t"] - o2.cam_t[i]).max() < 1e-6
        assert synth.quat_angle(cams[int(c) + 7]["q"], o2.cam_q[i]) < 1e-6
    # the covariance print-out of the final solve (src/TagReconstructor.cpp:744-783) against the oracle's dense (J^T J)^-1
    cov = oracle.marker_translation_covariance(o2.final_scene)
    std = np.sqrt(np.diagonal(cov, axis1=1, axis2=2))
    lines = {int(l.split()[3].rstrip(":")): l for l in r.stdout.splitlines() if l.startswith("StdDev of tag ")}
    assert sorted(lines) == [int(t) + 20 for t in keep_tag]
    for j, t in enumerate(keep_tag):
        f = lines[int(t) + 20].split()
        got = np.array([float(f[4]), float(f[5]), float(f[6])])
        assert np.abs(got - std[j]).max() <= 1e-4 * std.max() + 1e-12, (t, got, std[j])     # 6 printed digits
        assert abs(float(f[-1]) - np.sqrt(np.linalg.norm(std[j]))) <= 1e-4 * np.sqrt(np.linalg.norm(std[j])) + 1e-12
    rms = [float(l.split("=")[1]) for l in r.stdout.splitlines() if l.startswith("Marker Position RMS")]
    assert len(rms) == 1 and abs(rms[0] - np.linalg.norm(np.sqrt(np.diagonal(cov, axis1=1, axis2=2).mean(0)))) <= 1e-4 * rms[0]
    # reconstruction.json carries the derived world corners and the camera model as the reference writes them
    root = json.loads((tmp_path / "reconstruction.json").read_text())
    assert list(root) == ["reconstructed_tags", "reconstructed_marker_corners", "reconstructed_cameras", "camera_model"]
    assert len(root["reconstructed_marker_corners"]) == 4 * len(keep_tag)


@pytest.mark.gpu
def test_facade_reprojection_statistics(host, oracle, tmp_path, scene_c1):
    project_io.write_project(scene_c1, str(tmp_path))
    r = subprocess.run([DRIVER, "stats", str(tmp_path)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = json.loads(r.stdout)
    sums = oracle.reprojection_error_sums(scene_c1)
    n_tag = np.bincount(scene_c1.to_tag, minlength=scene_c1.n_markers)
    per_tag = np.bincount(scene_c1.to_tag, sums, scene_c1.n_markers) / (4 * n_tag)
    for j in range(scene_c1.n_markers):
        assert abs(float(got["per_tag"][str(j)]) - per_tag[j]) < 1e-9 * max(per_tag[j], 1.0)
    assert abs(float(got["avg"]) - sums.sum() / (4 * scene_c1.n_tag_obs)) < 1e-9 * float(got["avg"])
    per_img = np.bincount(scene_c1.to_cam, sums, scene_c1.n_cameras) / (4 * np.bincount(scene_c1.to_cam, minlength=scene_c1.n_cameras))
    for i in range(scene_c1.n_cameras):
        assert abs(float(got["per_img"][str(i)]) - per_img[i]) < 1e-9 * max(per_img[i], 1.0)


# ---- incremental PnP initialisation (src/TagReconstructor.cpp:146-267) -----------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("wave", [False, True])
def test_cli_reconstructs_from_detections_alone(host, oracle, tmp_path, scene_c1, wave):
    """No initial poses: the tool initialises every camera (robust PnP on the corners of the tags reconstructed so far) and
    every tag (PnP on its four corners), bundle-adjusts per added image (or per round with --wave) and runs the closing
    sequence.  The result is the least-squares solution of the final problem — the same one the oracle reaches from the
    synthetic initial guess — and close to ground truth."""
    s = scene_c1
    project_io.write_project(s, str(tmp_path), image_id_offset=3, tag_id_offset=0)
    os.remove(tmp_path / "initial_reconstruction.json")
    args = [CLI, "--project_path", str(tmp_path), "--parity", "--overwrite"] + (["--wave"] if wave else [])
    r = subprocess.run(args, capture_output=True, text=True)
    assert r.returncode == 0 and "Error" not in r.stderr, r.stderr
    n_ba = r.stdout.count("Solution ")
    assert ("Initialized camera with id" in r.stdout) and ("Initialized tag with id" in r.stdout)
    if wave:
        assert n_ba < 10                                  # a handful of rounds + the two closing solves
    else:
        assert n_ba == s.n_cameras + 2                    # one bundle adjustment per image, :243, plus :271,:277
    tags, cams = project_io.read_reconstruction(str(tmp_path / "reconstruction.json"))
    assert sorted(tags) == list(range(s.n_markers)) and sorted(cams) == [i + 3 for i in range(s.n_cameras)]
    o = oracle.solve(s, max_num_iterations=1500)          # final solve of the reference sequence runs without a loss
    for j in range(s.n_markers):
        assert np.abs(tags[j]["t"] - o.tag_t[j]).max() < 2e-5
        assert synth.quat_angle(tags[j]["q"], o.tag_q[j]) < 2e-5
        assert np.abs(tags[j]["t"] - s.tag_t_gt[j]).max() < 5e-3
    for i in range(s.n_cameras):
        assert np.abs(cams[i + 3]["t"] - o.cam_t[i]).max() < 2e-5
        assert synth.quat_angle(cams[i + 3]["q"], o.cam_q[i]) < 2e-5


@pytest.mark.gpu
def test_cli_origin_tag_errors(host, tmp_path, scene_c1):
    project_io.write_project(scene_c1, str(tmp_path))
    os.remove(tmp_path / "initial_reconstruction.json")
    r = subprocess.run([CLI, "--project_path", str(tmp_path), "--start_tag_id", "99", "--overwrite"], capture_output=True, text=True)
    assert r.returncode == 0                              # the reference's main swallows the exception, src/main_mapping.cpp:90-95
    assert "Could not use tag with id 99 as origin tag, because it was not detected." in r.stderr


def test_move_tag_into_origin_matches_reference_formulas(host, tmp_path, scene_c1):
    """src/TagReconstructor.cpp:314-338: q = quat(tag)^-1, every pose becomes (q * pose.q, R(q) (pose.t - tag.t))."""
    s = scene_c1
    project_io.write_project(s, str(tmp_path))
    out = tmp_path / "moved.json"
    r = subprocess.run([DRIVER, "moveorigin", str(tmp_path), "3", str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    tags, cams = project_io.read_reconstruction(str(out))

    def qmul(a, b):
        return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                         a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])

    q3 = s.tag_q[3] / np.dot(s.tag_q[3], s.tag_q[3]) * np.array([1, -1, -1, -1])
    R = synth.quat_to_rot(q3[None] / np.linalg.norm(q3))[0] if abs(np.linalg.norm(q3) - 1) > 1e-9 else synth.quat_to_rot(q3[None])[0]
    assert np.abs(tags[3]["t"]).max() < 1e-12 and synth.quat_angle(tags[3]["q"], np.array([1.0, 0, 0, 0])) < 1e-9
    for j in range(s.n_markers):
        assert np.abs(tags[j]["q"] - qmul(q3, s.tag_q[j])).max() < 1e-9
        assert np.abs(tags[j]["t"] - R @ (s.tag_t[j] - s.tag_t[3])).max() < 1e-9
    for i in range(s.n_cameras):
        assert np.abs(cams[i]["q"] - qmul(q3, s.cam_q[i])).max() < 1e-9
        assert np.abs(cams[i]["t"] - R @ (s.cam_t[i] - s.tag_t[3])).max() < 1e-9
    bad = subprocess.run([DRIVER, "moveorigin", str(tmp_path), "77", str(out)], capture_output=True, text=True)
    assert bad.returncode == 1 and "Tag with id 77 is not reconstructed." in bad.stderr
