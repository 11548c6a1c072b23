"""vmm_ba_cost_no_alloc (include/vmm_ba.h): the entry point that allocates nothing — scratch in the library's static
device array — against vmm_ba_evaluate on the same problem, through the C ABI, on two calls in a row (the second call
finds the scratch full of the first call's, different, problem).  No counterpart in the reference.
This file is collected last on purpose."""
import numpy as np
import pytest

import visual_marker_mapping_b200 as pkg
from visual_marker_mapping_b200 import engine, synth

pytestmark = pytest.mark.gpu

COST_RTOL = 1e-12   # the tolerance of the evaluate tests (tests/test_gpu_parity.py): the sum order differs, nothing else


def _device_copy(torch, scene):
    dev = torch.device("cuda", 0)
    f = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
    i = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).to(dev)
    return {"cam_q": f(scene.cam_q), "cam_t": f(scene.cam_t), "tag_q": f(scene.tag_q), "tag_t": f(scene.tag_t), "tag_wh": f(scene.tag_wh),
            "to_cam": i(scene.to_cam), "to_tag": i(scene.to_tag), "to_uv": f(scene.to_uv)}


def _cost_no_alloc(torch, scene, d, out, loss, a):
    engine.cost_no_alloc(0, 0, scene.intr, scene.n_cameras, d["cam_q"].data_ptr(), d["cam_t"].data_ptr(), scene.n_markers, d["tag_q"].data_ptr(),
                         d["tag_t"].data_ptr(), d["tag_wh"].data_ptr(), scene.n_tag_obs, d["to_cam"].data_ptr(), d["to_tag"].data_ptr(),
                         d["to_uv"].data_ptr(), loss, a, out.data_ptr())
    torch.cuda.synchronize()
    return float(out.cpu()[0])


def test_cost_no_alloc_matches_evaluate_on_two_calls_in_a_row():
    import torch
    scenes = [synth.make_config("c2"), synth.make_config("c1")]   # 2000 cameras / 500 markers, then 40 / 12: the scratch is reused
    assert scenes[0].n_cameras <= engine.NO_ALLOC_MAX_CAMERAS and scenes[0].n_tag_obs <= engine.NO_ALLOC_MAX_TAG_OBS
    want = []
    for s in scenes:
        with pkg.BundleAdjuster(0) as ba:
            ba.set_scene(s)
            want.append([ba.evaluate(loss, a, residuals=False, jacobians=False, gradient=False).cost
                         for loss, a in ((pkg.LOSS_HUBER, 1.0), (pkg.LOSS_CAUCHY, 3.0), (pkg.LOSS_NONE, 1.0))])
    dev = [_device_copy(torch, s) for s in scenes]
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda:0")   # the caller's output buffer, allocated before the calls
    torch.cuda.synchronize()
    for rep in range(2):   # c2, c1, c2, c1: every call after the first starts from another problem's scratch contents
        for s, d, w in zip(scenes, dev, want):
            for (loss, a), expect in zip(((pkg.LOSS_HUBER, 1.0), (pkg.LOSS_CAUCHY, 3.0), (pkg.LOSS_NONE, 1.0)), w):
                got = _cost_no_alloc(torch, s, d, out, loss, a)
                assert abs(got - expect) <= COST_RTOL * abs(expect), (rep, s.n_cameras, loss, got, expect)


def test_cost_no_alloc_rejects_sizes_beyond_the_static_scratch_and_handles_empty_input():
    import torch
    s = synth.make_config("c1")
    d = _device_copy(torch, s)
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda:0")
    for kw in ({"n_cameras": engine.NO_ALLOC_MAX_CAMERAS + 1}, {"n_markers": engine.NO_ALLOC_MAX_MARKERS + 1}, {"n_tag_obs": engine.NO_ALLOC_MAX_TAG_OBS + 1}):
        n = {"n_cameras": s.n_cameras, "n_markers": s.n_markers, "n_tag_obs": s.n_tag_obs, **kw}
        with pytest.raises(engine.VmmBaError, match="exceed the static scratch"):
            engine.cost_no_alloc(0, 0, s.intr, n["n_cameras"], d["cam_q"].data_ptr(), d["cam_t"].data_ptr(), n["n_markers"], d["tag_q"].data_ptr(),
                                 d["tag_t"].data_ptr(), d["tag_wh"].data_ptr(), n["n_tag_obs"], d["to_cam"].data_ptr(), d["to_tag"].data_ptr(),
                                 d["to_uv"].data_ptr(), pkg.LOSS_HUBER, 1.0, out.data_ptr())
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == -1.0   # nothing was launched
    # no tag observations: cost 0, whatever the scratch held
    engine.cost_no_alloc(0, 0, s.intr, s.n_cameras, d["cam_q"].data_ptr(), d["cam_t"].data_ptr(), s.n_markers, d["tag_q"].data_ptr(),
                         d["tag_t"].data_ptr(), d["tag_wh"].data_ptr(), 0, 0, 0, 0, pkg.LOSS_HUBER, 1.0, out.data_ptr())
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 0.0
