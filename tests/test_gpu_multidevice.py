"""In-process multi-GPU (PtSys_SetDevices): one process, a replica of the scene per device, a host thread per device.
A call spread over two devices has to give what one device gives: a path's random stream follows (pixel or texel,
sample, seed), not the device, and the light-learning counters are merged after every call. Needs two GPUs
(`gpurun --gpus 2`); skipped on a one-GPU box."""
import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from pim_b200 import api as a
    a.load()
    yield a
    a.set_devices(1)


def _render(api, devices, lut):
    api.set_devices(devices)
    s = api.Scene(scenes.cornell_box("spheres", brdf_lut=lut, seed=1))
    cam, dof = api.make_camera([-4, 0, 4], [0, -1, 0]), api.default_dof(auto_focus=False)
    W, H = 192, 108
    out = {}
    tr = api.Trace(W, H)
    tr.trace_spp(s, cam, dof, 0, 6, sums=True)               # by sample index: 3 + 3 samples
    out["spp6"] = [tr.image(k).copy() for k in ("color", "albedo", "normal")]
    tr.trace_spp(s, cam, dof, 6, 1, sums=True)               # fewer samples than devices: by pixel band
    out["spp1"] = [tr.image(k).copy() for k in ("color", "albedo", "normal")]
    tr2 = api.Trace(W, H)
    for f in range(3):                                       # Pt_Trace: one sample per call, running mean on the device
        tr2.trace(s, cam, dof, 1.0 / (f + 1))
    out["frames"] = tr2.image("color").copy()
    out["live"] = np.concatenate([s.dist(c)[3] for c in range(0, int(np.prod(s.grid_size)), 3)])
    tr.close(); tr2.close(); s.close()
    return out


def test_two_devices_render_what_one_renders(api):
    lut = parity.reference_lut()
    one, two = _render(api, 1, lut), _render(api, 2, lut)
    # one sample per pixel lands in each pixel's sum exactly once whichever device traced it: equal bit for bit
    for a, b in zip(one["spp1"], two["spp1"]):
        assert np.array_equal(a, b)
    # several samples per pixel: the same samples in another order of float additions
    for a, b in zip(one["spp6"], two["spp6"]):
        assert np.allclose(a, b, rtol=1e-5, atol=1e-5)
    assert np.allclose(one["frames"], two["frames"], rtol=1e-5, atol=1e-5)
    # the merged light-learning counters are the single device's (integer adds, any order)
    assert np.array_equal(one["live"], two["live"]) and one["live"].sum() > 0
    assert one["spp6"][0].mean() > 0.01


def test_two_devices_bake_what_one_bakes(api, oracle):
    lut = parity.reference_lut()
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    assert oracle.lm_pack(256, 4.0) == 1
    probes, pos, nrm, cnt = oracle.lm_page(0)
    start = ([p.copy() for p in probes], pos.copy(), nrm.copy(), cnt.copy())
    results = []
    for devices in (1, 2):
        for mode in (0, 1):
            api.set_devices(devices)
            d = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
            d.bake_mode = mode
            s = api.Scene(d)
            p = [x.copy() for x in start[0]]; c = start[3].copy()
            for pass_index in range(2):
                s.bake_lightmap(p, start[1], start[2], c, 4.0, 1.0, 4, pass_index)
            s.close()
            results.append((devices, mode, p, c))
    for mode in (0, 1):
        (_, _, p1, c1), (_, _, p2, c2) = [r for r in results if r[1] == mode]
        assert np.array_equal(c1, c2) and (c1[start[3] > 0] == start[3][start[3] > 0] + 8).all()
        for k in range(5):
            assert np.array_equal(p1[k], p2[k]), (mode, k)


def test_refit_on_update_with_a_replica_per_device(api):
    """The "refit" option in multi-GPU mode: every replica keeps its own tree and refits it on its own device; a frame
    spread over the two devices afterwards equals the frame of a scene built from the moved description."""
    lut = parity.reference_lut()

    def desc(shift):
        d = scenes.cornell_box("spheres", brdf_lut=lut, seed=1)
        for e in d.entities:
            if len(e.positions) > 600:
                m = np.array(e.matrix, np.float32).copy(); m[12] += np.float32(shift); e.matrix = m
        return d

    api.set_devices(2)
    api.set_option("refit", 1)
    try:
        api.set_scene_desc(desc(0.0))
        s = api.Scene()
        api.set_scene_desc(desc(0.3))
        assert s.update().bvhRefitted == 1
        fresh = api.Scene(desc(0.3))
        cam, dof = api.make_camera([-4, 0, 4], [0, -1, 0]), api.default_dof(auto_focus=False)
        a, b = api.Trace(160, 90), api.Trace(160, 90)
        a.trace_spp(s, cam, dof, 0, 1, sums=True)            # one sample per pixel: equal bit for bit
        b.trace_spp(fresh, cam, dof, 0, 1, sums=True)
        assert np.array_equal(a.image("color"), b.image("color")) and a.image("color").mean() > 0.0
        a.trace_spp(s, cam, dof, 1, 4, sums=True)            # by sample index over the two devices
        b.trace_spp(fresh, cam, dof, 1, 4, sums=True)
        assert np.allclose(a.image("color"), b.image("color"), rtol=1e-5, atol=1e-5)
        a.close(); b.close(); fresh.close(); s.close()
    finally:
        api.set_option("refit", 0)
        api.set_scene_desc(None)
