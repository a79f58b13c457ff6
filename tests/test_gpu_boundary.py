"""The drop-in boundary driven the way the engine drives it (render_system.c:118-134, 150-159, 181-213,
277-303): PtSys_Init -> PtSys_SetSceneDesc (the adapter's stand-in for the engine tables) -> PtScene_New()
-> PtTrace_New -> Pt_Trace per frame -> Pt_TraceRay -> LmPack_Bake over the registered pack ->
PtScene_Update after the entities changed -> PtScene_Del -> PtSys_Shutdown; every step checked against the
reference's own functions (oracle/_ref) with the same seeds."""
import ctypes as C

import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def api():
    from pim_b200 import api as a
    a.load()
    return a


def test_engine_call_sequence(api, oracle):
    L = api.load()
    lut = parity.reference_lut()
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    desc.bake_mode = 1                              # sequential bake: comparable texel for texel
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    assert oracle.lm_pack(256, 4.0) == 1            # LmPack_Pack by the reference: the engine's pack
    probes, pos, nrm, cnt = oracle.lm_page(0)
    ref = oracle.Scene()

    L.PtSys_Init()
    api.set_scene_desc(desc)                        # copies: the numpy arrays behind it are gone after this line
    s = api.Scene()                                 # PtScene_New(void)
    assert s.stats.vertCount == ref.info.vertCount and s.stats.emissiveCount == ref.info.emissiveCount
    for c in range(int(np.prod(s.grid_size))):
        n, pdf, _, _ = ref.dist(c)
        if n:
            s.set_dist(c, pdf)

    # frames: Pt_Trace with sampleWeight = 1/n as PathTrace() does (render_system.c:286-303)
    W, H, frames = 96, 54, 3
    cam, dof = api.make_camera([-4, 0, 4], [0, -1, 0]), api.default_dof(auto_focus=False)
    ocam, odof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    odof.autoFocus = 0
    tr = api.Trace(W, H)
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
    for f in range(frames):
        tr.trace(s, cam, dof, 1.0 / (f + 1))
        ref.trace(col, alb, nor, 1.0 / (f + 1), ocam, odof, f)
        assert parity.path_agreement(tr.image("color"), col) >= 0.985, f
        assert parity.path_agreement(tr.image("normal"), nor) >= 0.985, f

    # single-ray forms (correctness only; the scene's pass counter is at `frames`)
    ro, rd = parity.box_rays(8, seed=2, extent=3.0)
    hits = ref.intersect(ro, rd)
    for i in range(8):
        one = L.Pt_Intersect(s.h, api._f4(ro[i]), api._f4(rd[i]), 0.0, 8388608.0)
        assert one.iVert == hits["iVert"][i] and one.type == hits["type"][i]
        if one.type:
            assert abs(one.wuvt.w - hits["wuvt"][i, 3]) <= 1e-5 * hits["wuvt"][i, 3]
    r = L.Pt_TraceRay(s.h, api._f4(ro[0]), api._f4(rd[0]))
    want = ref.trace_rays(ro[:1], rd[:1], frames)[0]
    assert np.allclose([r.color[0], r.color[1], r.color[2]], want[0:3], rtol=1e-3, atol=1e-3)
    assert np.allclose([r.normal[0], r.normal[1], r.normal[2]], want[6:9], rtol=1e-3, atol=1e-3)

    # LmPack_Bake(scene, timeSlice, spp) over the registered pack, results in the engine's arrays
    g_probes = [np.ascontiguousarray(p.copy()) for p in probes]
    g_pos, g_nrm, g_cnt = pos.copy(), nrm.copy(), cnt.copy()
    api.set_lm_pack([(g_probes, g_pos, g_nrm, g_cnt)], 4.0)
    s.bake(1.0, 4)                                   # pass = the scene's counter: frames + 1
    oracle.lm_bake(ref, 1.0, 4, frames + 1)
    covered = cnt > 0
    assert np.array_equal(g_cnt, cnt) and (g_cnt[~covered] == 0).all()
    for k in range(5):
        a, b = g_probes[k][covered], probes[k][covered]
        close = np.abs(a - b).max(axis=1) <= 1e-3 * (np.abs(b).max(axis=1) + 1e-2)
        assert close.mean() >= 0.95, (k, close.mean())
    s.bake(1.0, 0)                                   # spp is clamped to 1 (lightmap.c:1217)
    assert (g_cnt[covered] == cnt[covered] + 1).all()
    api.set_lm_pack([], 0.0)
    s.bake(1.0, 1)                                   # empty pack: nothing to do, no error

    # the entities change: registering a new description makes PtScene_Update rebuild in place
    desc2 = scenes.cornell_box("spheres", brdf_lut=lut, seed=1)
    api.set_scene_desc(desc2)
    before = s.stats.vertCount
    st = s.update()
    oracle.load_scene_desc(desc2)
    ref2 = oracle.Scene()
    assert st.vertCount == ref2.info.vertCount and st.vertCount != before
    ro, rd = parity.box_rays(20000, seed=4)
    assert parity.hit_agreement(s.intersect(ro, rd), ref2.intersect(ro, rd)) >= 0.9999
    tr.trace(s, cam, dof, 1.0)                       # the same handle keeps working after the rebuild
    assert tr.image("color").mean() > 0.0
    ref2.close(); ref.close()

    tr.close()
    s.close()                                        # PtScene_Del
    api.set_scene_desc(None)
    with pytest.raises(api.PtError):
        api.Scene()                                  # PtScene_New with nothing registered: NULL + message
    L.PtSys_Shutdown()
    L.PtSys_Init()


def test_overlapping_batches_on_two_streams(api):
    """Two closest-hit (and two any-hit) batches in flight on different streams of one scene: each call has its
    own ray cursor, so the results are those of the calls made one after the other."""
    import torch
    L = api.load()
    s = api.Scene(scenes.cornell_box("spheres", brdf_lut=parity.reference_lut(), seed=1))
    assert s.stats.bvhNodeCount >= 2048              # the persistent batch kernels (they fetch rays through the cursor)
    dev = torch.device("cuda", 0)
    n = 1 << 20
    rays = []
    for seed in (41, 42):
        ro, rd = parity.box_rays(n, seed=seed)
        rd[:, 3] = np.random.default_rng(seed).uniform(0.5, 8.0, n)
        rays.append((torch.from_numpy(ro).to(dev), torch.from_numpy(rd).to(dev)))
    serial_hits, serial_occ = [], []
    for ro, rd in rays:
        h = torch.zeros((n, 12), dtype=torch.float32, device=dev); o = torch.zeros(n, dtype=torch.uint8, device=dev)
        assert L.Pt_IntersectN_Device(s.h, ro.data_ptr(), rd.data_ptr(), 0.0, h.data_ptr(), n, None) == 0
        assert L.Pt_OccludedN_Device(s.h, ro.data_ptr(), rd.data_ptr(), 0.0, o.data_ptr(), n, None) == 0
        torch.cuda.synchronize()
        serial_hits.append(h.cpu().numpy()); serial_occ.append(o.cpu().numpy())
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    for rep in range(3):
        hs = [torch.full((n, 12), -7.0, dtype=torch.float32, device=dev) for _ in range(2)]
        os_ = [torch.full((n,), 9, dtype=torch.uint8, device=dev) for _ in range(2)]
        torch.cuda.synchronize()
        for k in range(2):
            assert L.Pt_IntersectN_Device(s.h, rays[k][0].data_ptr(), rays[k][1].data_ptr(), 0.0, hs[k].data_ptr(), n,
                                          streams[k].cuda_stream) == 0
        for k in range(2):
            assert L.Pt_OccludedN_Device(s.h, rays[k][0].data_ptr(), rays[k][1].data_ptr(), 0.0, os_[k].data_ptr(), n,
                                         streams[k].cuda_stream) == 0
        torch.cuda.synchronize()
        for k in range(2):
            assert np.array_equal(hs[k].cpu().numpy().view(np.uint32), serial_hits[k].view(np.uint32)), (rep, k)
            assert np.array_equal(os_[k].cpu().numpy(), serial_occ[k]), (rep, k)
    s.close()


def test_occluded_and_point_query_match_reference(api, oracle):
    """any hit against the reference's RtcOccluded16 (path_tracer.c:524-553 over the Embree restatement) and the
    nearest-surface query against its RtcPointQuery (:597-616), on cornell spheres and a 200 K-triangle soup"""
    lut = parity.reference_lut()
    for name in ("spheres", "soup"):
        if name == "spheres":
            desc = scenes.cornell_box("spheres", brdf_lut=lut, seed=1)
            ro, rd = parity.box_rays(200000, seed=11)
            rd[:, 3] = np.random.default_rng(5).uniform(0.05, 8.0, len(rd))
            pts = np.zeros((20000, 4), np.float32)
            pts[:, :3] = np.random.default_rng(6).uniform(-4.9, 4.9, (len(pts), 3))
            pts[:, 3] = np.random.default_rng(7).uniform(0.05, 2.0, len(pts))
        else:
            desc = scenes.random_soup(200_000, seed=5, extent=20.0)
            desc.brdf_lut = lut
            ro, rd = scenes.random_rays(200000, seed=6, extent=20.0, occlusion=True)
            pts = np.zeros((20000, 4), np.float32)
            pts[:, :3] = np.random.default_rng(6).uniform(0, 20, (len(pts), 3))
            pts[:, 3] = np.random.default_rng(7).uniform(0.05, 1.0, len(pts))
        s = api.Scene(desc)
        oracle.set_seed(1)
        oracle.load_scene_desc(desc)
        ref = oracle.Scene()
        got, want = s.occluded(ro, rd).astype(bool), ref.occluded(ro, rd).astype(bool)
        assert (got == want).mean() >= 0.9999, (name, (got == want).mean())
        assert 0.02 < want.mean() < 0.98
        d_got = s.point_query(pts)
        d_want, prim = ref.point_query(pts)
        none = prim < 0
        assert ((d_got >= 8388608.0) == none).all(), name
        assert np.abs(d_got[~none] - d_want[~none]).max() <= 1e-5, name
        assert 0.05 < none.mean() < 0.95 or name == "spheres"
        ref.close(); s.close()


def test_light_counters_summed_over_two_replicas_equal_one_scene(api):
    """PtScene_LightCounters, the hook one process per GPU uses to keep its replicas learning as one (SURVEY.md 8(e),
    Dist1D.live path_tracer.c:1709-1733): two scenes stand in for two ranks, each renders its own sample indices from
    the same light grid, and base + (what A added) + (what B added) equals, counter for counter, what ONE scene
    rendering both ranges in a single call holds. Written back, both replicas carry exactly those counters."""
    import torch
    L = api.load()
    dev = torch.device("cuda", 0)
    w, h, spp = 192, 128, 4
    cam = api.make_camera([-4.0, 0.0, 4.0], [0.0, -1.0, 0.0], 90.0)
    dof = api.default_dof(auto_focus=False)

    def scene():
        return api.Scene(scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1))

    def trace(s, begin, count):
        img = [torch.zeros(w * h * 3, dtype=torch.float32, device=dev) for _ in range(3)]
        assert L.Pt_TraceSpp_Device(s.h, C.byref(cam), C.byref(dof), w, h, begin, count,
                                    img[0].data_ptr(), img[1].data_ptr(), img[2].data_ptr(), None) == 0, api.last_error()
        torch.cuda.synchronize()
        return img

    def counters(s):
        t = torch.zeros(s.light_counters(), dtype=torch.int32, device=dev)
        assert s.light_counters(t.data_ptr(), False) == t.numel()
        torch.cuda.synchronize()
        return t

    a, b, one = scene(), scene(), scene()
    n = a.light_counters()
    assert n == one.light_counters() and n > 0
    base = counters(a)
    assert torch.equal(base, counters(b)) and torch.equal(base, counters(one))
    img_a = trace(a, 0, spp); img_b = trace(b, spp, spp); img_one = trace(one, 0, 2 * spp)
    da, db = counters(a) - base, counters(b) - base
    assert int(da.sum()) > 0 and int(db.sum()) > 0 and not torch.equal(da, db)
    merged = base + da + db
    assert torch.equal(merged, counters(one))
    for s in (a, b):
        assert s.light_counters(merged.data_ptr(), True) == n
        assert torch.equal(counters(s), merged)
    # the two halves are the one call's paths: same light grid during the call, disjoint sample indices
    for k in range(3):
        assert torch.allclose(img_a[k] + img_b[k], img_one[k], rtol=1e-5, atol=1e-5)
    # and from here on the replicas pick lights exactly as the one scene does
    nxt_a, nxt_one = trace(a, 2 * spp, spp), trace(one, 2 * spp, spp)
    for k in range(3):
        assert torch.allclose(nxt_a[k], nxt_one[k], rtol=1e-5, atol=1e-5)
    assert torch.equal(counters(a), counters(one))
    for s in (a, b, one):
        s.close()


def test_pageable_host_buffers_through_the_copy_ring_equal_the_direct_forms(api):
    """Host-buffer calls with pageable arrays of 2 MB or more go through the page-locked ring (hostcopy.h: parallel
    streaming memcpy into two slots, graded chunk sizes); page-locked arrays and device pointers do not. Same bytes
    either way: an odd-sized ray batch against the device form, a screenshot against the same call on page-locked
    memory, and sizes around the ring's threshold and slot size."""
    import torch
    L = api.load()
    dev = torch.device("cuda", 0)
    s = api.Scene(scenes.cornell_box("spheres", brdf_lut=parity.reference_lut(), seed=1))
    n = 3_000_001                                    # 48 MB of origins, 144 MB of hits: several slots, a ragged tail
    ro, rd = parity.box_rays(n, seed=77)
    got = s.intersect(ro, rd)
    d_ro, d_rd = torch.from_numpy(ro).to(dev), torch.from_numpy(rd).to(dev)
    d_hits = torch.zeros((n, 12), dtype=torch.float32, device=dev)
    assert L.Pt_IntersectN_Device(s.h, d_ro.data_ptr(), d_rd.data_ptr(), 0.0, d_hits.data_ptr(), n, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(got.view(np.uint32).reshape(n, 12), d_hits.cpu().numpy().view(np.uint32))
    occ = s.occluded(ro, rd)
    d_occ = torch.zeros(n, dtype=torch.uint8, device=dev)
    assert L.Pt_OccludedN_Device(s.h, d_ro.data_ptr(), d_rd.data_ptr(), 0.0, d_occ.data_ptr(), n, None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(occ, d_occ.cpu().numpy())
    s.close()
    # tonemap: pageable numpy arrays against page-locked torch tensors, sizes around 2 MB (threshold) and 8 MB (slot)
    rng = np.random.default_rng(5)
    for pixels in (131_071, 131_072, 131_073, 524_288 + 3, 1920 * 1080, 2_097_152 + 5):
        light = (rng.random((pixels, 4), np.float32) * 4.0).astype(np.float32)
        out = np.zeros(pixels, np.uint32)
        assert L.Pt_Tonemap(light.ctypes.data, pixels, 1.0, 1, 0, out.ctypes.data) == 0, api.last_error()
        p_light = torch.from_numpy(light).pin_memory(); p_out = torch.zeros(pixels, dtype=torch.int32).pin_memory()
        assert L.Pt_Tonemap(p_light.data_ptr(), pixels, 1.0, 1, 0, p_out.data_ptr()) == 0, api.last_error()
        assert np.array_equal(out, p_out.numpy().view(np.uint32)), pixels


def test_refit_on_update_returns_the_hits_of_a_fresh_build(api):
    """Pt_SetOption("refit", 1) ("refit path for unchanged topology"): after the entities moved,
    PtScene_Update keeps the tree's topology, rewrites the packed triangles and refits the node boxes level by level.
    Closest hits do not depend on the tree, so they must equal, bit for bit, those of a scene built from the moved
    description; a description of another shape falls back to the rebuild."""
    L = api.load()
    lut = parity.reference_lut()

    def moved(step):
        d = scenes.cornell_box("spheres", brdf_lut=lut, seed=1)
        rng = np.random.default_rng(100 + step)
        for k, e in enumerate(d.entities):
            m = np.array(e.matrix, np.float32).copy()
            if step and len(e.positions) > 600:         # the spheres: move them about, the second one a lot
                m[12:15] += np.float32(0.35 * step) * rng.uniform(-1.0, 1.0, 3).astype(np.float32)
            e.matrix = m
            if step:                                    # and every vertex of every entity a little
                e.positions = e.positions.copy()
                e.positions[:, :3] += rng.uniform(-0.01, 0.01, (len(e.positions), 3)).astype(np.float32)
        return d

    ro, rd = parity.box_rays(400_000, seed=21)
    assert L.Pt_SetOption(b"refit", 1) == 0
    try:
        api.set_scene_desc(moved(0))
        s = api.Scene()
        assert s.stats.bvhRefitted == 0
        first = s.intersect(ro, rd)
        nodes = s.stats.bvhNodeCount
        for step in (1, 2, 0):                          # moved, moved further, back where it was built
            d = moved(step)
            api.set_scene_desc(d)
            st = s.update()
            assert st.bvhRefitted == 1 and st.bvhNodeCount == nodes
            fresh = api.Scene(d)
            assert fresh.stats.bvhRefitted == 0
            got, want = s.intersect(ro, rd), fresh.intersect(ro, rd)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), step
            assert np.array_equal(s.occluded(ro, rd), fresh.occluded(ro, rd)), step
            if step == 0:
                assert np.array_equal(got.view(np.uint32), first.view(np.uint32))
            fresh.close()
        # a frame after the refit: the light grid, emissive tables and pool were set up anew around the kept tree
        tr = api.Trace(96, 54)
        cam, dof = api.make_camera([-4, 0, 4], [0, -1, 0]), api.default_dof(auto_focus=False)
        tr.trace(s, cam, dof, 1.0)
        assert np.isfinite(tr.image("color")).all() and tr.image("color").mean() > 0.0
        tr.close()
        # another shape: rebuilt, not refitted
        api.set_scene_desc(scenes.cornell_box("boxes", brdf_lut=lut, seed=1))
        st = s.update()
        assert st.bvhRefitted == 0 and st.bvhNodeCount != nodes
        s.close()
    finally:
        L.Pt_SetOption(b"refit", 0)
        api.set_scene_desc(None)
