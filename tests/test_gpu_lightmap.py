"""GPU parity of the lightmap bake (LmPack_BakeDesc, K12) against the reference's LmPack_Bake
(src/rendering/lightmap.c:1117-1223). The charting / atlas step LmPack_Pack runs as reference
code on the CPU and is the shared input of both sides."""
import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu


def test_lightmap_bake_matches_reference(oracle):
    from pim_b200 import api
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    desc.bake_mode = 1      # PtBake_Sequential: the integrator's RNG runs on from sample to sample as in BakeFn
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    texels_per_meter = 4.0
    assert oracle.lm_pack(256, texels_per_meter) == 1
    probes, pos, nrm, cnt = oracle.lm_page(0)
    covered = cnt > 0
    assert covered.sum() > 3000
    ref = oracle.Scene()
    scene = api.Scene(desc)
    for c in range(int(np.prod(scene.grid_size))):
        n, pdf, _, _ = ref.dist(c)
        if n:
            scene.set_dist(c, pdf)
    # device copies of the page, same starting state
    g_probes = [np.ascontiguousarray(p.copy()) for p in probes]
    g_pos, g_nrm, g_cnt = pos.copy(), nrm.copy(), cnt.copy()
    for pass_index, spp in ((0, 4), (1, 8)):
        oracle.lm_bake(ref, 1.0, spp, pass_index)
        scene.bake_lightmap(g_probes, g_pos, g_nrm, g_cnt, texels_per_meter, 1.0, spp, pass_index)
        assert np.array_equal(g_cnt, cnt)                      # sample counts advance identically
        assert (g_cnt[~covered] == 0).all()
        for k in range(5):
            a, b = g_probes[k][covered], probes[k][covered]
            # the SG fit is a sequential non-linear recurrence over the texel's samples: one path
            # that takes a different branch moves the texel, so compare texels, not paths
            close = np.abs(a - b).max(axis=1) <= 1e-3 * (np.abs(b).max(axis=1) + 1e-2)
            assert close.mean() >= 0.95, (pass_index, k, close.mean())
            assert np.abs(a.mean(axis=0) - b.mean(axis=0)).max() <= 0.02 * np.abs(b.mean(axis=0)).max()
    # time slicing: with timeSlice = 0.25 about a quarter of the covered texels advance, the same ones
    before = g_cnt.copy()
    oracle.lm_bake(ref, 0.25, 1, 2)
    scene.bake_lightmap(g_probes, g_pos, g_nrm, g_cnt, texels_per_meter, 0.25, 1, 2)
    assert np.array_equal(g_cnt, cnt)
    frac = (g_cnt > before).sum() / covered.sum()
    assert 0.15 < frac < 0.35
    ref.close(); scene.close()


def test_wavefront_bake_statistics(oracle):
    """PtBake_Wavefront: same rays per texel as BakeFn, integrator seeded per ray. Not comparable
    texel by texel with the seeded reference; the fitted lobes agree in the mean and the sample
    counts and time slicing are identical."""
    from pim_b200 import api
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    assert oracle.lm_pack(256, 4.0) == 1
    probes, pos, nrm, cnt = oracle.lm_page(0)
    covered = cnt > 0
    ref = oracle.Scene()
    scene = api.Scene(desc)
    g_probes = [np.ascontiguousarray(p.copy()) for p in probes]
    g_pos, g_nrm, g_cnt = pos.copy(), nrm.copy(), cnt.copy()
    spp = 64
    oracle.lm_bake(ref, 1.0, spp, 0)
    scene.bake_lightmap(g_probes, g_pos, g_nrm, g_cnt, 4.0, 1.0, spp, 0)
    assert np.array_equal(g_cnt, cnt)
    for k in range(5):
        a, b = g_probes[k][covered, :3], probes[k][covered, :3]
        assert np.isfinite(a).all() and (a >= 0).all()
        # the running SG fit lets single bright samples through: compare means with the top 1 % clipped
        q = np.percentile(b, 99)
        ma, mb = np.minimum(a, q).mean(), np.minimum(b, q).mean()
        assert abs(ma - mb) <= 0.08 * mb, (k, ma, mb)
        # lobe weights (w) depend only on the ray directions, which are BakeFn's own: identical
        assert np.allclose(g_probes[k][covered, 3], probes[k][covered, 3], rtol=1e-4, atol=1e-6)
    before = g_cnt.copy()
    oracle.lm_bake(ref, 0.25, 2, 1)
    scene.bake_lightmap(g_probes, g_pos, g_nrm, g_cnt, 4.0, 0.25, 2, 1)
    assert np.array_equal(g_cnt, cnt)
    assert 0.15 < (g_cnt > before).sum() / covered.sum() < 0.35
    ref.close(); scene.close()


def test_device_resident_bake_equals_host_form(oracle):
    """LmPack_BakeDesc_Device on device copies of a page gives bit-identical probes and counts to
    LmPack_BakeDesc on the host page (same kernels, no copies), in both bake modes, over two passes"""
    import torch
    from pim_b200 import api
    for mode in (0, 1):
        desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
        desc.bake_mode = mode
        oracle.set_seed(1)
        oracle.load_scene_desc(desc)
        assert oracle.lm_pack(256, 4.0) == 1
        probes, pos, nrm, cnt = oracle.lm_page(0)
        h_probes = [np.ascontiguousarray(p.copy()) for p in probes]
        h_pos, h_nrm, h_cnt = pos.copy(), nrm.copy(), cnt.copy()
        d_probes = [torch.from_numpy(p.copy()).cuda() for p in probes]
        d_pos, d_nrm, d_cnt = torch.from_numpy(pos.copy()).cuda(), torch.from_numpy(nrm.copy()).cuda(), torch.from_numpy(cnt.copy()).cuda()
        a, b = api.Scene(desc), api.Scene(desc)        # two scenes: a bake also feeds the light-grid learning
        for p in range(2):
            a.bake_lightmap(h_probes, h_pos, h_nrm, h_cnt, 4.0, 1.0, 4, p)
            b.bake_lightmap_device([t.data_ptr() for t in d_probes], d_pos.data_ptr(), d_nrm.data_ptr(), d_cnt.data_ptr(),
                                   256, 4.0, 1.0, 4, p)
        torch.cuda.synchronize()
        assert np.array_equal(d_cnt.cpu().numpy(), h_cnt)
        if mode == 1:       # sequential: deterministic per texel
            for k in range(5):
                assert np.array_equal(d_probes[k].cpu().numpy(), h_probes[k])
        else:               # wavefront: LightOnHit's atomics order may differ between runs; compare the means
            for k in range(5):
                x, y = d_probes[k].cpu().numpy()[h_cnt > 0], h_probes[k][h_cnt > 0]
                assert np.abs(x.mean(axis=0) - y.mean(axis=0)).max() <= 0.02 * np.abs(y.mean(axis=0)).max()
        a.close(); b.close()


def test_bake_writes_the_block_lmpack_save_stores(oracle):
    """LmPack_Save (lightmap.c:1225-1262) writes each lightmap as ONE blob: bytesPerLightmap = 108 * size^2 bytes from
    probes[0], i.e. Lightmap_New's allocation [5 x N float4 probes][N float3 position][N float3 normal][N float counts]
    (lightmap.c:81-109). The bake works in place on arrays laid out like that: here the page is one such block (a byte
    copy of the reference's), the PtLightmapDesc pointers are Lightmap_New's offsets into it, and after the same bake on
    both sides the two blobs are compared region by region — what LmPack_Save would store is the reference's file."""
    import ctypes as C
    from pim_b200 import api
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    desc.bake_mode = 1
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    size = 256
    assert oracle.lm_pack(size, 4.0) == 1
    probes, pos, nrm, cnt = oracle.lm_page(0)
    n = size * size
    # the reference's own block: its five probe arrays, position, normal and counts are consecutive
    base = probes[0].ctypes.data
    assert [p.ctypes.data - base for p in probes] == [16 * n * k for k in range(5)]
    assert pos.ctypes.data - base == 80 * n and nrm.ctypes.data - base == 92 * n and cnt.ctypes.data - base == 104 * n
    bytes_per_lightmap = 108 * n                      # DiskLmPack.bytesPerLightmap
    ref_blob = np.ctypeslib.as_array(C.cast(base, C.POINTER(C.c_uint8)), (bytes_per_lightmap,))
    blob = ref_blob.copy()
    f = blob.view(np.float32)
    views = ([f[4 * n * k:4 * n * (k + 1)].reshape(n, 4) for k in range(5)], f[20 * n:23 * n].reshape(n, 3),
             f[23 * n:26 * n].reshape(n, 3), f[26 * n:27 * n])
    ref = oracle.Scene()
    s = api.Scene(desc)
    for c in range(int(np.prod(s.grid_size))):
        k, pdf, _, _ = ref.dist(c)
        if k:
            s.set_dist(c, pdf)
    oracle.lm_bake(ref, 1.0, 4, 0)
    s.bake_lightmap(views[0], views[1], views[2], views[3], 4.0, 1.0, 4, 0)
    ref.close(); s.close()
    # position, normal and sample counts: byte for byte the reference's blob
    assert np.array_equal(blob[80 * n:], ref_blob[80 * n:])
    # probes: the uncovered texels byte for byte, the covered ones texel for texel within the bake's tolerance
    covered = views[3] > 0
    got = blob[:80 * n].view(np.float32).reshape(5, n, 4); want = ref_blob[:80 * n].view(np.float32).reshape(5, n, 4)
    assert np.array_equal(got[:, ~covered], want[:, ~covered])
    close = np.abs(got[:, covered] - want[:, covered]).max(axis=2) <= 1e-3 * (np.abs(want[:, covered]).max(axis=2) + 1e-2)
    assert close.mean() >= 0.95
