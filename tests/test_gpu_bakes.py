"""GPU parity of the callers either side of the path tracer, through the C ABI:
Cubemap_Bake / Cubemap_Convolve (src/rendering/cubemap.c), AmbCube_Bake (src/math/ambcube.c), the sky
bake (render_system.c:403-425 over atmosphere.h) and the screenshot tonemap loop (render_system.c:707-727),
each against the reference's own code in oracle/_ref driven with the same per-item seeds."""
import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu

PATH_AGREE = 0.99       # as in test_gpu_integrator.py: the rest flip a branch on libm-vs-CUDA rounding
SKY9 = [6360e3, 60e3, 1 / 192e3, 1 / 82e3, 1 / 34e3, 1 / 8.5e3, 1 / 48e3, 1 / 1.2e3, 0.758]
SUN = [0.882, 0.195, 0.429]


@pytest.fixture(scope="module")
def api():
    from pim_b200 import api as a
    a.load()
    return a


@pytest.fixture(scope="module")
def lut():
    return parity.reference_lut()


def _pair(api, oracle, lut, kind="boxes"):
    desc = scenes.cornell_box(kind, brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    cells = int(np.prod(s.grid_size))
    for c in range(cells):
        n, pdf = ref.dist(c)[:2]
        if n:
            s.set_dist(c, pdf)
    return s, ref


def test_sky_bake_matches_reference(api, oracle):
    """every texel of the 64^2 x 6 sky within 2e-4 of the reference's Atmosphere() (expf vs __expf)"""
    for size, steps in ((16, 1), (64, 4)):
        cm = api.Cubemap(size)
        cm.bake_sky(api.earth_sky(), SUN, 3800.0, steps)
        want = oracle.bake_sky(size, SKY9, SUN, 3800.0, steps)
        got = cm.colors()
        assert np.isfinite(got).all()
        assert want.max() > 1.0                       # the sun side of the sky is bright
        err = np.abs(got - want) / (np.abs(want) + 1e-3)
        assert err.max() < 2e-4, err.max()
        cm.close()


def test_tonemap_matches_reference(api, oracle):
    """RGBA8 within one code value of the reference's loop; the dither uses the same per-pixel stream"""
    rng = np.random.default_rng(5)
    light = np.zeros((4096, 4), np.float32)
    light[:, :3] = np.exp(rng.uniform(-8, 3, (4096, 3))).astype(np.float32)
    light[:16, :3] = 0.0
    light[16:32, :3] = -1.0                                # negative input clamps to 0
    light[32:48, :3] = 1e6
    for exposure in (1.0, 0.25):
        got = api.tonemap(light, exposure, seed=1, pass_index=3).view(np.uint8).reshape(-1, 4).astype(np.int32)
        want = oracle.tonemap(light, exposure, 3).view(np.uint8).reshape(-1, 4).astype(np.int32)
        assert (got[:, 3] == 255).all() and (want[:, 3] == 255).all()
        d = np.abs(got - want)
        assert d.max() <= 1
        assert (d == 0).mean() > 0.99
    assert len(api.tonemap(np.zeros((0, 4), np.float32))) == 0     # empty image


def test_cubemap_convolve_matches_reference(api, oracle):
    """GGX prefilter of a sky cubemap whose texels span 10..1800: the sample directions come from the same
    streams, and __sinf/__cosf (--use_fast_math) vs libm move the bilinear taps by ~1e-6 of a face, which
    next to the sun is ~1e-4 of the value. Tolerance: median 1e-4, 99.9 % of mip texels 2e-3, all 5e-2."""
    size = 32
    color = oracle.bake_sky(size, SKY9, SUN, 3800.0, 2)
    cm = api.Cubemap(size)
    for f in range(6):
        cm.color(f)[:] = color[f]
    n = oracle.cubemap_elem_count(size)
    assert n == cm.elem_count
    want = np.zeros((6, n, 4), np.float32)
    for p, w in ((0, 1.0), (1, 0.5)):                     # second pass blends into the first
        oracle.cubemap_convolve(color, want, 16, w, p)
        cm.convolve(16, w, seed=1, pass_index=p)
    got = np.stack([cm.convolved(f) for f in range(6)])
    err = np.abs(got - want) / (np.abs(want) + 1e-2)
    assert np.median(err) < 1e-4
    assert np.percentile(err, 99.9) < 2e-3
    assert err.max() < 5e-2
    assert np.allclose(got[:, :n - 1, 3], 1.0, atol=1e-5)   # w = sum(NoL) / sum(NoL)
    assert (got[:, n - 1] == 0).all() and (want[:, n - 1] == 0).all()   # 1x1 mip: direction is 0/0 on both sides
    cm.close()


def test_cubemap_bake_matches_reference_per_path(api, oracle, lut):
    """Cubemap_Bake inside cornell boxes: same jittered ray and same path per face texel"""
    s, ref = _pair(api, oracle, lut)
    size = 24
    cm = api.Cubemap(size)
    want = np.zeros((6, size, size, 3), np.float32)
    origin = [0.3, -0.2, 0.1, 0.0]
    for p, w in ((0, 1.0), (1, 0.5)):
        ref.cubemap_bake(want, origin, w, p)
        cm.bake(s, origin, w, pass_index=p)
        assert parity.path_agreement(cm.colors(), want) >= PATH_AGREE - 0.01 * p
    assert cm.colors().mean() > 0.01
    cm.close(); ref.close(); s.close()


def test_ambcube_bake_matches_reference(api, oracle, lut):
    """AmbCube_Bake: the ordered fit over the same Pt_RayGen probes; a path that differs moves the
    cube by its weight 6/(1+n), so compare the cube to a few of those"""
    s, ref = _pair(api, oracle, lut)
    cube = np.zeros((6, 4), np.float32)
    n_want, want = ref.ambcube_bake(cube, [0, 0, 0, 0], 2000, 0, 0)
    n_got, got = api.ambcube_bake(s, cube, [0, 0, 0, 0], 2000, 0)
    assert n_got == n_want == 1
    assert want[:, :3].max() > 0.01
    assert np.abs(got - want).max() < 0.02 * want[:, :3].max()
    ref.close(); s.close()


def test_bake_entry_points_reject_bad_input(api):
    L = api.load()
    assert L.Cubemap_ConvolvePass(None, 4, 1.0, 1, 0) == -1
    assert b"cubemap" in L.Pt_LastError()
    empty = api.CubemapStruct()
    assert L.Pt_BakeSky(empty, api.earth_sky(), api.Float3(0, 1, 0), 1.0, 1) == -1
    assert L.Pt_Tonemap(None, 8, 1.0, 1, 0, None) == -1


def _embed_case(api, oracle, kind, atlas, tpm):
    desc = scenes.cornell_box(kind, brdf_lut=np.zeros((4, 4, 2), np.float32), seed=1)
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    pages = oracle.lm_pack(atlas, tpm)
    assert pages >= 1
    meshes = oracle.lm_meshes()
    slots = [oracle.lm_slot(k) for k in range(pages)]
    assert len(set(slots)) == pages
    got = api.lm_embed(meshes, slots, atlas)
    covered = 0
    for k in range(pages):
        _, pos, nrm, cnt = oracle.lm_page(k)
        g_pos, g_nrm, g_cnt = got[k]
        # coverage: the 2-texel padding test flips only where a distance rounds across 2.0
        assert (g_cnt != cnt).mean() < 1e-4
        both = (g_cnt > 0) & (cnt > 0)
        covered += int(both.sum())
        # a padding texel exactly half way between two charts has two nearest triangles at distances equal
        # up to rounding; which one wins differs with the arithmetic (both sides build with fast math)
        bad = (np.abs(g_pos[both] - pos[both]).max(axis=1) > 2e-4) | (np.abs(g_nrm[both] - nrm[both]).max(axis=1) > 2e-4)
        assert (not both.any()) or bad.mean() < 2e-3          # metres; positions are O(1..5)
        assert (g_pos[g_cnt == 0] == 0).all() and (g_nrm[g_cnt == 0] == 0).all()
    return pages, covered


def test_lightmap_embed_matches_reference(api, oracle):
    """LmPack_EmbedDesc (EmbedTaskFn, lightmap.c:947-1034) on the atlas the reference's own LmPack_Pack laid
    out: one page, and a small atlas that forces several pages (each drawable's triangles carry their page's slot)"""
    pages, covered = _embed_case(api, oracle, "boxes", 256, 4.0)
    assert pages == 1 and covered > 3000
    pages, covered = _embed_case(api, oracle, "boxes", 256, 12.0)      # five pages, two of them empty
    assert pages > 1 and covered > 30000          # (the reference's packer never returns for 256 / 16.0 or 128 / 8.0)


def test_lightmap_embed_edge_cases(api):
    """no triangles: every texel is empty; unmapped (sky) drawables are skipped; a triangle with no area in
    lightmap space bids only inside its padded box (in the reference sdTriangle2D's sign term is 0 for it, the
    distance comes out negative everywhere and it floods the page; attributes are 0/0 on both sides)"""
    out = api.lm_embed([], [7], 16)
    assert (out[0][2] == 0).all() and (out[0][0] == 0).all()
    pos = np.array([[0, 0, 0, 1], [1, 0, 0, 1], [0, 1, 0, 1]], np.float32)
    nrm = np.tile(np.array([[0, 0, 1, 0]], np.float32), (3, 1))
    uv = np.array([[0, 0, 0.25, 0.25], [0, 0, 0.75, 0.25], [0, 0, 0.25, 0.75]], np.float32)
    tex = np.tile(np.array([[0, 0, 0, 7]], np.int32), (3, 1))
    sky = api.lm_embed([(pos, nrm, uv, tex, 1 << 1)], [7], 16)
    assert (sky[0][2] == 0).all()
    other = api.lm_embed([(pos, nrm, uv, tex, 0)], [3], 16)          # slot does not match any page
    assert (other[0][2] == 0).all()
    one = api.lm_embed([(pos, nrm, uv, tex, 0)], [7], 16)
    assert one[0][2].sum() > 30
    inside = one[0][2].reshape(16, 16)[5, 5] == 1.0                  # texel centre (5.5, 5.5) / 16 is inside
    assert inside
    p = one[0][0].reshape(16, 16, 3)[5, 5]
    assert np.allclose(p, [(5.5 / 16 - 0.25) / 0.5, (5.5 / 16 - 0.25) / 0.5, 0.0], atol=1e-5)
    assert np.allclose(one[0][1].reshape(16, 16, 3)[5, 5], [0, 0, 1], atol=1e-6)
    degenerate = uv.copy(); degenerate[:, 2:] = 0.5
    d = api.lm_embed([(pos, nrm, degenerate, tex, 0)], [7], 16)
    assert 4 <= d[0][2].sum() <= 49


def test_rome_repack_matches_reference(api, oracle):
    """Pt_RepackRome (CreateRomeTexture's loop, gltf_model.c:1118-1139): three float4 images of different
    sizes onto one RGBA8 texture; every channel within one code value, 99 % equal; 1x1 defaults work"""
    g = np.random.default_rng(11)
    mr = g.uniform(0, 1, (48, 64, 4)).astype(np.float32)
    occ = g.uniform(0, 1, (16, 16, 4)).astype(np.float32)
    em = (g.uniform(0, 1, (32, 8, 4)) * np.array([150.0, 80.0, 20.0, 0.0])).astype(np.float32)
    em[:8] = 0.0                                       # no emission: packs to exactly 0
    for size in (64, 256):
        got = api.repack_rome(mr, occ, em, size).view(np.uint8).reshape(-1, 4).astype(np.int32)
        want = oracle.repack_rome(mr, occ, em, size).view(np.uint8).reshape(-1, 4).astype(np.int32)
        d = np.abs(got - want)
        assert d.max() <= 1 and (d == 0).mean() > 0.99
        assert want[:, 3].max() > 200 and (want[:size, 3] == 0).all()
    d_mr = np.array([[[0.0, 0.5, 0.0, 0.0]]], np.float32); d_occ = np.array([[[1.0, 0, 0, 0]]], np.float32)
    d_em = np.zeros((1, 1, 4), np.float32)
    got = api.repack_rome(d_mr, d_occ, d_em, 4).view(np.uint8).reshape(-1, 4)
    assert (got == np.array([128, 255, 0, 0], np.uint8)).all()
    assert (got == oracle.repack_rome(d_mr, d_occ, d_em, 4).view(np.uint8).reshape(-1, 4)).all()
    assert api.load().Pt_RepackRome(None, api.Int2(1, 1), None, api.Int2(1, 1), None, api.Int2(1, 1), 4, None) == -1


def test_chart_masks_and_atlas_match_reference(oracle):
    """N1: ChartMaskFn and AtlasFn / atlas_search of LmPack_Pack on the device (LmPack_ChartMasksDesc, LmPack_AtlasDesc)
    against the reference's own run on one task thread, whose intermediate state the oracle records: the charts moved
    to (2, 2), their areas, mask sizes and masks texel for texel, and then, for the charts in the reference's sorted
    order, the atlas and translation every chart gets. Two scenes: cornell boxes at 16 texels/m (28 charts) and the
    cornell spheres (thousands of small charts)."""
    from pim_b200 import api
    for variant, size, tpm in (("boxes", 1024, 16.0), ("boxes", 256, 8.0)):
        desc = scenes.cornell_box(variant, brdf_lut=np.zeros((4, 4, 2), np.float32), seed=1)
        oracle.set_seed(1)
        oracle.load_scene_desc(desc)
        charts, ref_atlases = oracle.lm_pack_capture(size, tpm, 0.1, 15.0)
        assert len(charts) >= 20
        first = np.zeros(len(charts) + 1, np.int32)
        first[1:] = np.cumsum([len(c["tri_before"]) for c in charts])
        tri = np.concatenate([c["tri_before"] for c in charts]).astype(np.float32)
        moved, area, msize, masks = api.lm_chart_masks(tri, first)
        assert np.array_equal(moved, np.concatenate([c["tri_after"] for c in charts]))
        assert np.array_equal(area, np.array([c["area"] for c in charts], np.float32))
        differing = total = 0
        for c, ch in enumerate(charts):
            assert masks[c].shape == ch["mask"].shape
            differing += int((masks[c] != ch["mask"]).sum()); total += masks[c].size
            assert set(np.unique(masks[c])) <= {0, 255}
        assert differing <= 1e-5 * total, (differing, total)       # the same arithmetic, rounded the same way
        # AtlasFn's order: the reference's chart_sort
        order = np.argsort([c["sorted_index"] for c in charts])
        s_first = np.zeros(len(charts) + 1, np.int32)
        s_first[1:] = np.cumsum([len(charts[i]["tri_after"]) for i in order])
        s_tri = np.concatenate([charts[i]["tri_after"] for i in order]).astype(np.float32)
        idx, tr, used = api.lm_atlas(s_tri, s_first, area[order], msize[order], size)
        want_idx = np.array([charts[i]["atlas"] for i in order]); want_tr = np.array([charts[i]["translation"] for i in order])
        if differing == 0:
            assert np.array_equal(idx, want_idx) and np.array_equal(tr, want_tr)
        else:       # a rim texel apart: the first chart after it may shift, and everything packed later with it
            assert (idx == want_idx).mean() > 0.5
        assert used == max(want_idx) + 1 and used <= ref_atlases
        # nothing overlaps: paint the device's placements with the device's masks
        canvas = np.zeros((used, size, size), np.uint16)
        for k, i in enumerate(order):
            m = masks[i] != 0
            h, w = m.shape
            canvas[idx[k], tr[k, 1]:tr[k, 1] + h, tr[k, 0]:tr[k, 0] + w] += m
        assert canvas.max() == 1
