"""GPU parity of the integrator: Pt_Trace / Pt_TraceSpp / Pt_TraceRayN / Pt_RayGen / lightmap
bake through the C ABI vs the reference's Pt_TraceRay driven with the same per-path seeds."""
import os

import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu

PATH_AGREE = 0.99   # fraction of paths whose radiance matches the reference path for path
                    # (the rest flip a branch on libm-vs-CUDA rounding of sinf/cosf/logf)


@pytest.fixture(scope="module")
def api():
    from pim_b200 import api as a
    a.load()
    return a


@pytest.fixture(scope="module")
def lut():
    return parity.reference_lut()


def _copy_grid(scene, pdf, lens):
    """Give the device scene the oracle's light-grid pdfs, so that per-path comparisons test the
    integrator alone (the device-built grid is checked on its own in test_gpu_parity.py)."""
    for c in range(len(lens)):
        if lens[c]:
            scene.set_dist(c, pdf[c])


def _cams(api, oracle=None):
    cam, dof = api.make_camera([-4, 0, 4], [0, -1, 0]), api.default_dof(auto_focus=False)
    if oracle is None:
        return cam, dof
    ocam, odof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    odof.autoFocus = 0
    return cam, dof, ocam, odof


def test_frame_matches_golden_per_path(api, lut):
    """Two reference-style frames (1 spp each, relearn between) vs the committed oracle frames."""
    g = np.load(os.path.join(parity.GOLDEN, "cornell_boxes_64x36.npz"))
    s = api.Scene(scenes.cornell_box("boxes", brdf_lut=lut, seed=1))
    _copy_grid(s, g["pdf"], g["lens"])
    cam, dof = _cams(api)
    tr = api.Trace(64, 36)
    for p in range(2):
        tr.trace_spp(s, cam, dof, p, 1, sums=True)
        assert parity.path_agreement(tr.image("color"), g["color%d" % p]) >= PATH_AGREE
        assert parity.path_agreement(tr.image("albedo"), g["albedo%d" % p]) >= PATH_AGREE
        assert parity.path_agreement(tr.image("normal"), g["normal%d" % p]) >= PATH_AGREE
    tr.close()
    s.close()


@pytest.mark.parametrize("variant", ["boxes", "spheres"])
def test_frames_match_oracle_per_path(api, lut, oracle, variant):
    desc = scenes.cornell_box(variant, brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    cam, dof, ocam, odof = _cams(api, oracle)
    cells = int(np.prod(s.grid_size))
    dists = [ref.dist(c) for c in range(cells)]
    _copy_grid(s, [d[1] for d in dists], [d[0] for d in dists])
    W, H = 96, 54
    tr = api.Trace(W, H)
    for p in range(3):
        col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
        ref.trace(col, alb, nor, 1.0, ocam, odof, p)
        tr.trace_spp(s, cam, dof, p, 1, sums=True)
        frac = parity.path_agreement(tr.image("color"), col)
        assert frac >= (PATH_AGREE if variant == "boxes" else 0.98), (p, frac)
    # the adaptive light grid learned the same counters (integer atomics, order free)
    diff = tot = 0
    for c in range(0, cells, 7):
        n0, _, _, live0 = s.dist(c)
        n1, _, _, live1 = ref.dist(c)
        assert n0 == n1
        diff += np.abs(live0.astype(np.int64) - live1.astype(np.int64)).sum()
        tot += live1.sum()
    assert diff <= 0.02 * max(tot, 1), (diff, tot)
    tr.close(); ref.close(); s.close()


def test_pt_trace_running_mean_and_dof(api, lut):
    """Pt_Trace keeps the reference's semantics: lerp by sampleWeight, autofocus easing."""
    s = api.Scene(scenes.cornell_box("boxes", brdf_lut=lut, seed=1))
    cam = api.make_camera([-4, 0, 4], [0, -1, 0])
    dof = api.default_dof(auto_focus=True)
    f0 = dof.focalLength
    tr = api.Trace(64, 36)
    n = 8
    for i in range(n):
        tr.trace(s, cam, dof, 1.0 / (i + 1))
    mean_pt = tr.image("color").copy()
    assert dof.focalLength != f0            # DofUpdate eased toward the centre hit distance
    dof2 = api.default_dof(auto_focus=False)
    tr2 = api.Trace(64, 36)
    tr2.trace_spp(s, cam, dof2, 0, n)
    a, b = mean_pt.mean(), tr2.image("color").mean()
    assert abs(a - b) / b < 0.15
    tr.close(); tr2.close(); s.close()


def test_trace_spp_statistics_vs_oracle(api, lut, oracle):
    """Converged-image parity: 64 spp vs oracle 64 spp, within 1.5x the oracle's own
    two-seed noise (SURVEY.md 8(d) tolerance), plus CalcStdDev."""
    W, H, spp = 80, 45, 64
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    cam, dof, ocam, odof = _cams(api, oracle)
    imgs = []
    for seed in (1, 2):
        oracle.set_seed(seed)
        oracle.load_scene_desc(desc)
        ref = oracle.Scene()
        col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
        for p in range(spp):
            ref.trace(col, alb, nor, 1.0 / (p + 1), ocam, odof, p)
        ref.close()
        imgs.append(col)
    oracle.set_seed(1)
    desc3 = scenes.cornell_box("boxes", brdf_lut=lut, seed=3)
    s = api.Scene(desc3)
    tr = api.Trace(W, H)
    done = 0
    while done < spp:
        tr.trace_spp(s, cam, dof, done, 16)
        done += 16
    got = tr.image("color").copy()
    tr.close(); s.close()

    def rmse(a, b):
        return np.sqrt(np.mean((np.minimum(a, 4.0) - np.minimum(b, 4.0)) ** 2, axis=(0, 1)))
    floor = rmse(imgs[0], imgs[1])
    for ref_img in imgs:
        assert (rmse(got, ref_img) <= 1.5 * floor + 1e-3).all(), (rmse(got, ref_img), floor)
    m_ref = 0.5 * (imgs[0].mean() + imgs[1].mean())
    assert abs(got.mean() - m_ref) / m_ref <= 1.5 * abs(imgs[0].mean() - imgs[1].mean()) / m_ref + 0.02
    sd = [oracle.stddev(i) for i in imgs]
    assert abs(oracle.stddev(got) - np.mean(sd)) / np.mean(sd) < 0.25


def test_trace_rays_match_oracle(api, lut, oracle):
    """Pt_TraceRayN (both the one-thread-per-path kernel and the wavefront) vs Pt_TraceRay."""
    desc = scenes.cornell_box("spheres", brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    cells = int(np.prod(s.grid_size))
    dists = [ref.dist(c) for c in range(cells)]
    _copy_grid(s, [d[1] for d in dists], [d[0] for d in dists])
    for n in (1000, 20000):
        ro, rd = parity.box_rays(n, seed=21, extent=4.0)
        want = ref.trace_rays(ro, rd, 5)
        got = s.trace_rays(ro, rd, 5)
        assert parity.path_agreement(got[:, 0:3], want[:, 0:3]) >= 0.98
        assert parity.path_agreement(got[:, 6:9], want[:, 6:9]) >= 0.98
    ref.close(); s.close()


def test_raygen_matches_oracle(api, lut, oracle):
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    cells = int(np.prod(s.grid_size))
    dists = [ref.dist(c) for c in range(cells)]
    _copy_grid(s, [d[1] for d in dists], [d[0] for d in dists])
    want_c, want_d = ref.raygen([0, 0, 0, 0], 5000, 0)
    got_c, got_d = s.raygen([0, 0, 0, 0], 5000)
    assert np.abs(got_d - want_d).max() < 1e-5
    assert parity.path_agreement(got_c[:, :3], want_c[:, :3]) >= PATH_AGREE
    ref.close(); s.close()


def test_empty_scene_and_sky(api, lut):
    """no geometry: every path misses and returns the sky (GetSky) or zero"""
    desc = scenes.SceneDesc(brdf_lut=lut)
    s = api.Scene(desc)
    cam, dof = _cams(api)
    tr = api.Trace(16, 9)
    tr.trace_spp(s, cam, dof, 0, 2)
    assert (tr.image("color") == 0).all()
    tr.close(); s.close()
    sky = np.full((6, 8, 8, 3), 0.25, np.float32)
    desc = scenes.SceneDesc(brdf_lut=lut, sky=sky)
    s = api.Scene(desc)
    tr = api.Trace(16, 9)
    tr.trace_spp(s, cam, dof, 0, 2)
    assert np.allclose(tr.image("color"), 0.25, atol=1e-6)
    tr.close(); s.close()


def test_dense_medium_matches_oracle(api, lut, oracle):
    """Fog with a 4 m (constant) + 6 m (3-octave noise) mean free path and an anisotropic two-lobe phase
    function, so that most paths scatter inside the medium: ScatterRay, Media_Sample and its gradient noise,
    SamplePhaseDir, CalcTransmittance on the NEE rays, and the medium's AOV contributions — the default
    40 km fog almost never takes those branches. Per path against the reference's Pt_TraceRay, both through
    the wavefront and the one-thread-per-path kernel, and per pixel for the three images of a frame."""
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    cells = int(np.prod(s.grid_size))
    dists = [ref.dist(c) for c in range(cells)]
    _copy_grid(s, [d[1] for d in dists], [d[0] for d in dists])
    m = s.set_media(constantMfp=4.0, noiseMfp=6.0, constantColor=(0.3, 0.5, 0.8), noiseColor=(0.7, 0.5, 0.4),
                    absorption=0.2, noiseOctaves=3, noiseScale=1.5, noiseHeight=3.0, noiseFreq=0.7,
                    phaseDirA=0.6, phaseDirB=-0.3, phaseBlend=0.4)
    ref.set_media(m)
    # CalcMajorant: the red channel has the shortest free paths, 2 + 0.3 * 6 and 3 + 0.7 * 9 metres
    assert abs(1.0 / ref.info.rcpMajorant - 2.0 * 1.2 * (1 / 3.8 + 1 / 9.3)) < 1e-3
    for n in (1000, 20000):
        ro, rd = parity.box_rays(n, seed=33, extent=4.0)
        want = ref.trace_rays(ro, rd, 7)
        got = s.trace_rays(ro, rd, 7)
        # more random decisions per path than in clear air: a few more paths flip a branch on rounding
        assert parity.path_agreement(got[:, 0:3], want[:, 0:3]) >= 0.97
        assert parity.path_agreement(got[:, 3:6], want[:, 3:6]) >= 0.97
        assert parity.path_agreement(got[:, 6:9], want[:, 6:9]) >= 0.97
    # the fog does scatter: the albedo AOV of paths differs from the clear-air run
    clear = api.Scene(desc)
    _copy_grid(clear, [d[1] for d in dists], [d[0] for d in dists])
    ro, rd = parity.box_rays(2000, seed=33, extent=4.0)
    assert np.abs(clear.trace_rays(ro, rd, 7)[:, 3:6] - s.trace_rays(ro, rd, 7)[:, 3:6]).mean() > 0.01
    clear.close()
    cam, dof, ocam, odof = _cams(api, oracle)
    W, H = 64, 36
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
    ref.trace(col, alb, nor, 1.0, ocam, odof, 0)
    tr = api.Trace(W, H)
    tr.trace_spp(s, cam, dof, 0, 1, sums=True)
    assert parity.path_agreement(tr.image("color"), col) >= 0.97
    assert parity.path_agreement(tr.image("albedo"), alb) >= 0.97
    assert parity.path_agreement(tr.image("normal"), nor) >= 0.97
    tr.close(); ref.close(); s.close()


def test_textured_room_matches_oracle(api, lut, oracle):
    """Every material input of GetSurface at once (scenes.textured_room): varying RGBA8 albedo and ROME
    textures (bilinear + gamma decode, metallic and glossy texels), RG16 normal maps (TanToWorld,
    FixShadingNormal, the shading normal handed from stage A to stage B), a light whose emission texture is a
    checkerboard (EmissionPdf's texel probes decide which triangles are emissive) and a sky-flagged ceiling
    over a non-constant sky cubemap. Scene build and per-path radiance against the reference."""
    desc = scenes.textured_room(brdf_lut=lut, seed=1)
    s = api.Scene(desc)
    oracle.set_seed(1)
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    d = s.download()
    assert ref.info.emissiveCount == 24                      # 12 light triangles + the 12 of the sky ceiling
    assert (d["vertToEmit"] == ref.array("vertToEmit")).all()
    assert (d["emitToVert"] == ref.array("emitToVert")).all()
    assert (d["matIds"] == ref.array("matIds")).all()
    cells = int(np.prod(s.grid_size))
    dists = [ref.dist(c) for c in range(cells)]
    _copy_grid(s, [x[1] for x in dists], [x[0] for x in dists])
    for n in (1000, 20000):
        ro, rd = parity.box_rays(n, seed=9, extent=4.0)
        want = ref.trace_rays(ro, rd, 3)
        got = s.trace_rays(ro, rd, 3)
        assert parity.path_agreement(got[:, 0:3], want[:, 0:3]) >= 0.98
        assert parity.path_agreement(got[:, 3:6], want[:, 3:6]) >= 0.98
        assert parity.path_agreement(got[:, 6:9], want[:, 6:9]) >= 0.98
    # paths that leave through the ceiling pick up the sky: the scene is brighter than with a black sky
    cam, dof, ocam, odof = _cams(api, oracle)
    W, H = 64, 36
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
    ref.trace(col, alb, nor, 1.0, ocam, odof, 0)
    tr = api.Trace(W, H)
    tr.trace_spp(s, cam, dof, 0, 1, sums=True)
    assert parity.path_agreement(tr.image("color"), col) >= 0.98
    assert parity.path_agreement(tr.image("albedo"), alb) >= 0.98
    assert parity.path_agreement(tr.image("normal"), nor) >= 0.98
    tr.close(); ref.close(); s.close()
