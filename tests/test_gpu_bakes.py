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
