"""CPU suite: the oracle against its own golden fixtures and ground truth (no GPU)."""
import ctypes as C
import os

import numpy as np

import parity
from pim_b200 import scenes


def _desc():
    return scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)


def test_reference_reproduces_golden_hits_and_frames(oracle):
    """oracle/_ref rebuilt here gives the committed fixtures bit for bit (deterministic seeding)."""
    oracle.load_scene_desc(_desc())
    s = oracle.Scene()
    g = np.load(os.path.join(parity.GOLDEN, "cornell_boxes_hits.npz"))
    hits = s.intersect(g["ro"], g["rd"])
    assert (hits["iVert"] == g["hits"]["iVert"]).all() and (hits["type"] == g["hits"]["type"]).all()
    assert np.array_equal(hits["wuvt"], g["hits"]["wuvt"])
    assert np.array_equal(s.array("positions"), g["positions"])
    f = np.load(os.path.join(parity.GOLDEN, "cornell_boxes_64x36.npz"))
    cam, dof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    dof.autoFocus = 0
    for p in range(2):
        col = np.zeros((36, 64, 3), np.float32); alb = col.copy(); nor = col.copy()
        s.trace(col, alb, nor, 1.0, cam, dof, p)
        assert parity.path_agreement(col, f["color%d" % p], rel=1e-6) == 1.0
    s.close()


def test_thread_count_does_not_change_results(oracle):
    oracle.load_scene_desc(_desc())
    cam, dof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    dof.autoFocus = 0
    imgs = []
    for threads in (1, 4):
        oracle.set_threads(threads)
        s = oracle.Scene()
        col = np.zeros((18, 32, 3), np.float32); alb = col.copy(); nor = col.copy()
        s.trace(col, alb, nor, 1.0, cam, dof, 0)
        imgs.append(col)
        s.close()
    oracle.set_threads(min(os.cpu_count() or 1, 64))
    assert np.array_equal(imgs[0], imgs[1])


def test_cpu_bvh_against_fp64_brute_force(oracle):
    """The Embree restatement (cpu_bvh.c) is pinned by the all-triangles fp64 loop."""
    oracle.load_scene_desc(scenes.cornell_box("spheres", brdf_lut=parity.reference_lut(), seed=1))
    s = oracle.Scene()
    ro, rd = parity.box_rays(20000, seed=17)
    hits = s.intersect(ro, rd)
    tuv, prim, flags = s.brute(ro, rd)
    s.close()
    ok = flags == 2
    ok |= (flags == 0) & (hits["type"] == 0)
    good = (flags == 1) & (hits["iVert"] == prim * 3)
    err = np.abs(hits["wuvt"][good, 3] - tuv[good, 0]) / np.maximum(tuv[good, 0], 0.05)
    assert (ok | good).mean() >= 0.9999
    assert err.max() <= 4e-5


def test_cornell_box_python_matches_reference_math(oracle):
    """scenes.cornell_box (numpy float32) vs the C restatement on the reference's own headers."""
    L = oracle.lib()
    for variant, flag in (("boxes", 0), ("spheres", 1)):
        d = scenes.cornell_box(variant, brdf_lut=parity.reference_lut())
        L.oracle_cornell_box(flag)
        assert L.oracle_entity_count() == len(d.entities)
        for i, e in enumerate(d.entities):
            mid = C.c_int32(); M = np.zeros(16, np.float32); t3 = np.zeros(3, np.int32)
            fl = C.c_uint32(); mfp = np.zeros(4, np.float32); ior = C.c_float(); b = C.c_float()
            L.oracle_entity_get(i, C.byref(mid), oracle._ptr(M), oracle._ptr(t3, oracle.i32p), C.byref(fl),
                                oracle._ptr(mfp), C.byref(ior), C.byref(b))
            assert np.abs(M - e.matrix).max() < 1e-6 and fl.value == e.flags and abs(ior.value - e.ior) < 1e-7
            for k, ti in enumerate((e.albedo, e.rome)):
                w = C.c_int32(); h = C.c_int32(); p = C.c_void_p(); f = C.c_int32()
                assert L.oracle_texture_get(int(t3[k]), C.byref(w), C.byref(h), C.byref(p), C.byref(f))
                tex = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint32)), (1,))
                assert tex[0] == d.textures[ti].texels[0]
            pp, pn, pu = oracle.f32p(), oracle.f32p(), oracle.f32p()
            n = L.oracle_mesh_get(mid.value, C.byref(pp), C.byref(pn), C.byref(pu))
            assert n == len(e.positions)
            assert np.abs(np.ctypeslib.as_array(pp, (n, 4))[:, :3] - e.positions[:, :3]).max() < 2e-6


def test_stddev_metric(oracle):
    """CalcStdDev restatement (render_system.c:1374-1394): unbiased, over avglum."""
    rng = np.random.default_rng(0)
    img = rng.uniform(0, 2, (9, 16, 3)).astype(np.float32)
    lum = img.mean(axis=2).astype(np.float64)
    assert abs(oracle.stddev(img) - lum.std(ddof=1)) < 1e-5


def test_pcg32_known_answer():
    """O'Neill's pcg32 demo vector (seed 42, stream 54): first outputs of pcg32-demo."""
    r = scenes.Pcg32(42, 54)
    assert [r.next_u32() for _ in range(3)] == [0xa15c02b7, 0x7b47f409, 0xba1d3330]


# ---- the reference's cubemap / sky / tonemap code in oracle/_ref ----------

SKY9 = [6360e3, 60e3, 1 / 192e3, 1 / 82e3, 1 / 34e3, 1 / 8.5e3, 1 / 48e3, 1 / 1.2e3, 0.758]


def test_oracle_sky_bake_is_a_sky(oracle):
    """BakeSkyFn over the reference's Atmosphere(): finite, positive, brightest towards the (low) sun"""
    sun = [0.882, 0.195, 0.429]
    c = oracle.bake_sky(16, SKY9, sun, 3800.0, 2)
    assert np.isfinite(c).all() and (c >= 0).all()
    assert c[0].mean() > 3 * c[1].mean()         # +X (towards the sun) much brighter than -X
    assert c[2].mean() > c[3].mean() > 0         # -Y starts on the crust: one march step, then it ends
    assert c.max() < 3800.0
    again = oracle.bake_sky(16, SKY9, sun, 3800.0, 2)
    assert (again == c).all()                    # no random numbers in this bake


def test_oracle_tonemap_known_answers(oracle):
    """screenshot loop: black stays black (up to the 1/255 dither), huge saturates, monotone in between"""
    x = np.zeros((256, 4), np.float32)
    x[:, :3] = np.linspace(0.0, 4.0, 256, dtype=np.float32)[:, None]
    out = oracle.tonemap(x, 1.0, 0).view(np.uint8).reshape(-1, 4).astype(np.int32)
    assert (out[:, 3] == 255).all()
    assert out[0, :3].max() <= 1
    assert out[-1, :3].min() >= 250
    assert (np.diff(out[:, 0]) >= -1).all()       # monotone up to the dither
    mid = oracle.tonemap(np.array([[0.18, 0.18, 0.18, 1.0]], np.float32), 1.0, 0).view(np.uint8)
    assert abs(int(mid[0]) - 106) <= 2            # sRGB(GT(0.18)): toe 0.5 * (0.18 / 0.5)^1.33 = 0.128 -> 0.415


def test_oracle_cubemap_convolve_of_a_constant_is_the_constant(oracle):
    size = 8
    color = np.full((6, size, size, 3), 0.5, np.float32)
    n = oracle.cubemap_elem_count(size)
    assert n == 64 + 16 + 4 + 1
    conv = np.zeros((6, n, 4), np.float32)
    oracle.cubemap_convolve(color, conv, 8, 1.0, 0)
    assert np.allclose(conv[:, :n - 1, :3], 0.5, atol=1e-5)
    assert np.allclose(conv[:, :n - 1, 3], 1.0, atol=1e-5)
    # the 1x1 mip: Cubemap_CalcDir divides by size - 1 = 0, the direction is NaN, no sample passes NoL > 0
    assert (conv[:, n - 1] == 0).all()


def test_fp64_grid_ground_truth_equals_all_triangles_loop():
    """oracle/cpu_bvh.c f64grid_*: the uniform-grid fp64 ground truth the full-size GPU tests use must give, ray for
    ray, what the all-triangles fp64 loop gives (hit flag incl. ties, primitive, t bit for bit), for long rays and for
    occlusion-length rays; and the CPU BVH's occlusion query must agree with both."""
    import truth
    from pim_b200 import scenes
    desc = scenes.random_soup(100_000, seed=5, extent=20.0)
    verts = desc.entities[0].positions
    grid = truth.Grid(verts, 1.0)
    bvh = truth.CpuBvh(verts)
    for occlusion in (False, True):
        ro, rd = scenes.random_rays(1500, seed=6, extent=20.0, occlusion=occlusion)
        bh, bf = truth.brute(verts, ro, rd)
        gh, gf = grid.intersect(ro, rd)
        assert np.array_equal(bf, gf) and np.array_equal(bh["prim"], gh["prim"]) and np.array_equal(bh["t"], gh["t"])
        assert 0.05 < (bf != 0).mean() < 0.95
        if occlusion:
            assert np.array_equal(bvh.occluded(ro, rd).astype(bool), bf != 0)
        else:
            prim, nodes, tris = bvh.intersect(ro, rd)
            assert ((prim == bh["prim"]) | (bf == 2)).all() and nodes > 0 and tris > 0
    grid.close(); bvh.close()


def test_lm_pack_capture_is_repeatable(oracle):
    """the packer's recorded intermediate state (one task thread, the reference's chunking) is the same on every run:
    what makes it usable as the expected value of the device's chart masks and atlas placement"""
    from pim_b200 import scenes
    desc = scenes.cornell_box("boxes", brdf_lut=np.zeros((4, 4, 2), np.float32), seed=1)
    runs = []
    for _ in range(2):
        oracle.set_seed(1)
        oracle.load_scene_desc(desc)
        charts, atlases = oracle.lm_pack_capture(256, 8.0, 0.1, 15.0)
        runs.append((charts, atlases))
    (a, na), (b, nb) = runs
    assert na == nb and len(a) == len(b) and len(a) >= 20
    for x, y in zip(a, b):
        assert np.array_equal(x["tri_after"], y["tri_after"]) and np.array_equal(x["mask"], y["mask"])
        assert x["atlas"] == y["atlas"] and x["translation"] == y["translation"] and x["sorted_index"] == y["sorted_index"]
    # every chart sits at (2, 2) after ChartMaskFn and got a place
    assert all(abs(c["tri_after"].reshape(-1, 2).min(axis=0) - 2.0).max() < 1e-3 for c in a)
    assert all(c["atlas"] >= 0 for c in a) and max(c["atlas"] for c in a) == 1


def test_reference_chunking_gives_the_same_estimate(oracle):
    """task mode 1 (threading/task.c's chunks, one random stream per thread: the timing arm) renders the same
    estimator as task mode 0 (one seeded stream per pixel: the parity arm): equal mean luminance within noise"""
    from pim_b200 import scenes
    desc = scenes.cornell_box("boxes", brdf_lut=oracle.bake_brdf_lut(128, 256), seed=1)
    cam, dof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    dof.autoFocus = 0
    means = []
    try:
        for mode in (0, 1):
            oracle.set_task_mode(mode)
            oracle.set_seed(1)
            oracle.load_scene_desc(desc)
            ref = oracle.Scene()
            col = np.zeros((90, 160, 3), np.float32); alb = col.copy(); nor = col.copy()
            for p in range(16):
                ref.trace(col, alb, nor, 1.0 / (p + 1), cam, dof, p)
            ref.close()
            means.append(float(np.minimum(col, 4.0).mean()))
    finally:
        oracle.set_task_mode(0)
    assert abs(means[0] - means[1]) <= 0.05 * means[0], means
