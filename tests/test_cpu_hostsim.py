"""CPU suite: the library's device headers (pt_core.h, bvh.h, traverse.h, integrator.h) compiled
for the host and stepped one thread at a time (tests/hostsim), against the oracle. This checks
the kernels' arithmetic and the GPU builder's node encoding without a GPU; nothing of it is
linked into libpimpt.so."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import parity
from pim_b200 import scenes

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostsim", "hostsim.cpp")
LIB = os.path.join(HERE, "hostsim", "libhostsim.so")


class Tex(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("w", C.c_int32), ("h", C.c_int32)]


class Mat(C.Structure):
    _fields_ = [("albedo", C.c_int32), ("rome", C.c_int32), ("normal", C.c_int32), ("flags", C.c_uint32),
                ("mfp", C.c_float * 4), ("ior", C.c_float), ("bump", C.c_float)]


class Cam(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("fwd", C.c_float * 3),
                ("slope", C.c_float * 2), ("w", C.c_int32), ("h", C.c_int32), ("aperture", C.c_float),
                ("focal", C.c_float), ("blades", C.c_int32), ("rot", C.c_float), ("curv", C.c_float)]


@pytest.fixture(scope="module")
def hs():
    deps = [SRC] + [os.path.join(HERE, "..", "pim_b200", "csrc", f) for f in
                    ("pt_core.h", "bvh.h", "traverse.h", "integrator.h")]
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-o", LIB, SRC])
    L = C.CDLL(LIB)
    L.hs_new.restype = C.c_void_p
    return L


def _build(hs, oracle, desc):
    oracle.load_scene_desc(desc)
    ref = oracle.Scene()
    info = ref.info
    keep = []
    texs = (Tex * len(desc.textures))()
    for i, t in enumerate(desc.textures):
        a = np.ascontiguousarray(t.texels, np.uint32); keep.append(a)
        texs[i].texels, texs[i].w, texs[i].h = a.ctypes.data, t.width, t.height
    mats = (Mat * len(desc.entities))()
    for i, e in enumerate(desc.entities):
        mats[i].albedo, mats[i].rome, mats[i].normal, mats[i].flags, mats[i].ior = e.albedo, e.rome, e.normal, e.flags, e.ior
    pos, nrm, uv = ref.array("positions"), ref.array("normals"), ref.array("uvs")
    mi, v2e, e2v = ref.array("matIds"), ref.array("vertToEmit"), ref.array("emitToVert")
    E, cells = info.emissiveCount, int(np.prod(ref.grid_size))
    pdf = np.zeros((cells, E), np.float32); cdf = np.zeros((cells, E + 1), np.float32); ln = np.zeros(cells, np.int32)
    for c in range(cells):
        n, p, cd, _ = ref.dist(c)
        ln[c] = n
        if n:
            pdf[c], cdf[c] = p, cd
    gs = np.array(ref.grid_size, np.int32); glo = np.array(info.gridLo, np.float32); ghi = np.array(info.gridHi, np.float32)
    lut = np.ascontiguousarray(desc.brdf_lut, np.float32)
    sky = np.ascontiguousarray(desc.sky, np.float32) if desc.sky is not None else None
    keep.append(sky)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    hs.hs_new.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                             C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
    h = hs.hs_new(P(pos), P(nrm), P(uv), P(mi), info.vertCount, C.addressof(mats), len(desc.entities), C.addressof(texs),
                  len(desc.textures), P(v2e), P(e2v), E, P(glo), P(ghi), P(gs), info.cellsPerMeter, P(pdf), P(cdf), P(ln),
                  P(lut), lut.shape[1], lut.shape[0], P(sky) if sky is not None else None, sky.shape[1] if sky is not None else 0)
    return h, ref, keep


@pytest.mark.parametrize("variant", ["boxes", "spheres", "textured"])
def test_device_headers_match_reference(hs, oracle, variant):
    if variant == "textured":       # varying albedo / ROME / normal textures, textured emission, sky ceiling
        desc = scenes.textured_room(brdf_lut=parity.reference_lut(), seed=1)
    else:
        desc = scenes.cornell_box(variant, brdf_lut=parity.reference_lut(), seed=1)
    h, ref, keep = _build(hs, oracle, desc)
    # traversal of the GPU node format, built by the GPU builder's own helpers
    n = 50000
    ro, rd = parity.box_rays(n, seed=9)
    want = ref.intersect(ro, rd)
    wuvt = np.zeros((n, 4), np.float32); nor = np.zeros((n, 4), np.float32); tvf = np.zeros((n, 3), np.int32)
    hs.hs_intersect_n.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int64] + [C.c_void_p] * 3
    hs.hs_intersect_n(h, ro.ctypes.data, rd.ctypes.data, 0.0, n, wuvt.ctypes.data, nor.ctypes.data, tvf.ctypes.data)
    assert ((tvf[:, 0] == want["type"]) & (tvf[:, 1] == want["iVert"])).mean() >= 0.9999
    hit = want["type"] != 0
    assert np.abs(wuvt[hit, 3] - want["wuvt"][hit, 3]).max() <= 1e-5 * 20
    occ = np.zeros(n, np.uint8)
    rd2 = rd.copy(); rd2[:, 3] = 3.0
    hs.hs_occluded_n.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int64, C.c_void_p]
    hs.hs_occluded_n(h, ro.ctypes.data, rd2.ctypes.data, 0.0, n, occ.ctypes.data)
    assert (occ.astype(bool) == (want["wuvt"][:, 3] <= 3.0) & hit).mean() >= 0.9999

    # one frame, path for path
    W, H = 48, 27
    cam, dof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    dof.autoFocus = 0
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nr = col.copy()
    ref.trace(col, alb, nr, 1.0, cam, dof, 0)
    q = np.array(cam.rotation, np.float32)
    hc = Cam()
    hc.eye[:] = list(cam.position)[:3]
    hc.right[:] = scenes.quat_mul_dir(q, [1, 0, 0]); hc.up[:] = scenes.quat_mul_dir(q, [0, 1, 0])
    hc.fwd[:] = scenes.quat_mul_dir(q, [0, 0, -1])
    th = np.float32(np.tan(np.float32(np.radians(np.float32(90.0))) * np.float32(0.5)))
    hc.slope[:] = [np.float32(W / H) * th, th]
    hc.w, hc.h = W, H
    hc.aperture, hc.focal, hc.blades, hc.rot, hc.curv = dof.aperture, dof.focalLength, dof.bladeCount, dof.bladeRot, dof.focalPlaneCurvature
    hcol = np.zeros((H, W, 3), np.float32); halb = hcol.copy(); hnr = hcol.copy()
    hs.hs_trace_frame.argtypes = [C.c_void_p, C.POINTER(Cam), C.c_uint32, C.c_uint32] + [C.c_void_p] * 3
    hs.hs_trace_frame(h, C.byref(hc), 0, 1, hcol.ctypes.data, halb.ctypes.data, hnr.ctypes.data)
    assert parity.path_agreement(hcol, col) >= 0.99
    assert parity.path_agreement(halb, alb) >= 0.99
    assert parity.path_agreement(hnr, nr) >= 0.99
    hs.hs_del.argtypes = [C.c_void_p]
    hs.hs_del(h)
    ref.close()


def _frame_agreement(hs, oracle, desc, **dof_fields):
    h, ref, keep = _build(hs, oracle, desc)
    W, H = 48, 27
    cam, dof = oracle.camera_lookat([-4, 0, 4], [0, -1, 0]), oracle.default_dof()
    dof.autoFocus = 0
    for k, v in dof_fields.items():
        setattr(dof, k, v)
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nr = col.copy()
    ref.trace(col, alb, nr, 1.0, cam, dof, 0)
    q = np.array(cam.rotation, np.float32)
    hc = Cam()
    hc.eye[:] = list(cam.position)[:3]
    hc.right[:] = scenes.quat_mul_dir(q, [1, 0, 0]); hc.up[:] = scenes.quat_mul_dir(q, [0, 1, 0])
    hc.fwd[:] = scenes.quat_mul_dir(q, [0, 0, -1])
    th = np.float32(np.tan(np.float32(np.radians(np.float32(90.0))) * np.float32(0.5)))
    hc.slope[:] = [np.float32(W / H) * th, th]
    hc.w, hc.h = W, H
    hc.aperture, hc.focal, hc.blades, hc.rot, hc.curv = dof.aperture, dof.focalLength, dof.bladeCount, dof.bladeRot, dof.focalPlaneCurvature
    hcol = np.zeros((H, W, 3), np.float32); halb = hcol.copy(); hnr = hcol.copy()
    hs.hs_trace_frame.argtypes = [C.c_void_p, C.POINTER(Cam), C.c_uint32, C.c_uint32] + [C.c_void_p] * 3
    hs.hs_trace_frame(h, C.byref(hc), 0, 1, hcol.ctypes.data, halb.ctypes.data, hnr.ctypes.data)
    hs.hs_del.argtypes = [C.c_void_p]
    hs.hs_del(h)
    ref.close()
    return min(parity.path_agreement(hcol, col), parity.path_agreement(halb, alb), parity.path_agreement(hnr, nr))


@pytest.mark.parametrize("dof", [dict(aperture=0.3, focalLength=4.0), dict(aperture=0.3, bladeCount=666),
                                 dict(aperture=0.2, bladeCount=3, bladeRot=1.0),
                                 dict(aperture=0.2, focalPlaneCurvature=0.5), dict(aperture=0.0)],
                         ids=["wide", "pentagram", "triangle-rotated", "curved-focal-plane", "pinhole"])
def test_depth_of_field_variants(hs, oracle, dof):
    """CalculateDof (path_tracer.c:2418-2452): n-gon and pentagram apertures, blade rotation, focal plane curvature"""
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    assert _frame_agreement(hs, oracle, desc, **dof) >= 0.99


def test_missing_textures_and_sky_only_lighting(hs, oracle):
    """materials without albedo / ROME textures take SampleAlbedo / SampleRome's defaults (:1328-1352); a scene
    whose only light is a sky-flagged ceiling"""
    lut = parity.reference_lut()
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    for e in desc.entities[3:]:
        e.albedo = -1
    desc.entities[0].rome = -1
    assert _frame_agreement(hs, oracle, desc) >= 0.99
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    del desc.entities[2]
    desc.sky = np.full((6, 4, 4, 3), 0.5, np.float32)
    desc.entities[1].flags |= scenes.MATFLAG_SKY
    assert _frame_agreement(hs, oracle, desc) >= 0.99
