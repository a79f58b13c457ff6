"""ctypes binding of the CPU oracle (test infrastructure only).

oracle/_ref/libpim_ref.so is the reference's own path_tracer.c compiled from
/root/reference over oracle/cpu_bvh.c (see oracle/Makefile). Only tests/,
__graft_entry__.smoke() and bench.py's CPU baseline import this module.
"""
import ctypes as C
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libpim_ref.so")
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")

f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int32)
u32p = C.POINTER(C.c_uint32)


def available():
    return os.path.exists(REF_SO)


class SceneInfo(C.Structure):
    _fields_ = [("vertCount", C.c_int32), ("matCount", C.c_int32), ("emissiveCount", C.c_int32),
                ("gridSize", C.c_int32 * 3), ("gridLo", C.c_float * 4), ("gridHi", C.c_float * 4),
                ("cellsPerMeter", C.c_float), ("rcpMajorant", C.c_float),
                ("positions", f32p), ("normals", f32p), ("uvs", f32p),
                ("matIds", i32p), ("vertToEmit", i32p), ("emitToVert", i32p)]


class Camera(C.Structure):
    _fields_ = [("position", C.c_float * 4), ("rotation", C.c_float * 4),
                ("zNear", C.c_float), ("zFar", C.c_float), ("fovy", C.c_float)]


class Dof(C.Structure):
    _fields_ = [("aperture", C.c_float), ("focalLength", C.c_float), ("bladeCount", C.c_int32),
                ("bladeRot", C.c_float), ("focalPlaneCurvature", C.c_float),
                ("autoFocusSpeed", C.c_float), ("autoFocus", C.c_int32)]


HIT_DTYPE = np.dtype([("wuvt", np.float32, 4), ("normal", np.float32, 4),
                      ("type", np.int32), ("iVert", np.int32), ("flags", np.uint32), ("pad", np.uint32)])
assert HIT_DTYPE.itemsize == 48

_lib = None


def _ptr(a, t=f32p):
    return a.ctypes.data_as(t)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(REF_SO)
        L.oracle_scene_new.restype = C.c_void_p
        L.oracle_scene_del.argtypes = [C.c_void_p]
        L.oracle_scene_info.argtypes = [C.c_void_p, C.POINTER(SceneInfo)]
        L.oracle_scene_dist.argtypes = [C.c_void_p, C.c_int, f32p, f32p, u32p]
        L.oracle_scene_set_dist.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p]
        L.oracle_scene_set_live.argtypes = [C.c_void_p, C.c_int, u32p]
        L.oracle_trace.argtypes = [C.c_void_p, C.c_int, C.c_int, f32p, f32p, f32p, C.c_float,
                                   C.POINTER(Camera), C.POINTER(Dof), C.c_uint32]
        L.oracle_intersect_n.argtypes = [C.c_void_p, f32p, f32p, C.c_float, C.c_int64, C.c_void_p]
        L.oracle_trace_rays.argtypes = [C.c_void_p, f32p, f32p, C.c_int64, C.c_uint32, f32p]
        L.oracle_raygen.argtypes = [C.c_void_p, f32p, C.c_int, C.c_uint32, f32p, f32p]
        L.oracle_brute_n.argtypes = [C.c_void_p, f32p, f32p, C.c_float, C.c_int64, f32p, i32p, i32p]
        L.oracle_camera_lookat.argtypes = [C.POINTER(Camera), f32p, f32p, C.c_float]
        L.oracle_stddev.restype = C.c_float
        L.oracle_stddev.argtypes = [f32p, C.c_int, C.c_int]
        L.oracle_add_texture.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.oracle_add_mesh.argtypes = [f32p, f32p, f32p, C.c_int]
        L.oracle_add_entity.argtypes = [C.c_int, f32p, C.c_int, C.c_int, C.c_int, C.c_uint32, f32p,
                                        C.c_float, C.c_float]
        L.oracle_set_sky.argtypes = [C.c_int, f32p]
        L.oracle_set_brdf_lut.argtypes = [C.c_int, C.c_int, f32p]
        L.oracle_bake_brdf_lut.argtypes = [C.c_int, C.c_int, C.c_int, f32p]
        L.oracle_set_pt_dist_meters.argtypes = [C.c_float]
        L.oracle_counters_get.argtypes = [C.POINTER(C.c_uint64)] * 3
        L.oracle_entity_get.argtypes = [C.c_int, i32p, f32p, i32p, u32p, f32p, f32p, f32p]
        L.oracle_mesh_get.argtypes = [C.c_int, C.POINTER(f32p), C.POINTER(f32p), C.POINTER(f32p)]
        L.oracle_texture_get.argtypes = [C.c_int, i32p, i32p, C.POINTER(C.c_void_p), i32p]
        L.oracle_cubemap_bake.argtypes = [C.c_void_p, C.c_int, f32p, f32p, C.c_float, C.c_uint32]
        L.oracle_cubemap_convolve.argtypes = [C.c_int, f32p, f32p, C.c_uint32, C.c_float, C.c_uint32]
        L.oracle_ambcube_bake.argtypes = [C.c_void_p, f32p, f32p, C.c_int, C.c_int, C.c_uint32]
        L.oracle_bake_sky.argtypes = [C.c_int, f32p, f32p, f32p, C.c_float, C.c_int]
        L.oracle_tonemap.argtypes = [f32p, C.c_int, C.c_float, C.c_uint32, u32p]
        L.oracle_sys_init()
        L.oracle_set_threads(min(os.cpu_count() or 1, 64))
        _lib = L
    return _lib


def set_threads(n):
    lib().oracle_set_threads(int(n))


def set_task_mode(mode):
    """0: one call and one seeded random state per work item (parity runs, the default); 1: the reference's own
    chunking, threading/task.c:35-64 — chunks of worksize / T^2 items per call, one random state per thread (timing)"""
    lib().oracle_set_task_mode(int(mode))


def set_seed(seed):
    lib().oracle_set_seed(int(seed))


def load_scene_desc(desc):
    """Fill the oracle's entity/mesh/texture tables from a pim_b200 SceneDesc."""
    L = lib()
    L.oracle_tables_clear()
    tex_ids = []
    for t in desc.textures:
        tex = np.ascontiguousarray(t.texels)
        tex_ids.append(L.oracle_add_texture(t.width, t.height, tex.ctypes.data, tex.itemsize * (tex.shape[-1] if tex.ndim == 3 else 1), t.vk_format))
    mesh_ids = {}
    for e in desc.entities:
        key = id(e.positions)
        if key not in mesh_ids:
            mesh_ids[key] = L.oracle_add_mesh(_ptr(np.ascontiguousarray(e.positions, np.float32)),
                                              _ptr(np.ascontiguousarray(e.normals, np.float32)),
                                              _ptr(np.ascontiguousarray(e.uvs, np.float32)), len(e.positions))
        mfp = np.asarray(e.mean_free_path, np.float32)
        M = np.ascontiguousarray(e.matrix, np.float32)
        L.oracle_add_entity(mesh_ids[key], _ptr(M),
                            tex_ids[e.albedo] if e.albedo >= 0 else 0,
                            tex_ids[e.rome] if e.rome >= 0 else 0,
                            tex_ids[e.normal] if e.normal >= 0 else 0,
                            e.flags, _ptr(mfp), e.ior, e.bumpiness)
    if desc.sky is not None:
        sky = np.ascontiguousarray(desc.sky, np.float32)
        L.oracle_set_sky(sky.shape[1], _ptr(sky))
    else:
        L.oracle_set_sky(0, None)
    lut = np.ascontiguousarray(desc.brdf_lut, np.float32)
    L.oracle_set_brdf_lut(lut.shape[1], lut.shape[0], _ptr(lut))
    L.oracle_set_pt_dist_meters(desc.pt_dist_meters)


class Scene:
    def __init__(self):
        self.h = lib().oracle_scene_new()
        self.info = SceneInfo()
        lib().oracle_scene_info(self.h, C.byref(self.info))

    def close(self):
        if self.h:
            lib().oracle_scene_del(self.h)
            self.h = None

    def array(self, name):
        i = self.info
        n = i.vertCount
        if name == "positions":
            return np.ctypeslib.as_array(i.positions, (n, 4)).copy()
        if name == "normals":
            return np.ctypeslib.as_array(i.normals, (n, 4)).copy()
        if name == "uvs":
            return np.ctypeslib.as_array(i.uvs, (n, 2)).copy()
        if name == "matIds":
            return np.ctypeslib.as_array(i.matIds, (n,)).copy()
        if name == "vertToEmit":
            return np.ctypeslib.as_array(i.vertToEmit, (n,)).copy()
        if name == "emitToVert":
            return np.ctypeslib.as_array(i.emitToVert, (i.emissiveCount,)).copy() if i.emissiveCount else np.zeros(0, np.int32)
        raise KeyError(name)

    @property
    def grid_size(self):
        return tuple(self.info.gridSize)

    def dist(self, cell):
        E = self.info.emissiveCount
        pdf = np.zeros(E, np.float32); cdf = np.zeros(E + 1, np.float32); live = np.zeros(E, np.uint32)
        n = lib().oracle_scene_dist(self.h, cell, _ptr(pdf), _ptr(cdf), _ptr(live, u32p))
        return n, pdf, cdf, live

    def set_dist(self, cell, pdf):
        pdf = np.ascontiguousarray(pdf, np.float32)
        lib().oracle_scene_set_dist(self.h, cell, len(pdf), _ptr(pdf))

    def trace(self, color, albedo, normal, weight, cam, dof, pass_index):
        h, w, _ = color.shape
        lib().oracle_trace(self.h, w, h, _ptr(color), _ptr(albedo), _ptr(normal), weight,
                           C.byref(cam), C.byref(dof), pass_index)

    def intersect(self, ro, rd, t_near=0.0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        out = np.zeros(len(ro), HIT_DTYPE)
        lib().oracle_intersect_n(self.h, _ptr(ro), _ptr(rd), t_near, len(ro), out.ctypes.data)
        return out

    def brute(self, ro, rd, t_near=0.0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        n = len(ro)
        tuv = np.zeros((n, 3), np.float32); prim = np.zeros(n, np.int32); flags = np.zeros(n, np.int32)
        lib().oracle_brute_n(self.h, _ptr(ro), _ptr(rd), t_near, n, _ptr(tuv), _ptr(prim, i32p), _ptr(flags, i32p))
        return tuv, prim, flags

    def occluded(self, ro, rd, t_near=0.0):
        """the reference's RtcOccluded16 (path_tracer.c:524-553), rd[:, 3] = tFar; 1 = occluded"""
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        out = np.zeros(len(ro), np.uint8)
        L = lib()
        L.oracle_occluded_n.argtypes = [C.c_void_p, f32p, f32p, C.c_float, C.c_int64, C.c_void_p]
        L.oracle_occluded_n(self.h, _ptr(ro), _ptr(rd), t_near, len(ro), out.ctypes.data)
        return out

    def point_query(self, pts):
        """the reference's RtcPointQuery (path_tracer.c:597-616), pts[:, 3] = radius -> (distance, primitive)"""
        pts = np.ascontiguousarray(pts, np.float32)
        dist = np.zeros(len(pts), np.float32); prim = np.zeros(len(pts), np.int32)
        L = lib()
        L.oracle_point_query_n.argtypes = [C.c_void_p, f32p, C.c_int64, f32p, i32p]
        L.oracle_point_query_n(self.h, _ptr(pts), len(pts), _ptr(dist), _ptr(prim, i32p))
        return dist, prim

    def trace_rays(self, ro, rd, pass_index=0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        out = np.zeros((len(ro), 9), np.float32)
        lib().oracle_trace_rays(self.h, _ptr(ro), _ptr(rd), len(ro), pass_index, _ptr(out))
        return out

    def raygen(self, origin, count, pass_index=0):
        o = np.asarray(origin, np.float32)
        colors = np.zeros((count, 4), np.float32); dirs = np.zeros((count, 4), np.float32)
        lib().oracle_raygen(self.h, _ptr(o), count, pass_index, _ptr(colors), _ptr(dirs))
        return colors, dirs


    def set_media(self, m):
        """the same PtMediaParams (pim_b200.api) on the reference scene, then media_desc_update"""
        L = lib()
        L.oracle_scene_set_media.argtypes = [C.c_void_p, f32p, C.c_int, f32p]
        p = np.array([m.constantColor.x, m.constantColor.y, m.constantColor.z, m.noiseColor.x, m.noiseColor.y, m.noiseColor.z,
                      m.constantMfp, m.noiseMfp, m.absorption, m.noiseGain, m.noiseLacunarity, m.noiseFreq, m.noiseScale,
                      m.noiseHeight], np.float32)
        ph = np.array([m.phaseDirA, m.phaseDirB, m.phaseBlend], np.float32)
        L.oracle_scene_set_media(self.h, _ptr(p), int(m.noiseOctaves), _ptr(ph))
        L.oracle_scene_info(self.h, C.byref(self.info))      # rcpMajorant changed

    def cubemap_bake(self, color, origin, weight, pass_index=0):
        """Cubemap_Bake (cubemap.c:162-187) into color[6, size, size, 3], in place."""
        assert color.dtype == np.float32 and color.flags.c_contiguous and color.shape[0] == 6
        o = np.asarray(list(origin) + [0.0] * (4 - len(origin)), np.float32)
        lib().oracle_cubemap_bake(self.h, color.shape[1], _ptr(color), _ptr(o), weight, pass_index)
        return color

    def ambcube_bake(self, cube, origin, samples, prev_sample_count, pass_index=0):
        cube = np.ascontiguousarray(cube, np.float32).copy()
        o = np.asarray(list(origin) + [0.0] * (4 - len(origin)), np.float32)
        n = lib().oracle_ambcube_bake(self.h, _ptr(cube), _ptr(o), samples, prev_sample_count, pass_index)
        return n, cube


def cubemap_elem_count(size):
    return lib().oracle_cubemap_elem_count(size)


def cubemap_convolve(color, convolved, sample_count, weight, pass_index=0):
    """Cubemap_Convolve (cubemap.c:262-303): color[6, size, size, 3] -> convolved[6, elemCount, 4], in place."""
    assert color.dtype == np.float32 and convolved.dtype == np.float32
    lib().oracle_cubemap_convolve(color.shape[1], _ptr(color), _ptr(convolved), sample_count, weight, pass_index)
    return convolved


def bake_sky(size, sky9, sun_dir, sun_lum, steps):
    """BakeSkyFn (render_system.c:403-425) over the reference's Atmosphere(); returns color[6, size, size, 3]."""
    color = np.zeros((6, size, size, 3), np.float32)
    s = np.asarray(sky9, np.float32); d = np.asarray(sun_dir, np.float32)
    lib().oracle_bake_sky(size, _ptr(color), _ptr(s), _ptr(d), sun_lum, steps)
    return color


def tonemap(light, exposure=1.0, pass_index=0):
    light = np.ascontiguousarray(light, np.float32).reshape(-1, 4)
    out = np.zeros(light.shape[0], np.uint32)
    lib().oracle_tonemap(_ptr(light), light.shape[0], exposure, pass_index, _ptr(out, u32p))
    return out


def repack_rome(mr, occ, em, size):
    """CreateRomeTexture's texel loop (gltf_model.c:1118-1139) over float4 images [h, w, 4]"""
    L = lib()
    L.oracle_repack_rome.argtypes = [f32p, C.c_int, C.c_int, f32p, C.c_int, C.c_int, f32p, C.c_int, C.c_int, C.c_int, u32p]
    imgs = [np.ascontiguousarray(t, np.float32) for t in (mr, occ, em)]
    out = np.zeros(size * size, np.uint32)
    args = []
    for t in imgs:
        args += [_ptr(t), t.shape[1], t.shape[0]]
    L.oracle_repack_rome(*args, size, _ptr(out, u32p))
    return out


def default_dof():
    d = Dof()
    lib().oracle_dof_default(C.byref(d))
    return d


def camera_lookat(eye, at, fovy=90.0):
    cam = Camera()
    e = np.asarray(eye, np.float32); a = np.asarray(at, np.float32)
    lib().oracle_camera_lookat(C.byref(cam), _ptr(e), _ptr(a), fovy)
    return cam


def counters():
    a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
    lib().oracle_counters_get(C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def bake_brdf_lut(size=128, samples=4096):
    out = np.zeros((size, size, 2), np.float32)
    lib().oracle_bake_brdf_lut(size, size, samples, _ptr(out))
    return out


def stddev(color):
    h, w, _ = color.shape
    c = np.ascontiguousarray(color, np.float32)
    return float(lib().oracle_stddev(_ptr(c), w, h))


# ---- lightmaps ---------------------------------------------------------------------------

def lm_pack(atlas_size, texels_per_unit, dist_thresh=0.1, deg_thresh=15.0):
    """The reference's LmPack_Pack on the loaded entity tables; returns the page count."""
    L = lib()
    L.oracle_lm_pack.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float]
    return L.oracle_lm_pack(atlas_size, texels_per_unit, dist_thresh, deg_thresh)


def lm_page(i):
    """numpy views (no copy) of page i: probes [5][S*S,4], position [S*S,3], normal [S*S,3], counts [S*S]"""
    L = lib()
    L.oracle_lm_page.argtypes = [C.c_int, C.POINTER(f32p * 5), C.POINTER(f32p), C.POINTER(f32p), C.POINTER(f32p)]
    probes = (f32p * 5)(); pos = f32p(); nrm = f32p(); cnt = f32p()
    L.oracle_lm_page(i, C.byref(probes), C.byref(pos), C.byref(nrm), C.byref(cnt))
    n = L.oracle_lm_size() ** 2
    return ([np.ctypeslib.as_array(probes[k], (n, 4)) for k in range(5)], np.ctypeslib.as_array(pos, (n, 3)),
            np.ctypeslib.as_array(nrm, (n, 3)), np.ctypeslib.as_array(cnt, (n,)))


def lm_embed():
    lib().oracle_lm_embed()


def lm_slot(i):
    return lib().oracle_lm_slot(i)


def lm_meshes():
    """The drawables as EmbedTaskFn reads them after LmPack_Pack, in Entities order:
    (positions[V,4], normals[V,4], uvs[V,4] (zw = lightmap uv), texIndices[V,4] int32, material flags)."""
    L = lib()
    L.oracle_mesh_texindices.restype = i32p
    L.oracle_mesh_texindices.argtypes = [C.c_int]
    out = []
    for i in range(L.oracle_entity_count()):
        mesh_id = np.zeros(1, np.int32); mat = np.zeros(16, np.float32); tex = np.zeros(3, np.int32)
        flags = np.zeros(1, np.uint32); mfp = np.zeros(4, np.float32); ior = np.zeros(1, np.float32); bump = np.zeros(1, np.float32)
        L.oracle_entity_get(i, _ptr(mesh_id, i32p), _ptr(mat), _ptr(tex, i32p), _ptr(flags, u32p), _ptr(mfp), _ptr(ior), _ptr(bump))
        pos = f32p(); nrm = f32p(); uv = f32p()
        n = L.oracle_mesh_get(int(mesh_id[0]), C.byref(pos), C.byref(nrm), C.byref(uv))
        ti = L.oracle_mesh_texindices(int(mesh_id[0]))
        out.append((np.ctypeslib.as_array(pos, (n, 4)).copy(), np.ctypeslib.as_array(nrm, (n, 4)).copy(),
                    np.ctypeslib.as_array(uv, (n, 4)).copy(), np.ctypeslib.as_array(ti, (n, 4)).copy(), int(flags[0])))
    return out


def lm_bake(scene, time_slice, spp, pass_index):
    L = lib()
    L.oracle_lm_bake.argtypes = [C.c_void_p, C.c_float, C.c_int, C.c_uint32]
    L.oracle_lm_bake(scene.h, time_slice, spp, pass_index)


def lm_pack_capture(atlas_size, texels_per_unit, dist_thresh=0.1, deg_thresh=15.0):
    """LmPack_Pack on one task thread with ChartMaskFn's and AtlasFn's inputs and outputs recorded. Returns a list of
    charts in chart_group's order: dict(tri_before [n,3,2], tri_after, area, mask [h,w] uint8, atlas, translation (x, y),
    sorted_index = the chart's position in AtlasFn's order) and the number of atlases the reference estimated."""
    L = lib()
    L.oracle_lm_pack_capture.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float]
    L.oracle_lm_cap_chart.argtypes = [C.c_int, i32p, f32p, i32p, i32p, i32p]
    L.oracle_lm_cap_chart_data.argtypes = [C.c_int, f32p, f32p, C.c_void_p]
    n = L.oracle_lm_pack_capture(atlas_size, texels_per_unit, dist_thresh, deg_thresh)
    charts = []
    for c in range(n):
        nc = np.zeros(1, np.int32); area = np.zeros(1, np.float32); ms = np.zeros(2, np.int32)
        ai = np.zeros(1, np.int32); tr = np.zeros(2, np.int32)
        L.oracle_lm_cap_chart(c, _ptr(nc, i32p), _ptr(area), _ptr(ms, i32p), _ptr(ai, i32p), _ptr(tr, i32p))
        before = np.zeros((int(nc[0]), 3, 2), np.float32); after = np.zeros_like(before)
        mask = np.zeros((int(ms[1]), int(ms[0])), np.uint8)
        L.oracle_lm_cap_chart_data(c, _ptr(before), _ptr(after), mask.ctypes.data)
        charts.append({"tri_before": before, "tri_after": after, "area": float(area[0]), "mask": mask, "atlas": int(ai[0]),
                       "translation": (int(tr[0]), int(tr[1])), "sorted_index": L.oracle_lm_cap_sorted_index(c)})
    return charts, L.oracle_lm_cap_atlas_count()
