"""ctypes binding of libpimpt.so — the names and argument meaning of include/pim_pt.h,
which mirrors /root/reference/src/rendering/path_tracer.h:67-103.

Everything computes on the GPU through the C ABI; there is no Python or CPU
implementation behind these wrappers. Loading fails loudly when the library has
not been built (python -m pim_b200.build, or __graft_entry__.build()).
"""
import ctypes as C
import os

import numpy as np

from . import scenes

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PIMPT_LIB", os.path.join(HERE, "libpimpt.so"))   # PIMPT_LIB: tuning builds

f32p = C.POINTER(C.c_float)


class Float4(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


class Int2(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32)]


class PtRayHit(C.Structure):
    _fields_ = [("wuvt", Float4), ("normal", Float4), ("type", C.c_int32), ("iVert", C.c_int32),
                ("flags", C.c_uint32), ("_pad", C.c_uint32)]


class PtDofInfo(C.Structure):
    _fields_ = [("aperture", C.c_float), ("focalLength", C.c_float), ("bladeCount", C.c_int32),
                ("bladeRot", C.c_float), ("focalPlaneCurvature", C.c_float), ("autoFocusSpeed", C.c_float),
                ("autoFocus", C.c_bool)]


class PtTrace(C.Structure):
    _fields_ = [("color", f32p), ("albedo", f32p), ("normal", f32p), ("denoised", f32p),
                ("imageSize", Int2), ("sampleWeight", C.c_float)]


class PtResult(C.Structure):
    _fields_ = [("color", C.c_float * 3), ("albedo", C.c_float * 3), ("normal", C.c_float * 3)]


class PtResults(C.Structure):
    _fields_ = [("colors", C.POINTER(Float4)), ("directions", C.POINTER(Float4))]


class Camera(C.Structure):
    _fields_ = [("position", Float4), ("rotation", Float4), ("zNear", C.c_float), ("zFar", C.c_float),
                ("fovy", C.c_float), ("_pad", C.c_float)]


class PtTextureDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("format", C.c_int32), ("texels", C.c_void_p)]


class PtMaterialDesc(C.Structure):
    _fields_ = [("albedo", C.c_int32), ("rome", C.c_int32), ("normal", C.c_int32), ("flags", C.c_uint32),
                ("meanFreePath", C.c_float * 4), ("ior", C.c_float), ("bumpiness", C.c_float)]


class PtEntityDesc(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("normals", C.c_void_p), ("uvs", C.c_void_p), ("length", C.c_int32),
                ("matrix", C.c_float * 16), ("material", PtMaterialDesc)]


class PtConfig(C.Structure):
    _fields_ = [("ptDistMeters", C.c_float), ("seed", C.c_uint32), ("pathPoolSize", C.c_int32), ("bakeMode", C.c_int32)]


class PtSceneDesc(C.Structure):
    _fields_ = [("entities", C.POINTER(PtEntityDesc)), ("entityCount", C.c_int32),
                ("textures", C.POINTER(PtTextureDesc)), ("textureCount", C.c_int32),
                ("skyFaces", C.c_void_p), ("skySize", C.c_int32),
                ("brdfLut", C.c_void_p), ("lutWidth", C.c_int32), ("lutHeight", C.c_int32),
                ("config", PtConfig)]


class PtLightmapDesc(C.Structure):
    _fields_ = [("probes", C.c_void_p * 5), ("position", C.c_void_p), ("normal", C.c_void_p),
                ("sampleCounts", C.c_void_p), ("size", C.c_int32)]


class PtMediaParams(C.Structure):       # PtMediaDesc's user-facing fields (path_tracer.c:91-111)
    _fields_ = [("constantColor", Float4), ("noiseColor", Float4), ("constantMfp", C.c_float), ("noiseMfp", C.c_float),
                ("absorption", C.c_float), ("noiseOctaves", C.c_int32), ("noiseGain", C.c_float),
                ("noiseLacunarity", C.c_float), ("noiseFreq", C.c_float), ("noiseScale", C.c_float),
                ("noiseHeight", C.c_float), ("phaseDirA", C.c_float), ("phaseDirB", C.c_float), ("phaseBlend", C.c_float)]


class PtEmbedMesh(C.Structure):
    _fields_ = [("positions", C.c_void_p), ("normals", C.c_void_p), ("uvs", C.c_void_p), ("texIndices", C.c_void_p),
                ("length", C.c_int32), ("materialFlags", C.c_uint32)]


class PtSceneStats(C.Structure):
    _fields_ = [("vertCount", C.c_int32), ("matCount", C.c_int32), ("emissiveCount", C.c_int32),
                ("gridSize", C.c_int32 * 3), ("gridLo", C.c_float * 4), ("gridHi", C.c_float * 4),
                ("cellsPerMeter", C.c_float), ("bvhNodeCount", C.c_int32), ("bvhTriCount", C.c_int32),
                ("buildMs", C.c_float), ("phaseMs", C.c_float * 6), ("bvhRefitted", C.c_int32)]


class PtCounters(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("occlusionRays", C.c_uint64),
                ("kernelLaunches", C.c_uint64), ("iterations", C.c_uint64)]


class PtKernelStats(C.Structure):
    _fields_ = [("ms", C.c_double * 5), ("launches", C.c_uint64 * 5), ("extendRays", C.c_uint64),
                ("connectRays", C.c_uint64)]


class Float3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class CubemapStruct(C.Structure):       # Cubemap, src/rendering/cubemap.h:28-34
    _fields_ = [("size", C.c_int32), ("mipCount", C.c_int32), ("color", C.POINTER(C.c_float) * 6),
                ("convolved", C.POINTER(C.c_float) * 6)]


class SkyMedium(C.Structure):           # src/math/types.h:531-542
    _fields_ = [("rCrust", C.c_float), ("rAtmos", C.c_float), ("muR", Float3), ("rhoR", C.c_float),
                ("muM", C.c_float), ("rhoM", C.c_float), ("gM", C.c_float)]


KERNEL_NAMES = ["regen", "extend", "shadeA", "connect", "shadeB"]

HIT_DTYPE = np.dtype([("wuvt", np.float32, 4), ("normal", np.float32, 4), ("type", np.int32),
                      ("iVert", np.int32), ("flags", np.uint32), ("pad", np.uint32)])

EXPORTS = [
    "PtSys_Init", "PtSys_Update", "PtSys_Shutdown", "PtSys_SetSceneDesc", "PtScene_New", "PtScene_NewFromDesc",
    "PtScene_Update", "PtScene_Del", "PtScene_Gui", "PtTrace_New", "PtTrace_Del", "DofInfo_New", "DofInfo_Gui",
    "Pt_Intersect", "Pt_TraceRay", "Pt_Trace", "Pt_RayGen", "Pt_IntersectN", "Pt_OccludedN", "Pt_TraceRayN",
    "Pt_TraceSpp", "Pt_IntersectN_Device", "Pt_OccludedN_Device", "Pt_TraceSpp_Device", "LmPack_BakeDesc",
    "PtScene_GetStats", "PtScene_Download", "PtScene_GetDist", "PtScene_SetDist", "PtScene_GetBrdfLut",
    "Pt_GetCounters", "Pt_ResetCounters", "Pt_LastKernelMs", "Pt_LastError",
    "Pt_SetKernelTiming", "Pt_GetKernelStats", "Pt_IntersectCountN", "Pt_TraceRayN_NoAlloc",
    "Cubemap_New", "Cubemap_Del", "Cubemap_Bake", "Cubemap_Convolve", "Cubemap_BakePass", "Cubemap_ConvolvePass",
    "AmbCube_Bake", "Pt_BakeSky", "Pt_Tonemap", "LmPack_EmbedDesc",
    "LmPack_BakeDesc_Device", "PtMediaParams_Default", "PtScene_SetMedia",
    "Pt_RepackRome", "PtSys_SetLmPack", "LmPack_Bake", "Pt_PointQueryN", "Pt_SetOption", "PtSys_SetDevices", "PtSys_GetDevices", "LmPack_ChartMasksDesc", "LmPack_AtlasDesc", "Pt_MeasureL2ReadGBs", "PtScene_LightCounters",
]

NOALLOC_MAX_PATHS = 32768

_lib = None


def load():
    """dlopen libpimpt.so and declare prototypes. Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("pim_b200/libpimpt.so is not built: run `python -m pim_b200.build` "
                          "(there is no CPU fallback for this library)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.Pt_LastError.restype = C.c_char_p
    L.PtScene_New.restype = vp
    L.PtScene_NewFromDesc.restype = vp
    L.PtScene_NewFromDesc.argtypes = [C.POINTER(PtSceneDesc)]
    L.PtSys_SetSceneDesc.argtypes = [C.POINTER(PtSceneDesc)]
    L.PtScene_Update.argtypes = [vp]
    L.PtScene_Del.argtypes = [vp]
    L.PtTrace_New.argtypes = [C.POINTER(PtTrace), Int2]
    L.PtTrace_Del.argtypes = [C.POINTER(PtTrace)]
    L.DofInfo_New.argtypes = [C.POINTER(PtDofInfo)]
    L.Pt_Intersect.restype = PtRayHit
    L.Pt_Intersect.argtypes = [vp, Float4, Float4, C.c_float, C.c_float]
    L.Pt_TraceRay.restype = PtResult
    L.Pt_TraceRay.argtypes = [vp, Float4, Float4]
    L.Pt_Trace.argtypes = [C.POINTER(PtTrace), C.POINTER(PtDofInfo), vp, C.POINTER(Camera)]
    L.Pt_RayGen.restype = PtResults
    L.Pt_RayGen.argtypes = [vp, Float4, C.c_int32]
    L.Pt_IntersectN.argtypes = [vp, vp, vp, C.c_float, vp, C.c_int64]
    L.Pt_OccludedN.argtypes = [vp, vp, vp, C.c_float, vp, C.c_int64]
    L.Pt_TraceRayN.argtypes = [vp, vp, vp, C.c_uint32, vp, C.c_int64]
    L.Pt_TraceSpp.argtypes = [C.POINTER(PtTrace), C.POINTER(PtDofInfo), vp, C.POINTER(Camera), C.c_int32, C.c_int32]
    L.Pt_IntersectN_Device.argtypes = [vp, vp, vp, C.c_float, vp, C.c_int64, vp]
    L.Pt_OccludedN_Device.argtypes = [vp, vp, vp, C.c_float, vp, C.c_int64, vp]
    L.Pt_TraceSpp_Device.argtypes = [vp, C.POINTER(Camera), C.POINTER(PtDofInfo), C.c_int32, C.c_int32, C.c_int32,
                                     C.c_int32, vp, vp, vp, vp]
    L.LmPack_BakeDesc.argtypes = [vp, C.POINTER(PtLightmapDesc), C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_uint32]
    L.PtScene_GetStats.argtypes = [vp, C.POINTER(PtSceneStats)]
    L.PtScene_Download.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.PtScene_GetDist.argtypes = [vp, C.c_int32, vp, vp, vp]
    L.PtScene_SetDist.argtypes = [vp, C.c_int32, C.c_int32, vp]
    L.PtScene_GetBrdfLut.argtypes = [vp, vp, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.Pt_GetCounters.argtypes = [C.POINTER(PtCounters)]
    L.Pt_LastKernelMs.restype = C.c_float
    L.Pt_TraceRayN_NoAlloc.argtypes = [vp, vp, vp, C.c_uint32, vp, vp, vp, C.c_int64, vp]
    L.Pt_TraceRayN_NoAlloc.restype = C.c_int32
    L.Pt_SetKernelTiming.argtypes = [C.c_int32]
    L.Pt_GetKernelStats.argtypes = [C.POINTER(PtKernelStats)]
    L.Pt_IntersectCountN.argtypes = [vp, vp, vp, C.c_float, C.c_int64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.Pt_IntersectCountN.restype = C.c_int32
    L.Cubemap_New.argtypes = [C.POINTER(CubemapStruct), C.c_int32]
    L.Cubemap_Del.argtypes = [C.POINTER(CubemapStruct)]
    L.Cubemap_Bake.argtypes = [C.POINTER(CubemapStruct), vp, Float4, C.c_float]
    L.Cubemap_Convolve.argtypes = [C.POINTER(CubemapStruct), C.c_uint32, C.c_float]
    L.Cubemap_BakePass.argtypes = [C.POINTER(CubemapStruct), vp, Float4, C.c_float, C.c_uint32]
    L.Cubemap_BakePass.restype = C.c_int32
    L.Cubemap_ConvolvePass.argtypes = [C.POINTER(CubemapStruct), C.c_uint32, C.c_float, C.c_uint32, C.c_uint32]
    L.Cubemap_ConvolvePass.restype = C.c_int32
    L.AmbCube_Bake.argtypes = [vp, vp, Float4, C.c_int32, C.c_int32]
    L.AmbCube_Bake.restype = C.c_int32
    L.Pt_BakeSky.argtypes = [C.POINTER(CubemapStruct), C.POINTER(SkyMedium), Float3, C.c_float, C.c_int32]
    L.Pt_BakeSky.restype = C.c_int32
    L.Pt_Tonemap.argtypes = [vp, C.c_int64, C.c_float, C.c_uint32, C.c_uint32, vp]
    L.Pt_Tonemap.restype = C.c_int32
    L.LmPack_EmbedDesc.argtypes = [C.POINTER(PtEmbedMesh), C.c_int32, C.POINTER(PtLightmapDesc), vp, C.c_int32]
    L.LmPack_EmbedDesc.restype = C.c_int32
    L.LmPack_BakeDesc_Device.argtypes = L.LmPack_BakeDesc.argtypes
    L.Pt_RepackRome.argtypes = [vp, Int2, vp, Int2, vp, Int2, C.c_int32, vp]
    L.Pt_RepackRome.restype = C.c_int32
    L.PtMediaParams_Default.argtypes = [C.POINTER(PtMediaParams)]
    L.PtScene_SetMedia.argtypes = [vp, C.POINTER(PtMediaParams)]
    L.PtScene_SetMedia.restype = C.c_int32
    L.LmPack_BakeDesc_Device.restype = C.c_int32
    L.Pt_PointQueryN.argtypes = [vp, vp, vp, C.c_int64]
    L.Pt_PointQueryN.restype = C.c_int32
    L.LmPack_ChartMasksDesc.argtypes = [vp, vp, C.c_int32, vp, vp, vp]
    L.LmPack_ChartMasksDesc.restype = C.c_int32
    L.LmPack_AtlasDesc.argtypes = [vp, vp, C.c_int32, vp, vp, C.c_int32, vp, vp, C.POINTER(C.c_int32)]
    L.LmPack_AtlasDesc.restype = C.c_int32
    L.PtScene_LightCounters.argtypes = [vp, vp, C.c_int32, vp]
    L.PtScene_LightCounters.restype = C.c_int64
    L.Pt_MeasureL2ReadGBs.argtypes = [C.c_int64, C.c_int32]
    L.Pt_MeasureL2ReadGBs.restype = C.c_float
    L.PtSys_SetDevices.argtypes = [C.c_int32]
    L.PtSys_SetDevices.restype = C.c_int32
    L.PtSys_GetDevices.restype = C.c_int32
    L.Pt_SetOption.argtypes = [C.c_char_p, C.c_int32]
    L.Pt_SetOption.restype = C.c_int32
    L.PtSys_SetLmPack.argtypes = [C.POINTER(PtLightmapDesc), C.c_int32, C.c_float]
    L.PtSys_SetLmPack.restype = None
    L.LmPack_Bake.argtypes = [vp, C.c_float, C.c_int32]
    L.LmPack_Bake.restype = None
    for name in ("Pt_IntersectN", "Pt_OccludedN", "Pt_TraceRayN", "Pt_TraceSpp", "Pt_IntersectN_Device",
                 "Pt_OccludedN_Device", "Pt_TraceSpp_Device", "LmPack_BakeDesc", "PtScene_GetStats",
                 "PtScene_Download", "PtScene_GetDist", "PtScene_SetDist", "PtScene_GetBrdfLut"):
        getattr(L, name).restype = C.c_int32
    _lib = L
    return L


def last_error():
    return (load().Pt_LastError() or b"").decode()


class PtError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise PtError("%s failed: %s" % (what, last_error()))


def _f4(v):
    v = list(v) + [0.0] * (4 - len(v))
    return Float4(float(v[0]), float(v[1]), float(v[2]), float(v[3]))


def make_camera(eye, at, fovy=90.0):
    pos, rot, fov = scenes.camera_lookat(eye, at, fovy)
    cam = Camera()
    cam.position = _f4(pos)
    cam.rotation = _f4(rot)
    cam.zNear, cam.zFar, cam.fovy = 0.1, 500.0, fov
    return cam


def default_dof(auto_focus=True):
    d = PtDofInfo()
    load().DofInfo_New(C.byref(d))
    d.autoFocus = auto_focus
    return d


def set_devices(count):
    """PtSys_SetDevices: scenes built afterwards are replicated on devices 0..count-1 and the frame / bake calls shard"""
    L = load()
    L.PtSys_Init()
    n = L.PtSys_SetDevices(int(count))
    if n < 0:
        raise PtError("PtSys_SetDevices failed: " + last_error())
    return n


def set_option(name, value):
    _check(load().Pt_SetOption(name.encode(), int(value)), "Pt_SetOption")


def counters():
    c = PtCounters()
    load().Pt_GetCounters(C.byref(c))
    return {k: getattr(c, k) for k, _ in PtCounters._fields_}


def kernel_stats():
    k = PtKernelStats()
    load().Pt_GetKernelStats(C.byref(k))
    out = {n: {"ms": k.ms[i], "launches": k.launches[i]} for i, n in enumerate(KERNEL_NAMES)}
    out["extendRays"], out["connectRays"] = k.extendRays, k.connectRays
    return out


def marshal_desc(desc):
    """scenes.SceneDesc -> (PtSceneDesc, objects that keep its arrays alive)"""
    keep = []
    ents = (PtEntityDesc * max(len(desc.entities), 1))()
    for i, e in enumerate(desc.entities):
        pos = np.ascontiguousarray(e.positions, np.float32)
        nrm = np.ascontiguousarray(e.normals, np.float32)
        uvs = np.ascontiguousarray(e.uvs, np.float32)
        # entities that share one mesh object share one upload (keyed by the source array)
        key = id(e.positions)
        cached = next((k for k in keep if isinstance(k, tuple) and k[0] == key), None)
        if cached is None:
            cached = (key, pos, nrm, uvs)
            keep.append(cached)
        _, pos, nrm, uvs = cached
        ents[i].positions = pos.ctypes.data
        ents[i].normals = nrm.ctypes.data
        ents[i].uvs = uvs.ctypes.data
        ents[i].length = len(pos)
        ents[i].matrix[:] = [float(x) for x in np.asarray(e.matrix, np.float32).reshape(-1)]
        m = ents[i].material
        m.albedo, m.rome, m.normal, m.flags = e.albedo, e.rome, e.normal, e.flags
        m.meanFreePath[:] = [float(x) for x in e.mean_free_path]
        m.ior, m.bumpiness = e.ior, e.bumpiness
    texs = (PtTextureDesc * max(len(desc.textures), 1))()
    for i, t in enumerate(desc.textures):
        arr = np.ascontiguousarray(t.texels, np.uint32)
        keep.append(arr)
        texs[i].width, texs[i].height, texs[i].format, texs[i].texels = t.width, t.height, t.format, arr.ctypes.data
    d = PtSceneDesc()
    d.entities = ents
    d.entityCount = len(desc.entities)
    d.textures = texs
    d.textureCount = len(desc.textures)
    if desc.sky is not None:
        sky = np.ascontiguousarray(desc.sky, np.float32)
        keep.append(sky)
        d.skyFaces, d.skySize = sky.ctypes.data, sky.shape[1]
    if desc.brdf_lut is not None:
        lut = np.ascontiguousarray(desc.brdf_lut, np.float32)
        keep.append(lut)
        d.brdfLut, d.lutWidth, d.lutHeight = lut.ctypes.data, lut.shape[1], lut.shape[0]
    d.config.ptDistMeters = desc.pt_dist_meters
    d.config.seed = desc.seed
    d.config.pathPoolSize = desc.path_pool_size
    d.config.bakeMode = desc.bake_mode
    keep += [ents, texs]
    return d, keep


def set_scene_desc(desc):
    """PtSys_SetSceneDesc: register (a copy of) the description PtScene_New(void) builds from; None clears it."""
    L = load()
    L.PtSys_Init()
    if desc is None:
        L.PtSys_SetSceneDesc(None)
        return
    d, keep = marshal_desc(desc)
    L.PtSys_SetSceneDesc(C.byref(d))
    del keep        # the library copied everything


def set_lm_pack(pages, texels_per_meter):
    """PtSys_SetLmPack. pages: list of (probes[5] arrays, position, normal, sample_counts) host arrays that stay alive
    (and are written) while the pack is registered; [] clears it."""
    arr = (PtLightmapDesc * max(len(pages), 1))()
    for i, (probes, position, normal, counts) in enumerate(pages):
        for k in range(5):
            arr[i].probes[k] = probes[k].ctypes.data
        arr[i].position, arr[i].normal, arr[i].sampleCounts = position.ctypes.data, normal.ctypes.data, counts.ctypes.data
        arr[i].size = int(round(np.sqrt(len(counts))))
    load().PtSys_SetLmPack(arr, len(pages), float(texels_per_meter))


class Scene:
    """PtScene built from a scenes.SceneDesc: PtScene_NewFromDesc, or with desc=None the reference's
    PtScene_New(void) over the description registered with set_scene_desc()."""

    def __init__(self, desc=None):
        L = load()
        L.PtSys_Init()
        if desc is None:
            self.h = L.PtScene_New()
            if not self.h:
                raise PtError("PtScene_New failed: " + last_error())
        else:
            d, keep = marshal_desc(desc)
            self.h = L.PtScene_NewFromDesc(C.byref(d))
            del keep
            if not self.h:
                raise PtError("PtScene_NewFromDesc failed: " + last_error())
        self.stats = PtSceneStats()
        _check(L.PtScene_GetStats(self.h, C.byref(self.stats)), "PtScene_GetStats")

    def refresh_stats(self):
        _check(load().PtScene_GetStats(self.h, C.byref(self.stats)), "PtScene_GetStats")
        return self.stats

    def update(self):
        load().PtScene_Update(self.h)
        return self.refresh_stats()

    def bake(self, time_slice, spp):
        """LmPack_Bake over the pack registered with set_lm_pack()."""
        load().LmPack_Bake(self.h, time_slice, spp)
        err = last_error()
        if err:
            raise PtError("LmPack_Bake failed: " + err)

    def close(self):
        if getattr(self, "h", None):
            load().PtScene_Del(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- introspection -------------------------------------------------------
    def download(self):
        n, E = self.stats.vertCount, self.stats.emissiveCount
        out = {"positions": np.zeros((n, 4), np.float32), "normals": np.zeros((n, 4), np.float32),
               "uvs": np.zeros((n, 2), np.float32), "matIds": np.zeros(n, np.int32),
               "vertToEmit": np.zeros(n, np.int32), "emitToVert": np.zeros(E, np.int32)}
        _check(load().PtScene_Download(self.h, *[out[k].ctypes.data for k in
                                                 ("positions", "normals", "uvs", "matIds", "vertToEmit", "emitToVert")]),
               "PtScene_Download")
        return out

    @property
    def grid_size(self):
        return tuple(self.stats.gridSize)

    def dist(self, cell):
        E = self.stats.emissiveCount
        pdf = np.zeros(E, np.float32); cdf = np.zeros(E + 1, np.float32); live = np.zeros(E, np.uint32)
        n = load().PtScene_GetDist(self.h, cell, pdf.ctypes.data, cdf.ctypes.data, live.ctypes.data)
        if n < 0:
            raise PtError("PtScene_GetDist failed: " + last_error())
        return n, pdf, cdf, live

    def set_dist(self, cell, pdf):
        pdf = np.ascontiguousarray(pdf, np.float32)
        _check(load().PtScene_SetDist(self.h, cell, len(pdf), pdf.ctypes.data), "PtScene_SetDist")

    def light_counters(self, d_ptr=None, to_scene=False, stream=None):
        """PtScene_LightCounters: number of counters; with a device pointer also copies them out of / into the scene"""
        n = load().PtScene_LightCounters(self.h, d_ptr, 1 if to_scene else 0, stream)
        if n < 0:
            raise PtError("PtScene_LightCounters failed: " + last_error())
        return int(n)

    def brdf_lut(self):
        w, h = C.c_int32(), C.c_int32()
        L = load()
        _check(L.PtScene_GetBrdfLut(self.h, None, C.byref(w), C.byref(h)), "PtScene_GetBrdfLut")
        out = np.zeros((h.value, w.value, 2), np.float32)
        _check(L.PtScene_GetBrdfLut(self.h, out.ctypes.data, C.byref(w), C.byref(h)), "PtScene_GetBrdfLut")
        return out

    # ---- ray batches -----------------------------------------------------------
    def intersect(self, ro, rd, t_near=0.0, out=None):
        """Pt_IntersectN with host arrays; `out` (HIT_DTYPE, len(ro)) lets a caller reuse its result buffer"""
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        if out is None:
            out = np.zeros(len(ro), HIT_DTYPE)
        assert out.dtype == HIT_DTYPE and len(out) == len(ro) and out.flags.c_contiguous
        _check(load().Pt_IntersectN(self.h, ro.ctypes.data, rd.ctypes.data, t_near, out.ctypes.data, len(ro)), "Pt_IntersectN")
        return out

    def occluded(self, ro, rd, t_near=0.0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        out = np.zeros(len(ro), np.uint8)
        _check(load().Pt_OccludedN(self.h, ro.ctypes.data, rd.ctypes.data, t_near, out.ctypes.data, len(ro)), "Pt_OccludedN")
        return out

    def point_query(self, pts):
        """Pt_PointQueryN: pts[:, 3] = radius -> distance to the nearest triangle within it (8388608 if none)"""
        pts = np.ascontiguousarray(pts, np.float32)
        out = np.zeros(len(pts), np.float32)
        _check(load().Pt_PointQueryN(self.h, pts.ctypes.data, out.ctypes.data, len(pts)), "Pt_PointQueryN")
        return out

    def fetch_counts(self, ro, rd, t_near=0.0):
        """(wide-node fetches, triangle fetches) summed over a closest-hit batch"""
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        a, b = C.c_uint64(), C.c_uint64()
        _check(load().Pt_IntersectCountN(self.h, ro.ctypes.data, rd.ctypes.data, t_near, len(ro), C.byref(a), C.byref(b)),
               "Pt_IntersectCountN")
        return a.value, b.value

    def trace_rays(self, ro, rd, pass_index=0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        out = np.zeros((len(ro), 9), np.float32)
        _check(load().Pt_TraceRayN(self.h, ro.ctypes.data, rd.ctypes.data, pass_index, out.ctypes.data, len(ro)), "Pt_TraceRayN")
        return out

    def raygen(self, origin, count):
        r = load().Pt_RayGen(self.h, _f4(origin), count)
        if not r.colors:
            raise PtError("Pt_RayGen failed: " + last_error())
        colors = np.ctypeslib.as_array(C.cast(r.colors, f32p), (count, 4)).copy()
        dirs = np.ctypeslib.as_array(C.cast(r.directions, f32p), (count, 4)).copy()
        return colors, dirs

    def bake_lightmap(self, probes, position, normal, sample_counts, texels_per_meter, time_slice=1.0, spp=1, pass_index=0):
        """LmPack_BakeDesc on one page. probes: float32 [5, size*size, 4] (updated in place)."""
        size = int(round(np.sqrt(len(sample_counts))))
        d = PtLightmapDesc()
        for k in range(5):
            d.probes[k] = probes[k].ctypes.data
        d.position, d.normal, d.sampleCounts, d.size = position.ctypes.data, normal.ctypes.data, sample_counts.ctypes.data, size
        _check(load().LmPack_BakeDesc(self.h, C.byref(d), 1, texels_per_meter, time_slice, spp, pass_index), "LmPack_BakeDesc")


    def set_media(self, **fields):
        """PtScene_SetMedia: the defaults of media_desc_new with the given PtMediaParams fields replaced."""
        m = PtMediaParams()
        load().PtMediaParams_Default(C.byref(m))
        for k, v in fields.items():
            if k in ("constantColor", "noiseColor"):
                setattr(m, k, Float4(float(v[0]), float(v[1]), float(v[2]), 2.0))
            else:
                setattr(m, k, v)
        _check(load().PtScene_SetMedia(self.h, C.byref(m)), "PtScene_SetMedia")
        return m

    def bake_lightmap_device(self, probes, position, normal, sample_counts, size, texels_per_meter, time_slice=1.0, spp=1,
                             pass_index=0):
        """LmPack_BakeDesc_Device on one page: the arguments are device pointers (ints), probes a list of five."""
        d = PtLightmapDesc()
        for k in range(5):
            d.probes[k] = probes[k]
        d.position, d.normal, d.sampleCounts, d.size = position, normal, sample_counts, size
        _check(load().LmPack_BakeDesc_Device(self.h, C.byref(d), 1, texels_per_meter, time_slice, spp, pass_index),
               "LmPack_BakeDesc_Device")


class Trace:
    """PtTrace: three pinned host float3 images the library writes (PtTrace_New / PtTrace_Del)."""

    def __init__(self, width, height):
        self.t = PtTrace()
        load().PtTrace_New(C.byref(self.t), Int2(width, height))
        if not self.t.color:
            raise PtError("PtTrace_New failed: " + last_error())
        self.width, self.height = width, height

    def image(self, name="color"):
        return np.ctypeslib.as_array(getattr(self.t, name), (self.height, self.width, 3))

    def close(self):
        if self.t.color:
            load().PtTrace_Del(C.byref(self.t))

    def trace(self, scene, cam, dof, sample_weight):
        """One reference-style frame: Pt_Trace (1 spp, lerp by sample_weight)."""
        self.t.sampleWeight = sample_weight
        load().Pt_Trace(C.byref(self.t), C.byref(dof), scene.h, C.byref(cam))
        err = last_error()
        if err:
            raise PtError("Pt_Trace failed: " + err)

    def trace_spp(self, scene, cam, dof, spp_begin, spp_count, sums=False):
        self.t.sampleWeight = -1.0 if sums else 1.0
        _check(load().Pt_TraceSpp(C.byref(self.t), C.byref(dof), scene.h, C.byref(cam), spp_begin, spp_count), "Pt_TraceSpp")


def cubemap_elem_count(size):
    """Texels of one face's mip chain (cubemap.c:104-105: offset of the last, 1x1, mip plus one)."""
    n, m = 0, 0
    while (size >> m) >= 1:
        n += (size >> m) ** 2
        m += 1
    return n


class Cubemap:
    """Cubemap (src/rendering/cubemap.h:28-34): six colour faces and their convolved mip chains in page-locked
    host memory owned by the library (Cubemap_New / Cubemap_Del)."""

    def __init__(self, size):
        self.c = CubemapStruct()
        load().Cubemap_New(C.byref(self.c), size)
        if self.c.size != size:
            raise PtError("Cubemap_New failed: " + last_error())
        self.size = size
        self.elem_count = cubemap_elem_count(size)

    def color(self, face):
        return np.ctypeslib.as_array(self.c.color[face], (self.size, self.size, 3))

    def colors(self):
        return np.stack([self.color(f) for f in range(6)])

    def convolved(self, face):
        return np.ctypeslib.as_array(self.c.convolved[face], (self.elem_count, 4))

    def mip(self, face, m):
        off = sum((self.size >> i) ** 2 for i in range(m))
        s = self.size >> m
        return self.convolved(face)[off:off + s * s].reshape(s, s, 4)

    def close(self):
        if self.c.size:
            load().Cubemap_Del(C.byref(self.c))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bake(self, scene, origin, weight, pass_index=None):
        """Cubemap_Bake: one jittered path per face texel, lerped into the colour faces."""
        if pass_index is None:
            load().Cubemap_Bake(C.byref(self.c), scene.h, _f4(origin), weight)
            err = last_error()
            if err:
                raise PtError("Cubemap_Bake failed: " + err)
        else:
            _check(load().Cubemap_BakePass(C.byref(self.c), scene.h, _f4(origin), weight, pass_index), "Cubemap_BakePass")

    def convolve(self, sample_count, weight, seed=1, pass_index=0):
        _check(load().Cubemap_ConvolvePass(C.byref(self.c), sample_count, weight, seed, pass_index), "Cubemap_ConvolvePass")

    def bake_sky(self, sky, sun_dir, sun_lum, steps):
        _check(load().Pt_BakeSky(C.byref(self.c), C.byref(sky), Float3(*[float(x) for x in sun_dir]), sun_lum, steps),
               "Pt_BakeSky")


def earth_sky():
    """BakeSky's medium from the cvar defaults (src/common/cvars.c sky_*; render_system.c:473-482)."""
    kilo = 1000.0
    mfp = (192.0, 82.0, 34.0)
    return SkyMedium(6360.0 * kilo, 60.0 * kilo, Float3(*[1.0 / (m * kilo) for m in mfp]), 1.0 / (8.5 * kilo),
                     1.0 / (48.0 * kilo), 1.0 / (1.2 * kilo), 0.758)


def ambcube_bake(scene, cube, origin, samples, prev_sample_count):
    """AmbCube_Bake: cube is a float32 array (6, 4) updated in place; returns the new sample count."""
    cube = np.ascontiguousarray(cube, np.float32)
    assert cube.shape == (6, 4)
    aligned = np.zeros(24 + 4, np.float32)          # AmbCube_t holds float4s: 16-byte aligned
    off = (-aligned.ctypes.data % 16) // 4
    a = aligned[off:off + 24]
    a[:] = cube.ravel()
    n = load().AmbCube_Bake(scene.h, a.ctypes.data, _f4(origin), samples, prev_sample_count)
    return n, a.reshape(6, 4).copy()


def tonemap(light, exposure=1.0, seed=1, pass_index=0):
    """Pt_Tonemap: float4 HDR image -> RGBA8 (uint32, R in the low byte)."""
    light = np.ascontiguousarray(light, np.float32).reshape(-1, 4)
    out = np.zeros(light.shape[0], np.uint32)
    _check(load().Pt_Tonemap(light.ctypes.data, light.shape[0], exposure, seed, pass_index, out.ctypes.data), "Pt_Tonemap")
    return out


def lm_embed(meshes, slot_indices, size):
    """LmPack_EmbedDesc. meshes: list of (positions[V,4], normals[V,4], uvs[V,4], texIndices[V,4] int32,
    materialFlags) in drawable order; returns per page (position[S*S,3], normal[S*S,3], sampleCounts[S*S])."""
    keep = []
    arr = (PtEmbedMesh * max(len(meshes), 1))()
    for i, (pos, nrm, uv, tex, flags) in enumerate(meshes):
        pos = np.ascontiguousarray(pos, np.float32); nrm = np.ascontiguousarray(nrm, np.float32)
        uv = np.ascontiguousarray(uv, np.float32); tex = np.ascontiguousarray(tex, np.int32)
        keep += [pos, nrm, uv, tex]
        arr[i] = PtEmbedMesh(pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, tex.ctypes.data, len(pos), int(flags))
    slots = np.ascontiguousarray(slot_indices, np.int32)
    pages = (PtLightmapDesc * max(len(slots), 1))()
    out = []
    for k in range(len(slots)):
        position = np.zeros((size * size, 3), np.float32); normal = np.zeros((size * size, 3), np.float32)
        counts = np.zeros(size * size, np.float32)
        pages[k].position, pages[k].normal, pages[k].sampleCounts, pages[k].size = (
            position.ctypes.data, normal.ctypes.data, counts.ctypes.data, size)
        out.append((position, normal, counts))
    _check(load().LmPack_EmbedDesc(arr, len(meshes), pages, slots.ctypes.data, len(slots)), "LmPack_EmbedDesc")
    return out


def lm_chart_masks(tri_coords, first_node, want_masks=True):
    """LmPack_ChartMasksDesc. tri_coords float32 [nodes, 3, 2] (chart after chart), first_node int32 [charts + 1].
    Returns (moved tri_coords, area [charts], mask_size [charts, 2], masks: list of uint8 [h, w] or None)."""
    tri = np.ascontiguousarray(tri_coords, np.float32).copy()
    first = np.ascontiguousarray(first_node, np.int32)
    n = len(first) - 1
    area = np.zeros(n, np.float32); size = np.zeros((n, 2), np.int32)
    L = load()
    L.PtSys_Init()
    _check(L.LmPack_ChartMasksDesc(tri.ctypes.data, first.ctypes.data, n, area.ctypes.data, size.ctypes.data, None), "LmPack_ChartMasksDesc")
    masks = None
    if want_masks:
        # sizes are known now: a second call (the charts are already at (2, 2), moving them again changes nothing) fills the masks
        masks = [np.zeros((int(size[c, 1]), int(size[c, 0])), np.uint8) for c in range(n)]
        ptrs = (C.c_void_p * max(n, 1))(*[m.ctypes.data for m in masks])
        again = tri.copy()
        _check(L.LmPack_ChartMasksDesc(again.ctypes.data, first.ctypes.data, n, area.ctypes.data, size.ctypes.data, ptrs), "LmPack_ChartMasksDesc")
    return tri, area, size, masks


def lm_atlas(tri_coords, first_node, area, mask_size, atlas_size):
    """LmPack_AtlasDesc: charts in the given order -> (atlas index [charts], translation [charts, 2], atlases in use)"""
    tri = np.ascontiguousarray(tri_coords, np.float32)
    first = np.ascontiguousarray(first_node, np.int32)
    area = np.ascontiguousarray(area, np.float32); size = np.ascontiguousarray(mask_size, np.int32)
    n = len(first) - 1
    idx = np.zeros(n, np.int32); tr = np.zeros((n, 2), np.int32); count = C.c_int32()
    L = load()
    L.PtSys_Init()
    _check(L.LmPack_AtlasDesc(tri.ctypes.data, first.ctypes.data, n, area.ctypes.data, size.ctypes.data, int(atlas_size), idx.ctypes.data,
                              tr.ctypes.data, C.byref(count)), "LmPack_AtlasDesc")
    return idx, tr, count.value


def repack_rome(mr, occ, em, size):
    """Pt_RepackRome: float4 images [h, w, 4] (metallic-roughness, occlusion, emission) -> RGBA8 ROME texture
    [size*size] (uint32, roughness in the low byte)."""
    imgs = [np.ascontiguousarray(t, np.float32) for t in (mr, occ, em)]
    out = np.zeros(size * size, np.uint32)
    args = []
    for t in imgs:
        args += [t.ctypes.data, Int2(t.shape[1], t.shape[0])]
    _check(load().Pt_RepackRome(*args, size, out.ctypes.data), "Pt_RepackRome")
    return out
