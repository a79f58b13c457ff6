"""fp64 ground truths and the CPU BVH of oracle/liboracle.so as numpy calls (test infrastructure)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")

CPUHIT = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("prim", np.int32), ("ng", np.float32, 3)])
assert CPUHIT.itemsize == 28

_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(ORACLE_SO)
        vp, i64, f = C.c_void_p, C.c_int64, C.c_float
        L.cpubvh_build.restype = vp
        L.cpubvh_build.argtypes = [vp, C.c_int, i64]
        L.cpubvh_free.argtypes = [vp]
        L.cpubvh_intersect_batch.argtypes = [vp, vp, vp, f, i64, C.c_int, vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.cpubvh_occluded_batch.argtypes = [vp, vp, vp, f, i64, C.c_int, vp]
        L.brute_intersect_f64_batch.argtypes = [vp, C.c_int, i64, vp, vp, f, i64, C.c_int, vp, vp]
        L.f64grid_build.restype = vp
        L.f64grid_build.argtypes = [vp, C.c_int, i64, C.c_double]
        L.f64grid_free.argtypes = [vp]
        L.f64grid_intersect_batch.argtypes = [vp, vp, vp, f, i64, C.c_int, vp, vp]
        _lib = L
    return _lib


def threads():
    return min(os.cpu_count() or 1, 64)


def brute(verts, ro, rd, t_near=0.0):
    """all-triangles fp64 Moller-Trumbore; verts float32 [3*T, 4]; returns (hits CPUHIT[n], flags int32[n])"""
    verts = np.ascontiguousarray(verts, np.float32); ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
    hits = np.zeros(len(ro), CPUHIT); flags = np.zeros(len(ro), np.int32)
    lib().brute_intersect_f64_batch(verts.ctypes.data, verts.shape[1], len(verts) // 3, ro.ctypes.data, rd.ctypes.data, t_near,
                                    len(ro), threads(), hits.ctypes.data, flags.ctypes.data)
    return hits, flags


class Grid:
    """uniform-grid fp64 ground truth (oracle/cpu_bvh.c f64grid_*): no BVH involved"""

    def __init__(self, verts, cell):
        self.verts = np.ascontiguousarray(verts, np.float32)
        self.h = lib().f64grid_build(self.verts.ctypes.data, self.verts.shape[1], len(self.verts) // 3, float(cell))

    def intersect(self, ro, rd, t_near=0.0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        hits = np.zeros(len(ro), CPUHIT); flags = np.zeros(len(ro), np.int32)
        lib().f64grid_intersect_batch(self.h, ro.ctypes.data, rd.ctypes.data, t_near, len(ro), threads(), hits.ctypes.data,
                                      flags.ctypes.data)
        return hits, flags

    def close(self):
        if self.h:
            lib().f64grid_free(self.h)
            self.h = None


class CpuBvh:
    """the canonical CPU BVH (binned SAH, 8-wide) the Embree shim runs on"""

    def __init__(self, verts):
        self.verts = np.ascontiguousarray(verts, np.float32)
        self.h = lib().cpubvh_build(self.verts.ctypes.data, self.verts.shape[1], len(self.verts) // 3)

    def intersect(self, ro, rd, t_near=0.0):
        """-> (primitive or -1 per ray, wide-node fetches, triangle fetches)"""
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        prim = np.zeros(len(ro), np.int32)
        a, b = C.c_uint64(), C.c_uint64()
        lib().cpubvh_intersect_batch(self.h, ro.ctypes.data, rd.ctypes.data, t_near, len(ro), threads(), prim.ctypes.data,
                                     C.byref(a), C.byref(b))
        return prim, a.value, b.value

    def occluded(self, ro, rd, t_near=0.0):
        ro = np.ascontiguousarray(ro, np.float32); rd = np.ascontiguousarray(rd, np.float32)
        occ = np.zeros(len(ro), np.uint8)
        lib().cpubvh_occluded_batch(self.h, ro.ctypes.data, rd.ctypes.data, t_near, len(ro), threads(), occ.ctypes.data)
        a, b = C.c_uint64(), C.c_uint64()
        lib().cpubvh_last_batch_counts(C.byref(a), C.byref(b))
        self.last_counts = (a.value, b.value)
        return occ

    def close(self):
        if self.h:
            lib().cpubvh_free(self.h)
            self.h = None
