"""Generates tests/golden/kat.npz: known-answer vectors of the SURVEY.md 8(a) helper functions, produced by
the reference's own inline functions (oracle/ref_kat.c, run in the build container where /root/reference is
mounted). Inputs are uniform floats from a seeded pcg32 stream (integer KATs read their bit patterns)."""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import _oracle as O  # noqa: E402
from pim_b200.scenes import Pcg32  # noqa: E402

ROWS = 256


def kat_inputs(name, kin, rows=ROWS):
    rng = Pcg32(1000 + sum(name.encode()))
    if name in ("pcg1", "pcg2", "pcg4", "gamma"):
        return np.array([rng.next_u32() for _ in range(rows * kin)], np.uint32).view(np.float32).reshape(rows, kin)
    x = np.array([(rng.next_u32() >> 8) * (1.0 / 16777216.0) for _ in range(rows * kin)], np.float32).reshape(rows, kin)
    if name in ("ngon", "cubeuv"):
        col = 2 if name == "ngon" else 3
        x[:, col] = np.array([rng.next_u32() for _ in range(rows)], np.uint32).view(np.float32)
    if name == "sgfit":
        axii = np.zeros(20, np.float32)
        O.lib().oracle_sg_axii(axii.ctypes.data_as(C.POINTER(C.c_float)))
        x[:, :2] = np.array([rng.next_u32() for _ in range(rows * 2)], np.uint32).view(np.float32).reshape(rows, 2)
        x[:, 2:] = axii
    if name == "dist1d":
        pdf = x[:, 9:17].copy() + 0.05
        pdf[:, 3] = 0.0                                   # an empty bin
        pdf = pdf / pdf.mean(axis=1, keepdims=True)
        cdf = np.concatenate([np.zeros((rows, 1), np.float32), np.cumsum(pdf, axis=1) / 8.0], axis=1).astype(np.float32)
        cdf[:, 8] = 1.0
        x[:, 0:9] = cdf
        x[:, 9:17] = pdf
    return np.ascontiguousarray(x, np.float32)


def oracle_kats():
    L = O.lib()
    L.oracle_kat_name.restype = C.c_char_p
    L.oracle_kat_name.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.oracle_kat_run.argtypes = [C.c_int, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float)]
    out = {}
    for k in range(L.oracle_kat_count()):
        kin, kout = C.c_int(), C.c_int()
        name = L.oracle_kat_name(k, C.byref(kin), C.byref(kout)).decode()
        x = kat_inputs(name, kin.value)
        y = np.zeros((len(x), kout.value), np.float32)
        L.oracle_kat_run(k, x.ctypes.data_as(C.POINTER(C.c_float)), len(x), y.ctypes.data_as(C.POINTER(C.c_float)))
        out[name] = (x, y)
    return out


def main():
    kats = oracle_kats()
    flat = {}
    for name, (x, y) in kats.items():
        flat[name + "_in"] = x
        flat[name + "_out"] = y
    np.savez_compressed(os.path.join(HERE, "kat.npz"), **flat)
    print("wrote", len(kats), "KATs x", ROWS, "rows")


if __name__ == "__main__":
    main()
