"""CPU suite: known-answer vectors of the SURVEY.md 8(a) helper functions (tests/golden/kat.npz, produced by
the reference's own inline functions through oracle/ref_kat.c) against
  * the oracle rebuilt on this machine (pins the oracle to the committed vectors), and
  * the library's device headers compiled for the host (tests/hostsim), i.e. the arithmetic the kernels run.
Integer results (Pcg hashes, interval / cell / face indices) are bit exact; floating point within 1e-6 of
each output's magnitude over the vector set (measured: <= 3e-7) — libm vs the reference's -ffast-math build,
nothing algorithmic. One exception, stated below: the GGX half vector's x, y near the pole."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = np.load(os.path.join(HERE, "golden", "kat.npz"))
NAMES = sorted(k[:-3] for k in GOLD.files if k.endswith("_in"))
EXACT = {"pcg1", "pcg2", "pcg4"}
# columns holding integers (indices) that must match exactly
INT_COLS = {"cubeuv": [0], "grid": [0, 4], "dist1d": [0]}
# SampleGGXMicrofacet: sin(theta) = sqrt(1 - cos^2) with cos ~ 0.996 loses 8 bits to cancellation
TOL = {"ggx": 1e-4}


@pytest.fixture(scope="module")
def hostsim():
    src = os.path.join(HERE, "hostsim", "hostsim.cpp")
    lib = os.path.join(HERE, "hostsim", "libhostsim.so")
    deps = [src] + [os.path.join(HERE, "..", "pim_b200", "csrc", f) for f in ("pt_core.h", "bvh.h", "traverse.h", "integrator.h")]
    if not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in deps):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-o", lib, src])
    L = C.CDLL(lib)
    L.hs_kat_run.argtypes = [C.c_char_p, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float)]
    return L


def _check(name, got, want):
    if name in EXACT:
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), name
        return
    for c in INT_COLS.get(name, []):
        assert np.array_equal(got[:, c], want[:, c]), (name, c)
    assert np.isfinite(want).all() and np.isfinite(got).all(), name
    scale = np.abs(want).max(axis=0) + 1e-6                 # per output column
    err = (np.abs(got - want) / scale).max()
    assert err <= TOL.get(name, 1e-6), (name, float(err))


def test_golden_covers_the_helpers():
    assert len(NAMES) == 24
    for n in NAMES:
        assert GOLD[n + "_in"].shape[0] == GOLD[n + "_out"].shape[0] == 256


@pytest.mark.parametrize("name", NAMES)
def test_device_headers_reproduce_reference_kats(hostsim, name):
    x = np.ascontiguousarray(GOLD[name + "_in"]); want = GOLD[name + "_out"]
    got = np.zeros_like(want)
    kout = hostsim.hs_kat_run(name.encode(), x.ctypes.data_as(C.POINTER(C.c_float)), len(x), got.ctypes.data_as(C.POINTER(C.c_float)))
    assert kout == want.shape[1]
    _check(name, got, want)


def test_oracle_reproduces_its_committed_kats(oracle):
    """the oracle built here gives the vectors that were committed (same reference sources, same flags)"""
    L = oracle.lib()
    L.oracle_kat_name.restype = C.c_char_p
    L.oracle_kat_name.argtypes = [C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.oracle_kat_run.argtypes = [C.c_int, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float)]
    seen = 0
    for k in range(L.oracle_kat_count()):
        kin, kout = C.c_int(), C.c_int()
        name = L.oracle_kat_name(k, C.byref(kin), C.byref(kout)).decode()
        x = np.ascontiguousarray(GOLD[name + "_in"]); want = GOLD[name + "_out"]
        assert x.shape[1] == kin.value and want.shape[1] == kout.value
        got = np.zeros_like(want)
        L.oracle_kat_run(k, x.ctypes.data_as(C.POINTER(C.c_float)), len(x), got.ctypes.data_as(C.POINTER(C.c_float)))
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), name
        seen += 1
    assert seen == len(NAMES)
