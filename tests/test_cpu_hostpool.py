"""The resident host threads behind LmPack_BakeDesc's page compaction (pim_b200/csrc/hostpool.h), on the CPU: the
two-pass compaction equals numpy's, shares cover every item once (also for fewer items than threads), and the pool
survives stop / start (PtSys_Shutdown followed by more calls)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "hostsim", "hostpool_test.cpp")
LIB = os.path.join(HERE, "hostsim", "libhostpool_test.so")
HDR = os.path.join(HERE, "..", "pim_b200", "csrc", "hostpool.h")


@pytest.fixture(scope="module")
def hp():
    if not os.path.exists(LIB) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in (SRC, HDR)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-pthread", "-o", LIB, SRC])
    L = C.CDLL(LIB)
    L.hp_compact.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    L.hp_compact.restype = C.c_int32
    L.hp_touch.argtypes = [C.c_void_p, C.c_int64]
    L.hp_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    L.hp_stream_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    return L


@pytest.mark.parametrize("n,density", [(0, 0.5), (1, 1.0), (5, 0.5), (1 << 20, 0.052), (1 << 20, 1.0), (999_983, 0.0)])
def test_compaction_equals_numpy(hp, n, density):
    rng = np.random.default_rng(n + 1)
    counts = np.where(rng.random(n) < density, rng.integers(1, 100, n), 0).astype(np.float32)
    idx = np.full(max(n, 1), -1, np.int32)
    m = hp.hp_compact(counts.ctypes.data, n, idx.ctypes.data)
    want = np.nonzero(counts)[0].astype(np.int32)
    assert m == len(want)
    assert np.array_equal(idx[:m], want)


def test_shares_cover_every_item_once_and_pool_restarts(hp):
    assert 1 <= hp.hp_workers() <= 8
    for n in (0, 1, 3, 7, 8, 9, 1000, 54_437):
        visits = np.zeros(max(n, 1), np.int32)
        for _ in range(3):
            hp.hp_touch(visits.ctypes.data, n)
        assert (visits[:n] == 3).all()
    hp.hp_stop()
    visits = np.zeros(100, np.int32)
    hp.hp_touch(visits.ctypes.data, 100)
    assert (visits == 1).all()


def test_parallel_streaming_copy_is_a_memcpy(hp):
    """hostCopyParallel / streamCopy (the ring's fill and drain): every size class and misalignment copies exactly the
    bytes asked for and nothing beyond them."""
    rng = np.random.default_rng(7)
    for n in (0, 1, 15, 63, 64, 255, 256, 257, 4097, 1 << 20, (1 << 23) + 13):
        for src_off, dst_off in ((0, 0), (1, 0), (0, 5), (3, 9)):
            src = rng.integers(0, 256, n + 32, dtype=np.uint8)
            for fn in (hp.hp_copy, hp.hp_stream_copy):
                dst = np.full(n + 64, 0xAB, np.uint8)
                fn(dst.ctypes.data + 16 + dst_off, src.ctypes.data + src_off, n)
                assert np.array_equal(dst[16 + dst_off:16 + dst_off + n], src[src_off:src_off + n]), (n, src_off, dst_off)
                assert (dst[:16 + dst_off] == 0xAB).all() and (dst[16 + dst_off + n:] == 0xAB).all()


def test_pool_after_fork(hp):
    """A forked child holds the parent's thread handles but not its threads: start() lets go of them and starts anew"""
    assert hp.hp_fork_child_runs() == 0
    visits = np.zeros(100, np.int32)
    hp.hp_touch(visits.ctypes.data, 100)          # the parent's pool still works
    assert (visits == 1).all()
