"""Shared helpers of the parity tests (test infrastructure)."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def reference_lut():
    """BRDF LUT baked by the reference's BakeBrdfFn (tests/golden/make_golden.py), 128x128 float2."""
    return np.load(os.path.join(GOLDEN, "brdf_lut_128_4096.npy"))


def box_rays(n, seed=3, extent=4.8, t_far=8388608.0):
    rng = np.random.default_rng(seed)
    ro = np.zeros((n, 4), np.float32)
    ro[:, :3] = rng.uniform(-extent, extent, (n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rd = np.zeros((n, 4), np.float32)
    rd[:, :3] = d
    rd[:, 3] = t_far
    return ro, rd


def path_agreement(a, b, rel=1e-3, floor=1e-3):
    """fraction of pixels (= paths at 1 spp) whose colour agrees within rel of its magnitude"""
    a = np.asarray(a, np.float64).reshape(-1, 3); b = np.asarray(b, np.float64).reshape(-1, 3)
    d = np.abs(a - b).max(axis=1)
    return float(np.mean(d <= rel * (np.abs(b).max(axis=1) + floor)))


def hit_agreement(got, want, exempt=None):
    same = (got["type"] == want["type"]) & (got["iVert"] == want["iVert"])
    if exempt is not None:
        same = same | exempt
    return float(np.mean(same))
