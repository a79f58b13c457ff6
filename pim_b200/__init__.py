"""pim_b200 — B200-native backend for pim's CPU path-tracing hot path.

The product is pim_b200/libpimpt.so (hand-written sm_100a CUDA behind the C ABI
of include/pim_pt.h). This package is the thin host layer around it: the ctypes
binding (api), the scene descriptions the C ABI takes (scenes) and the
multi-GPU driver (multigpu). There is no CPU implementation: importing
pim_b200.api without the built library raises.
"""
from . import scenes  # noqa: F401

__all__ = ["scenes", "api", "multigpu"]
