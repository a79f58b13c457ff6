"""CPU suite: the C-ABI library loads and exports every function include/pim_pt.h declares."""
import ctypes as C
import os
import re

from pim_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions():
    text = open(os.path.join(ROOT, "include", "pim_pt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"^[A-Za-z_][\w \*]*?\b(\w+)\s*\([^;{]*\)\s*;", text, flags=re.M)
    return sorted(set(n for n in names if n[0].isupper()))


def test_library_exports_every_declared_symbol():
    lib = api.load()
    declared = _declared_functions()
    assert len(declared) >= 35
    missing = [n for n in declared if not hasattr(lib, n)]
    assert not missing, missing
    assert not [n for n in api.EXPORTS if n not in declared]


def test_struct_layouts_match_the_reference_abi():
    assert C.sizeof(api.PtRayHit) == 48          # path_tracer.h:24-31 with float4 alignment
    assert C.sizeof(api.PtResult) == 36          # path_tracer.h:54-59
    assert C.sizeof(api.PtDofInfo) == 28         # path_tracer.h:33-42
    assert C.sizeof(api.PtTrace) == 48
    assert C.sizeof(api.Camera) == 48            # camera.h:8-15
    assert api.HIT_DTYPE.itemsize == 48
    assert C.sizeof(api.CubemapStruct) == 8 + 12 * 8      # cubemap.h:28-34
    assert C.sizeof(api.SkyMedium) == 36                  # types.h:531-542


def test_no_cuda_device_fails_loudly():
    """Without a GPU the library reports an error instead of computing anything on the CPU."""
    import torch
    if torch.cuda.is_available():
        return
    from pim_b200 import scenes
    import numpy as np
    import pytest
    with pytest.raises(api.PtError):
        api.Scene(scenes.cornell_box("boxes", brdf_lut=np.zeros((4, 4, 2), np.float32)))
    assert "CUDA" in api.last_error() or "device" in api.last_error()
