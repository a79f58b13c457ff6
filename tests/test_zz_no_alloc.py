"""Pt_TraceRayN_NoAlloc: the Pt_TraceRayN variant for hosts that may not allocate while they
trace (frame loops that ban cudaMalloc after start-up, threads holding locks an allocating call
could stall behind, embeddings that budget device memory at load time). Its scratch is a fixed
array inside the library, so the things worth checking are that it gives the existing entry
point's answers, that a second call is not disturbed by what the first left in the array, and
that oversized batches are refused. Collected last (conftest sorts test_zz_* to the end)."""
import numpy as np
import pytest

import parity
from pim_b200 import scenes

pytestmark = pytest.mark.gpu


def test_no_alloc_matches_trace_ray_n_twice():
    import torch
    from pim_b200 import api
    L = api.load()
    scene = api.Scene(scenes.cornell_box("spheres", brdf_lut=parity.reference_lut(), seed=1))
    dev = torch.device("cuda", 0)
    for call, (n, seed, pass_index) in enumerate([(20000, 31, 3), (7001, 32, 4)]):
        ro, rd = parity.box_rays(n, seed=seed, extent=4.0)
        want = scene.trace_rays(ro, rd, pass_index)              # existing entry point (host buffers)
        d_ro, d_rd = torch.from_numpy(ro).to(dev), torch.from_numpy(rd).to(dev)
        out = [torch.full((n, 3), -1.0, dtype=torch.float32, device=dev) for _ in range(3)]
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]          # cudaMemGetInfo: bytes the driver still has to give
        rc = L.Pt_TraceRayN_NoAlloc(scene.h, d_ro.data_ptr(), d_rd.data_ptr(), pass_index, out[0].data_ptr(),
                                    out[1].data_ptr(), out[2].data_ptr(), n, None)
        assert rc == 0, api.last_error()
        if call > 0:
            # the property in the name, on a warm call (the first one may have loaded the kernels' code):
            # the device's free memory is what it was before the call
            assert torch.cuda.mem_get_info(dev)[0] == free_before
        got = np.concatenate([o.cpu().numpy() for o in out], axis=1)
        # same kernels, same per-path seeds: the tolerance of the other path-parity tests, and tighter
        assert parity.path_agreement(got[:, 0:3], want[:, 0:3]) >= 0.999, call
        assert parity.path_agreement(got[:, 3:6], want[:, 3:6]) >= 0.999, call
        assert parity.path_agreement(got[:, 6:9], want[:, 6:9]) >= 0.999, call
        assert np.allclose(got, want, rtol=1e-5, atol=1e-6), call
    # beyond the static scratch: refused, with a message, nothing traced
    n = api.NOALLOC_MAX_PATHS + 1
    big = torch.zeros((n, 4), dtype=torch.float32, device=dev)
    out = torch.zeros((n, 3), dtype=torch.float32, device=dev)
    rc = L.Pt_TraceRayN_NoAlloc(scene.h, big.data_ptr(), big.data_ptr(), 0, out.data_ptr(), None, None, n, None)
    assert rc != 0 and "PT_NOALLOC_MAX_PATHS" in api.last_error()
    scene.close()
