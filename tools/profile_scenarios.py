#!/usr/bin/env python
"""Small workloads for ncu captures of the kernels outside the wavefront (profiles/README.md):

  python tools/profile_scenarios.py build3     C3 scene build: k_sahBinLarge / k_sahPickLarge / k_sahPartitionLarge / k_sahSplit,
                                               k_collapse, k_emissionPdf, k_lightGrid
  python tools/profile_scenarios.py lbvh       1 M-triangle soup through the Morton / radix-sort builder: k_sortScatter, k_radixTree, k_refit
  python tools/profile_scenarios.py bake4      C4 lightmap bake, wavefront (k_bakeRays, k_bakeFit) and sequential (k_lightmapBake)
  python tools/profile_scenarios.py pack4      C4 pack: k_chartMask, k_atlasFind, k_atlasWrite on the reference's charts for cornell at 16 texels/m
  python tools/profile_scenarios.py rays5      C5: k_closestHitPersistent / k_anyHitPersistent on 4 M rays, SAH build of 10 M triangles
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    what = sys.argv[1]
    from pim_b200 import api, scenes
    if what == "build3":
        s = api.Scene(scenes.interior(1_000_000, seed=7, emissive_fraction=0.001))
        print(dict(zip(["upload_flatten", "bvh", "emissive", "grid", "lut", "pool"], s.stats.phaseMs)))
        s.close()
    elif what == "lbvh":
        os.environ["PIMPT_SAH_MAX_TRIS"] = "1000"
        s = api.Scene(scenes.random_soup(1_000_000, seed=11))
        print(s.stats.bvhNodeCount, s.stats.phaseMs[1])
        s.close()
    elif what == "bake4":
        import _oracle as O
        import parity
        desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
        O.set_seed(1)
        O.load_scene_desc(desc)
        O.lm_pack(1024, 16.0, 0.1, 15.0)
        probes, pos, nrm, cnt = O.lm_page(0)
        for mode in (0, 1):
            desc.bake_mode = mode
            s = api.Scene(desc)
            p = [np.ascontiguousarray(x.copy()) for x in probes]
            s.bake_lightmap(p, pos.copy(), nrm.copy(), cnt.copy(), 16.0, 1.0, 64 if mode == 0 else 16, 0)
            s.close()
    elif what == "pack4":
        import _oracle as O
        desc = scenes.cornell_box("boxes", brdf_lut=np.zeros((4, 4, 2), np.float32), seed=1)
        O.set_seed(1)
        O.load_scene_desc(desc)
        charts, _ = O.lm_pack_capture(1024, 16.0, 0.1, 15.0)
        first = np.zeros(len(charts) + 1, np.int32)
        first[1:] = np.cumsum([len(c["tri_before"]) for c in charts])
        tri = np.concatenate([c["tri_before"] for c in charts]).astype(np.float32)
        moved, area, msize, masks = api.lm_chart_masks(tri, first)
        order = np.argsort([c["sorted_index"] for c in charts])
        s_first = np.zeros(len(charts) + 1, np.int32)
        s_first[1:] = np.cumsum([len(charts[i]["tri_after"]) for i in order])
        s_tri = np.concatenate([charts[i]["tri_after"] for i in order]).astype(np.float32)
        print(api.lm_atlas(s_tri, s_first, area[order], msize[order], 1024)[2])
    elif what == "rays5":
        s = api.Scene(scenes.random_soup(10_000_000, seed=11))
        n = 1 << 22
        ro, rd = scenes.random_rays(n, seed=13)
        _, rdo = scenes.random_rays(n, seed=13, occlusion=True)
        s.intersect(ro, rd)
        s.occluded(ro, rdo)
        print(s.stats.bvhNodeCount, s.stats.phaseMs[1])
        s.close()


if __name__ == "__main__":
    main()
