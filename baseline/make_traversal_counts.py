"""Counts wide-node and triangle fetches per ray on the canonical CPU BVH (oracle/cpu_bvh.c:
binned SAH, 16 bins, leaves <= 4, collapsed 8-wide, front to back) for the bench workloads:
all rays the reference's Pt_Trace shoots for cornell boxes from the pt_test camera (C1 / C2),
for the 1 M-triangle interior from the bench's camera (C3), the hemisphere rays of the C4
lightmap bake, and the C5 random rays against the 10 M-triangle soup (closest hit and any hit).
Writes baseline/traversal_counts.json (the N_node / N_tri of SURVEY.md 8(d)).

  python baseline/make_traversal_counts.py [cornell] [c3] [c4] [c5]      (default: all; c5 takes ~2 min)"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402
import parity  # noqa: E402
from pim_b200 import scenes  # noqa: E402


BVH_NOTE = "binned SAH 16 bins, leaf<=4, 8-wide collapse, front-to-back (oracle/cpu_bvh.c)"
OUT = os.path.join(ROOT, "baseline", "traversal_counts.json")


def c3():
    desc = scenes.interior(1_000_000, seed=7, emissive_fraction=0.001)
    desc.brdf_lut = parity.reference_lut()
    O.set_seed(1)
    O.load_scene_desc(desc)
    s = O.Scene()
    cam, dof = O.camera_lookat([3.0, 1.5, 3.0], [30.0, 1.2, 27.0], 90.0), O.default_dof()
    dof.autoFocus = 0
    W, H, spp = 480, 270, 4
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
    s.trace(col, alb, nor, 1.0, cam, dof, 0)
    O.lib().oracle_counters_reset()
    for p in range(spp):
        s.trace(col, alb, nor, 1.0 / (p + 2), cam, dof, p + 1)
    nodes, tris, rays = O.counters()
    s.close()
    samples = W * H * spp
    return {"interior_1m_camera_paths": {
        "nodes_per_ray": nodes / rays, "tris_per_ray": tris / rays, "rays_per_sample": rays / samples,
        "rays": rays, "samples": samples, "triangles": desc.tri_count, "bvh": BVH_NOTE,
        "camera": "eye (3, 1.5, 3) at (30, 1.2, 27), fovy 90, 480x270"}}


def c4():
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    O.set_seed(1)
    O.load_scene_desc(desc)
    assert O.lm_pack(1024, 16.0, 0.1, 15.0) == 1
    _, _, _, cnt = O.lm_page(0)
    covered = int((cnt > 0).sum())
    s = O.Scene()
    spp = 8
    O.lib().oracle_counters_reset()
    O.lm_bake(s, 1.0, spp, 0)
    nodes, tris, rays = O.counters()
    s.close()
    return {"cornell_lightmap_16tpm_bake_paths": {
        "nodes_per_ray": nodes / rays, "tris_per_ray": tris / rays, "rays_per_sample": rays / (covered * spp),
        "rays": rays, "samples": covered * spp, "covered_texels": covered, "bvh": BVH_NOTE}}


def c5():
    import truth
    desc = scenes.random_soup(10_000_000, seed=11)
    verts = desc.entities[0].positions
    bvh = truth.CpuBvh(verts)
    n = 1 << 20
    ro, rd = scenes.random_rays(n, seed=13)
    prim, nodes, tris = bvh.intersect(ro, rd)
    ro2, rd2 = scenes.random_rays(n, seed=13, occlusion=True)
    occ = bvh.occluded(ro2, rd2)
    onodes, otris = bvh.last_counts
    bvh.close()
    return {"soup_10m_random_rays": {
        "nodes_per_ray": nodes / n, "tris_per_ray": tris / n, "hit_fraction": float((prim >= 0).mean()),
        "anyhit_nodes_per_ray": onodes / n, "anyhit_tris_per_ray": otris / n, "occluded_fraction": float(occ.mean()),
        "rays": n, "triangles": 10_000_000, "bvh": BVH_NOTE,
        "rays_spec": "scenes.random_rays(seed 13): closest hit tFar 8388608; any hit tFar uniform in [1, 50]"}}


def cornell():
    O.set_seed(1)
    O.load_scene_desc(scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1))
    s = O.Scene()
    cam, dof = O.camera_lookat([-4, 0, 4], [0, -1, 0], 90.0), O.default_dof()
    dof.autoFocus = 0
    W, H, spp = 480, 270, 8
    col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
    O.lib().oracle_counters_reset()
    for p in range(spp):
        s.trace(col, alb, nor, 1.0 / (p + 1), cam, dof, p)
    nodes, tris, rays = O.counters()
    samples = W * H * spp
    out = {"cornell_boxes_camera_paths": {
        "nodes_per_ray": nodes / rays, "tris_per_ray": tris / rays, "rays_per_sample": rays / samples,
        "rays": rays, "samples": samples,
        "bvh": "binned SAH 16 bins, leaf<=4, 8-wide collapse, front-to-back (oracle/cpu_bvh.c)"}}
    ro, rd = parity.box_rays(200000, seed=3)
    O.lib().oracle_counters_reset()
    s.intersect(ro, rd)
    nodes, tris, rays = O.counters()
    out["cornell_boxes_random_rays"] = {"nodes_per_ray": nodes / rays, "tris_per_ray": tris / rays}
    s.close()
    return out


def main():
    which = sys.argv[1:] or ["cornell", "c3", "c4", "c5"]
    out = {}
    if os.path.exists(OUT):
        with open(OUT) as f:
            out = json.load(f)
    for name in which:
        out.update({"cornell": cornell, "c3": c3, "c4": c4, "c5": c5}[name]())
    with open(OUT, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
