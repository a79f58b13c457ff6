"""Counts wide-node and triangle fetches per ray on the canonical CPU BVH (oracle/cpu_bvh.c:
binned SAH, 16 bins, leaves <= 4, collapsed 8-wide, front to back) for the bench workload:
all rays the reference's Pt_Trace shoots for cornell boxes from the pt_test camera.
Writes baseline/traversal_counts.json (the N_node / N_tri of SURVEY.md 8(d))."""
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


def main():
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
    with open(os.path.join(ROOT, "baseline", "traversal_counts.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
