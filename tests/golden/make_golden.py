"""Generates the committed fixtures from the reference itself (run in the build container,
where /root/reference is mounted and oracle/_ref is built):

  brdf_lut_128_4096.npy   g_BrdfLut baked by the reference's BakeBrdfFn, 4096 samples/texel
  cornell_boxes_hits.npz  Pt_Intersect on 20000 seeded rays in the cornell boxes scene
  cornell_boxes_64x36.npz two 1-spp passes of Pt_Trace (seed 1) + CalcStdDev
  (kat.npz, the helper known-answer vectors, comes from make_kat.py)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import _oracle as O  # noqa: E402
import parity  # noqa: E402
from pim_b200 import scenes  # noqa: E402


def main():
    O.set_seed(1)
    lut_path = os.path.join(HERE, "brdf_lut_128_4096.npy")
    if not os.path.exists(lut_path):
        np.save(lut_path, O.bake_brdf_lut(128, 4096))
    lut = np.load(lut_path)
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=1)
    O.load_scene_desc(desc)
    s = O.Scene()
    ro, rd = parity.box_rays(20000, seed=3)
    hits = s.intersect(ro, rd)
    np.savez_compressed(os.path.join(HERE, "cornell_boxes_hits.npz"), ro=ro, rd=rd, hits=hits,
                        positions=s.array("positions"), normals=s.array("normals"), uvs=s.array("uvs"),
                        matIds=s.array("matIds"), vertToEmit=s.array("vertToEmit"), emitToVert=s.array("emitToVert"),
                        grid=np.array(s.grid_size))
    cells = int(np.prod(s.grid_size))
    E = s.info.emissiveCount
    pdf = np.zeros((cells, E), np.float32); lens = np.zeros(cells, np.int32)
    for c in range(cells):
        n, p, _, _ = s.dist(c)
        lens[c] = n
        pdf[c] = p
    W, H = 64, 36
    cam, dof = O.camera_lookat([-4, 0, 4], [0, -1, 0]), O.default_dof()
    dof.autoFocus = 0
    frames = []
    for p in range(2):
        col = np.zeros((H, W, 3), np.float32); alb = col.copy(); nor = col.copy()
        s.trace(col, alb, nor, 1.0, cam, dof, p)
        frames.append((col, alb, nor))
    np.savez_compressed(os.path.join(HERE, "cornell_boxes_64x36.npz"), pdf=pdf, lens=lens,
                        color0=frames[0][0], albedo0=frames[0][1], normal0=frames[0][2],
                        color1=frames[1][0], albedo1=frames[1][1], normal1=frames[1][2],
                        stddev0=O.stddev(frames[0][0]))
    s.close()
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
