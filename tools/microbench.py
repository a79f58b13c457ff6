#!/usr/bin/env python
"""Incoherent-ray intersection microbench (BASELINE.json configs[4]): random rays against a
random triangle soup, closest hit (Pt_IntersectN_Device) and any hit (Pt_OccludedN_Device),
next to the CPU restatement of Embree (oracle/cpu_bvh.c) on the host cores.

  python tools/microbench.py --tris 10000000 --rays 100000000 [--cpu-rays 1000000]
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=10_000_000)
    ap.add_argument("--rays", type=int, default=100_000_000)
    ap.add_argument("--batch", type=int, default=1 << 24)
    ap.add_argument("--cpu-rays", type=int, default=0, help="rays for the CPU baseline and canonical fetch counts")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--extent", type=float, default=100.0)
    args = ap.parse_args()

    import torch
    from pim_b200 import api, scenes
    L = api.load()
    dev = torch.device("cuda", 0)
    t0 = time.perf_counter()
    desc = scenes.random_soup(args.tris, seed=11, extent=args.extent)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    scene = api.Scene(desc)
    build_s = time.perf_counter() - t0
    out = {"tris": args.tris, "rays": args.rays, "scene_gen_s": gen_s, "scene_build_s": build_s,
           "scene_build_ms_device": scene.stats.buildMs, "bvh_nodes": scene.stats.bvhNodeCount, 
           "build_phase_ms": dict(zip(["upload_flatten", "bvh", "emissive", "grid", "lut", "pool"], [round(x, 1) for x in scene.stats.phaseMs]))}

    batch = min(args.batch, args.rays)
    ro, rd = scenes.random_rays(batch, seed=13, extent=args.extent)
    ro_o, rd_o = scenes.random_rays(batch, seed=13, occlusion=True, extent=args.extent)
    d_ro, d_rd = torch.from_numpy(ro).to(dev), torch.from_numpy(rd).to(dev)
    d_rdo = torch.from_numpy(rd_o).to(dev)
    d_hits = torch.empty((batch, 12), dtype=torch.float32, device=dev)
    d_occ = torch.empty(batch, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    nbatches = max(1, args.rays // batch)

    def timed(fn):
        best = None
        with torch.cuda.stream(stream):
            fn()
            stream.synchronize()
            for _ in range(args.reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(nbatches):
                    fn()
                e1.record(stream)
                e1.synchronize()
                ms = e0.elapsed_time(e1)
                best = ms if best is None else min(best, ms)
        return best

    ms_c = timed(lambda: L.Pt_IntersectN_Device(scene.h, d_ro.data_ptr(), d_rd.data_ptr(), 0.0, d_hits.data_ptr(), batch, stream.cuda_stream))
    ms_o = timed(lambda: L.Pt_OccludedN_Device(scene.h, d_ro.data_ptr(), d_rdo.data_ptr(), 0.0, d_occ.data_ptr(), batch, stream.cuda_stream))
    total = nbatches * batch
    nodes, tris = scene.fetch_counts(ro[:1 << 20], rd[:1 << 20])
    n_node, n_tri = nodes / (1 << 20), tris / (1 << 20)
    hits = d_hits.cpu().numpy().view(np.uint32)
    out.update({"closest_mrays_s": total / ms_c / 1e3, "anyhit_mrays_s": total / ms_o / 1e3,
                "hit_fraction": float((hits[:, 8] != 0).mean()), "occluded_fraction": float(d_occ.float().mean().item()),
                "gpu_nodes_per_ray": n_node, "gpu_tris_per_ray": n_tri,
                "gpu_bytes_per_ray": 48 + n_node * 80 + n_tri * 48})
    peak = 6527.5
    peaks = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks):
        peak = float(json.load(open(peaks))["hbm_gbs"])
    out["closest_gbs_own_counts"] = out["closest_mrays_s"] * 1e6 * out["gpu_bytes_per_ray"] / 1e9
    out["closest_frac_hbm_own_counts"] = out["closest_gbs_own_counts"] / peak

    if args.cpu_rays > 0:
        lib = C.CDLL(os.path.join(ROOT, "oracle", "liboracle.so"))
        lib.cpubvh_build.restype = C.c_void_p
        lib.cpubvh_build.argtypes = [C.c_void_p, C.c_int, C.c_int64]
        lib.cpubvh_intersect_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int64, C.c_int,
                                               C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        verts = desc.entities[0].positions
        t0 = time.perf_counter()
        bvh = lib.cpubvh_build(verts.ctypes.data, 4, args.tris)
        out["cpu_bvh_build_s"] = time.perf_counter() - t0
        n = args.cpu_rays
        threads = min(os.cpu_count() or 1, 64)
        prim = np.zeros(n, np.int32)
        a, b = C.c_uint64(), C.c_uint64()
        t0 = time.perf_counter()
        lib.cpubvh_intersect_batch(bvh, ro[:n].ctypes.data, rd[:n].ctypes.data, 0.0, n, threads, prim.ctypes.data, C.byref(a), C.byref(b))
        dt = time.perf_counter() - t0
        gpu_prim = np.where(hits[:n, 8] != 0, hits[:n, 9].view(np.int32) // 3, -1)
        out.update({"cpu_closest_mrays_s": n / dt / 1e6, "cpu_threads": threads, "cpu_kind": "CPU BVH restatement of Embree (binary unavailable)",
                    "canonical_nodes_per_ray": a.value / n, "canonical_tris_per_ray": b.value / n,
                    "agreement_vs_cpu": float((gpu_prim == prim).mean())})
        b_ray = 48 + out["canonical_nodes_per_ray"] * 80 + out["canonical_tris_per_ray"] * 48
        out["bytes_per_ray_canonical"] = b_ray
        out["closest_gbs"] = out["closest_mrays_s"] * 1e6 * b_ray / 1e9
        out["closest_frac_hbm"] = out["closest_gbs"] / peak
    print(json.dumps(out))
    scene.close()


if __name__ == "__main__":
    main()
