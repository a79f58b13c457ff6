#!/usr/bin/env python
"""Secondary BASELINE.json configs, measured the same way as bench.py measures configs[1]:

  python tools/configs.py c3 [--spp 256] [--cpu-spp 1]   ~1M-triangle interior, 1080p (configs[2])
  python tools/configs.py c4 [--spp 64] [--cpu]          cornell lightmap bake, 16 texels/m (configs[3])
  python tools/configs.py bakes [--cpu]                  SURVEY 8(f) rows: sky bake, cubemap bake / convolve,
                                                         ambient cube, screenshot tonemap

GPU numbers are CUDA-event / wall times around the C-ABI calls; the CPU numbers are the
reference's own code (oracle/_ref) on the host cores.
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


def c3(args):
    import torch
    from pim_b200 import api, scenes
    L = api.load()
    t0 = time.perf_counter()
    desc = scenes.interior(args.tris, seed=7, emissive_fraction=args.emissive)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    scene = api.Scene(desc)
    build_s = time.perf_counter() - t0
    st = scene.stats
    out = {"config": "c3 interior", "tris": desc.tri_count, "emissive_tris": st.emissiveCount,
           "grid": list(st.gridSize), "bvh_nodes": st.bvhNodeCount, "scene_gen_s": gen_s, "scene_build_s": build_s,
           "build_phase_ms": dict(zip(["flatten", "bvh", "emissive", "grid", "lut", "pool"], [round(x, 1) for x in st.phaseMs]))}
    W, H = 1920, 1080
    eye, at = [3.0, 1.5, 3.0], [30.0, 1.2, 27.0]
    cam, dof = api.make_camera(eye, at, 90.0), api.default_dof(auto_focus=False)
    dev = torch.device("cuda", 0)
    sums = [torch.zeros(W * H * 3, dtype=torch.float32, device=dev) for _ in range(3)]
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    torch.cuda.set_stream(stream)
    step = 16
    L.Pt_TraceSpp_Device(scene.h, C.byref(cam), C.byref(dof), W, H, 0, 4, sums[0].data_ptr(), sums[1].data_ptr(),
                         sums[2].data_ptr(), stream.cuda_stream)
    L.Pt_ResetCounters(); L.Pt_SetKernelTiming(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    done = 4
    while done < 4 + args.spp:
        rc = L.Pt_TraceSpp_Device(scene.h, C.byref(cam), C.byref(dof), W, H, done, step, sums[0].data_ptr(),
                                  sums[1].data_ptr(), sums[2].data_ptr(), stream.cuda_stream)
        assert rc == 0, api.last_error()
        done += step
    e1.record(stream); e1.synchronize()
    ms = e0.elapsed_time(e1)
    k = api.kernel_stats(); c = api.counters()
    rays = k["extendRays"] + k["connectRays"]
    out.update({"spp": args.spp, "msamples_s": W * H * args.spp / ms / 1e3, "mrays_s": rays / ms / 1e3,
                "rays_per_sample": rays / max(c["paths"], 1),
                "kernel_us": {n: round(k[n]["ms"] / max(k[n]["launches"], 1) * 1e3, 1) for n in api.KERNEL_NAMES},
                "mean_luminance": float((sums[0] / (args.spp + 4)).mean().item())})
    if args.cpu_spp > 0:
        import _oracle as O
        if O.available():
            O.set_seed(1)
            desc.brdf_lut = scene.brdf_lut()
            t0 = time.perf_counter()
            O.load_scene_desc(desc)
            ref = O.Scene()
            out["cpu_scene_build_s"] = time.perf_counter() - t0
            ocam, odof = O.camera_lookat(eye, at, 90.0), O.default_dof()
            odof.autoFocus = 0
            w, h = W // 4, H // 4
            col = np.zeros((h, w, 3), np.float32); alb = col.copy(); nor = col.copy()
            ref.trace(col, alb, nor, 1.0, ocam, odof, 0)
            t0 = time.perf_counter()
            for p in range(args.cpu_spp):
                ref.trace(col, alb, nor, 1.0 / (p + 2), ocam, odof, p + 1)
            dt = time.perf_counter() - t0
            out.update({"cpu_msamples_s": w * h * args.cpu_spp / dt / 1e6, "cpu_threads": O.lib().oracle_get_threads(),
                        "cpu_sample": "%d spp at %dx%d" % (args.cpu_spp, w, h), "cpu_mean_luminance": float(col.mean())})
            ref.close()
    print(json.dumps(out))
    scene.close()


def c4(args):
    import _oracle as O
    from pim_b200 import api, scenes
    import parity
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    O.set_seed(1)
    O.load_scene_desc(desc)
    t0 = time.perf_counter()
    pages = O.lm_pack(1024, 16.0)
    pack_s = time.perf_counter() - t0
    out = {"config": "c4 cornell lightmap 16 texels/m", "pages": pages, "lm_pack_cpu_s": pack_s, "spp": args.spp,
           "bake_mode": "sequential" if args.sequential else "wavefront"}
    probes, pos, nrm, cnt = O.lm_page(0)
    covered = int((cnt > 0).sum())
    out["covered_texels"] = covered
    desc.bake_mode = 1 if args.sequential else 0
    scene = api.Scene(desc)
    g_probes = [np.ascontiguousarray(p.copy()) for p in probes]
    g_pos, g_nrm, g_cnt = pos.copy(), nrm.copy(), cnt.copy()
    ref = O.Scene()
    for c in range(int(np.prod(scene.grid_size))):      # same light-grid pdfs on both sides
        n, pdf, _, _ = ref.dist(c)
        if n:
            scene.set_dist(c, pdf)
    warm = api.Scene(desc)     # warm-up on another scene object: a bake also feeds the light-grid learning
    warm.bake_lightmap([p.copy() for p in g_probes], g_pos, g_nrm, g_cnt.copy(), 16.0, 1.0, 1, 99)
    warm.close()
    t0 = time.perf_counter()
    scene.bake_lightmap(g_probes, g_pos, g_nrm, g_cnt, 16.0, 1.0, args.spp, 0)
    dt = time.perf_counter() - t0
    out["gpu_texel_msamples_s"] = covered * args.spp / dt / 1e6
    out["gpu_bake_s"] = dt
    # device-resident pack (LmPack_BakeDesc_Device): no page copies inside the call
    import torch
    d_probes = [torch.from_numpy(p.copy()).cuda() for p in probes]
    d_pos, d_nrm, d_cnt = torch.from_numpy(pos.copy()).cuda(), torch.from_numpy(nrm.copy()).cuda(), torch.from_numpy(cnt.copy()).cuda()
    dev = api.Scene(desc)
    best = 1e30
    for p in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dev.bake_lightmap_device([t.data_ptr() for t in d_probes], d_pos.data_ptr(), d_nrm.data_ptr(), d_cnt.data_ptr(),
                                 1024, 16.0, 1.0, args.spp, p)
        best = min(best, time.perf_counter() - t0)
    out["gpu_device_resident_texel_msamples_s"] = covered * args.spp / best / 1e6
    out["gpu_device_resident_bake_s"] = best
    dev.close()
    if args.cpu:
        t0 = time.perf_counter()
        O.lm_bake(ref, 1.0, args.spp, 0)
        dt = time.perf_counter() - t0
        out["cpu_texel_msamples_s"] = covered * args.spp / dt / 1e6
        out["cpu_bake_s"] = dt
        out["cpu_threads"] = O.lib().oracle_get_threads()
        k = 0
        a, b = g_probes[k][cnt > 0], probes[k][cnt > 0]
        out["probe0_rel_rmse"] = float(np.sqrt(np.mean((a - b) ** 2)) / np.sqrt(np.mean(b ** 2)))
        out["probe0_texels_within_1e-3"] = float((np.abs(a - b).max(axis=1) <= 1e-3 * (np.abs(b).max(axis=1) + 1e-2)).mean())
    ref.close()
    print(json.dumps(out))
    scene.close()


def bakes(args):
    """SURVEY 8(f) N2-N4. GPU: kernel time from CUDA events inside the library (Pt_LastKernelMs) and wall time
    of the whole C-ABI call with host buffers; CPU: the reference's code in oracle/_ref on the host threads."""
    import _oracle as O
    from pim_b200 import api, scenes
    import parity
    L = api.load()
    sky9 = [6360e3, 60e3, 1 / 192e3, 1 / 82e3, 1 / 34e3, 1 / 8.5e3, 1 / 48e3, 1 / 1.2e3, 0.758]
    sun = [0.882, 0.195, 0.429]
    out = {"config": "bakes (SURVEY 8(f) N2-N4)", "cpu_threads": O.lib().oracle_get_threads() if args.cpu else 0}

    def timed(fn, reps=3):
        fn()
        best_wall, best_k = 1e30, 1e30
        for _ in range(reps):
            t0 = time.perf_counter(); fn(); dt = time.perf_counter() - t0
            best_wall = min(best_wall, dt); best_k = min(best_k, L.Pt_LastKernelMs() * 1e-3)
        return best_wall, best_k

    # N3 sky: the reference's default (size 64, 4 steps) and a large one
    for size, steps in ((64, 4), (512, 4)):
        cm = api.Cubemap(size)
        wall, k = timed(lambda: cm.bake_sky(api.earth_sky(), sun, 3800.0, steps))
        r = {"texels": 6 * size * size, "gpu_kernel_ms": k * 1e3, "gpu_call_ms": wall * 1e3, "gpu_mtexels_s": 6 * size * size / k / 1e6}
        if args.cpu and size <= 128:
            t0 = time.perf_counter(); O.bake_sky(size, sky9, sun, 3800.0, steps); dt = time.perf_counter() - t0
            r["cpu_ms"] = dt * 1e3; r["cpu_mtexels_s"] = 6 * size * size / dt / 1e6
        out["sky_%d" % size] = r
        cm.close()

    # N4 tonemap at 1080p: 16 B in + 4 B out per pixel
    n = 1920 * 1080
    light = np.random.default_rng(1).random((n, 4), np.float32) * 4.0
    wall, k = timed(lambda: api.tonemap(light, 1.0, 1, 0))
    out["tonemap_1080p"] = {"gpu_kernel_ms": k * 1e3, "gpu_call_ms": wall * 1e3, "gpu_mpixels_s": n / k / 1e6,
                            "hbm_gbs": n * 20 / k / 1e9}
    if args.cpu:
        t0 = time.perf_counter(); O.tonemap(light, 1.0, 0); dt = time.perf_counter() - t0
        out["tonemap_1080p"].update({"cpu_ms": dt * 1e3, "cpu_mpixels_s": n / dt / 1e6})
    # the same kernel on an image larger than L2 (8192 x 4096): its HBM rate
    nbig = 8192 * 4096
    big = np.tile(light[:1 << 20], (nbig >> 20, 1))
    wall, k = timed(lambda: api.tonemap(big, 1.0, 1, 0), reps=2)
    out["tonemap_8192x4096"] = {"gpu_kernel_ms": k * 1e3, "gpu_call_ms": wall * 1e3, "gpu_mpixels_s": nbig / k / 1e6,
                                "hbm_gbs": nbig * 20 / k / 1e9}
    del big

    # N2 convolve: 64^2 faces, 64 samples per mip texel (reference: render_system.c uses 64)
    size, spt = 64, 64
    color = O.bake_sky(size, sky9, sun, 3800.0, 2) if args.cpu else np.ones((6, size, size, 3), np.float32)
    cm = api.Cubemap(size)
    for f in range(6):
        cm.color(f)[:] = color[f]
    wall, k = timed(lambda: cm.convolve(spt, 0.5, 1, 0))
    texels = 6 * sum((size >> m) ** 2 for m in range(7))
    out["convolve_64"] = {"mip_texels": texels, "samples_per_texel": spt, "gpu_kernel_ms": k * 1e3, "gpu_call_ms": wall * 1e3,
                          "gpu_msamples_s": texels * spt / k / 1e6}
    if args.cpu:
        conv = np.zeros((6, O.cubemap_elem_count(size), 4), np.float32)
        t0 = time.perf_counter(); O.cubemap_convolve(color, conv, spt, 0.5, 0); dt = time.perf_counter() - t0
        out["convolve_64"].update({"cpu_ms": dt * 1e3, "cpu_msamples_s": texels * spt / dt / 1e6})
    cm.close()

    # N1 embed: the C4 atlas (1024^2 at 16 texels/m) the reference's packer lays out for cornell boxes
    if args.cpu:
        desc = scenes.cornell_box("boxes", brdf_lut=np.zeros((4, 4, 2), np.float32), seed=1)
        O.set_seed(1)
        O.load_scene_desc(desc)
        pages = O.lm_pack(1024, 16.0)
        meshes = O.lm_meshes(); slots = [O.lm_slot(k) for k in range(pages)]
        tris = sum(len(m[0]) // 3 for m in meshes)
        wall, k = timed(lambda: api.lm_embed(meshes, slots, 1024))
        t0 = time.perf_counter(); O.lm_embed(); dt = time.perf_counter() - t0
        out["embed_1024"] = {"pages": pages, "triangles": tris, "texels": pages * 1024 * 1024, "gpu_kernel_ms": k * 1e3,
                             "gpu_call_ms": wall * 1e3, "cpu_ms": dt * 1e3}

    # N2 cubemap bake and ambient cube inside cornell boxes
    desc = scenes.cornell_box("boxes", brdf_lut=parity.reference_lut(), seed=1)
    scene = api.Scene(desc)
    for size in (64, 512):
        cm = api.Cubemap(size)
        state = {"p": 0}
        def bake():
            cm.bake(scene, [0.3, -0.2, 0.1, 0.0], 0.5, pass_index=state["p"]); state["p"] += 1
        wall, k = timed(bake)
        out["cubemap_bake_%d" % size] = {"paths": 6 * size * size, "gpu_kernel_ms": k * 1e3, "gpu_call_ms": wall * 1e3,
                                         "gpu_mpaths_s": 6 * size * size / k / 1e6}
        cm.close()
    cube = np.zeros((6, 4), np.float32)
    wall, k = timed(lambda: api.ambcube_bake(scene, cube, [0, 0, 0, 0], 4096, 0))
    out["ambcube_4096"] = {"gpu_call_ms": wall * 1e3}
    if args.cpu:
        O.set_seed(1)
        O.load_scene_desc(desc)
        ref = O.Scene()
        want = np.zeros((6, 64, 64, 3), np.float32)
        t0 = time.perf_counter(); ref.cubemap_bake(want, [0.3, -0.2, 0.1, 0.0], 0.5, 0); dt = time.perf_counter() - t0
        out["cubemap_bake_64"].update({"cpu_ms": dt * 1e3, "cpu_mpaths_s": 6 * 64 * 64 / dt / 1e6})
        t0 = time.perf_counter(); ref.ambcube_bake(cube, [0, 0, 0, 0], 4096, 0, 0); dt = time.perf_counter() - t0
        out["ambcube_4096"]["cpu_ms"] = dt * 1e3
        ref.close()
    scene.close()
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["c3", "c4", "bakes"])
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--emissive", type=float, default=0.001)
    ap.add_argument("--cpu-spp", type=int, default=0)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--sequential", action="store_true", help="c4: PtBake_Sequential")
    args = ap.parse_args()
    if args.config == "bakes":
        bakes(args)
    elif args.config == "c3":
        args.spp = args.spp or 256
        c3(args)
    else:
        args.spp = args.spp or 64
        c4(args)


if __name__ == "__main__":
    main()
