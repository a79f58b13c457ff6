#!/usr/bin/env python
"""bench.py — Msamples/s of the path-tracing hot path on cornell_box at 1920x1080.

A step is one progressive pass of the hot path: `--spp-per-step` samples of every
pixel of the 1080p frame (BASELINE.json configs[1]; the full config is 1024 spp =
the default 16 steps of 64 spp). `value` times Pt_TraceSpp_Device with the scene and the
accumulation images resident in HBM; `e2e` times Pt_TraceSpp, the reference-facing
call with host PtTrace buffers (device->host copies of the three images inside).
With N > 1 every rank renders its own sample indices of the full frame (scene
replicated) and the per-rank sums are reduced to rank 0 with one NCCL reduce per step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, HEIGHT = 1920, 1080
EYE, AT, FOVY = [-4.0, 0.0, 4.0], [0.0, -1.0, 0.0], 90.0    # pt_test camera (render_system.c:1360-1361)
METRIC, UNIT = "Msamples/s (1080p cornell_box)", "Msamples/s"
WORKLOAD = "cornell_box boxes 1920x1080 progressive accumulation (BASELINE configs[1])"
TRAVERSAL_COUNTS = os.path.join(ROOT, "baseline", "traversal_counts.json")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=64)
    ap.add_argument("--pool", type=int, default=0, help="path pool slots (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traversal_counts():
    """mean wide-node / triangle fetches per ray and rays per sample of the canonical CPU
    BVH on this workload; written by baseline/make_traversal_counts.py"""
    with open(TRAVERSAL_COUNTS) as f:
        return json.load(f)["cornell_boxes_camera_paths"]


class ClockSampler:
    """nvidia-smi clocks line of B200_PROFILING.md, sampled while the timed region runs."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
              "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.tmp = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(device_index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=self.tmp, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        self.tmp.flush()
        self.tmp.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.tmp.read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        self.tmp.close()
        os.unlink(self.tmp.name)
        if sm:
            out["sm_mhz"] = float(np.median(sm))
            out["sm_max_mhz"] = float(max(mx))
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def cpu_baseline(width, height, min_seconds=10.0, max_frames=64):
    """The reference's own Pt_Trace (oracle/_ref) on the host cores, same scene and camera."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    from pim_b200 import scenes
    if not O.available():
        return None
    cores = min(os.cpu_count() or 1, 64)
    O.set_threads(cores)
    O.set_seed(1)
    lut = O.bake_brdf_lut(128, 1024)
    O.load_scene_desc(scenes.cornell_box("boxes", brdf_lut=lut, seed=1))
    ref = O.Scene()
    cam, dof = O.camera_lookat(EYE, AT, FOVY), O.default_dof()
    dof.autoFocus = 0
    col = np.zeros((height, width, 3), np.float32); alb = col.copy(); nor = col.copy()
    ref.trace(col, alb, nor, 1.0, cam, dof, 0)          # warm-up frame (page faults, caches)
    L = O.lib()
    L.oracle_counters_reset()
    frames, t0 = 0, time.perf_counter()
    while True:
        ref.trace(col, alb, nor, 1.0 / (frames + 2), cam, dof, frames + 1)
        frames += 1
        dt = time.perf_counter() - t0
        if dt >= min_seconds or frames >= max_frames:
            break
    nodes, tris, rays = O.counters()
    ref.close()
    samples = frames * width * height
    return {"value": samples / dt / 1e6, "unit": UNIT, "cores": cores,
            "kind": "reference",
            "sample": "%d frames of 1 spp at %dx%d, reference path_tracer.c over the CPU BVH restatement "
                      "(Embree 3.13.5 binary unavailable)" % (frames, width, height),
            "seconds": dt, "mrays_per_s": rays / dt / 1e6, "rays_per_sample": rays / samples}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores.
    A step is 1 spp of a 480x270 frame (1/16 of the 1080p pixels, same scene and camera),
    so K steps + W warm-ups end within minutes whatever K is."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    from pim_b200 import scenes
    if not O.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpim_ref.so not built"}))
        return
    w, h = WIDTH // 4, HEIGHT // 4
    cores = min(os.cpu_count() or 1, 64)
    O.set_threads(cores)
    O.set_seed(1)
    lut = O.bake_brdf_lut(128, 1024)
    O.load_scene_desc(scenes.cornell_box("boxes", brdf_lut=lut, seed=1))
    ref = O.Scene()
    cam, dof = O.camera_lookat(EYE, AT, FOVY), O.default_dof()
    dof.autoFocus = 0
    col = np.zeros((h, w, 3), np.float32); alb = col.copy(); nor = col.copy()
    for i in range(args.warmup):
        ref.trace(col, alb, nor, 1.0 / (i + 1), cam, dof, i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        ref.trace(col, alb, nor, 1.0 / (args.warmup + i + 1), cam, dof, args.warmup + i)
    dt = time.perf_counter() - t0
    ref.close()
    value = args.steps * w * h / dt / 1e6
    sample = "per step: 1 spp of %dx%d (1/16 of the 1080p frame), same scene and camera" % (w, h)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample, "threads": cores},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args, rank, world, local_rank):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from pim_b200 import api, scenes, multigpu

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = api.load()
    spp = args.spp_per_step
    desc = scenes.cornell_box("boxes", brdf_lut=None, seed=1 + rank)   # LUT baked on the device
    desc.path_pool_size = args.pool
    scene = api.Scene(desc)
    cam = api.make_camera(EYE, AT, FOVY)
    dof = api.default_dof(auto_focus=False)
    n3 = WIDTH * HEIGHT * 3
    sums = [torch.zeros(n3, dtype=torch.float32, device=dev) for _ in range(3)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)     # > 126 MB L2
    stream = torch.cuda.Stream(device=dev)      # the library launches on this stream; events are recorded on it
    torch.cuda.synchronize()
    torch.cuda.set_stream(stream)

    def step(i):
        # every rank renders its own sample indices of the whole frame
        begin, _ = multigpu.sample_range(i, rank, world, spp)
        rc = L.Pt_TraceSpp_Device(scene.h, C.byref(cam), C.byref(dof), WIDTH, HEIGHT, begin, spp,
                                  sums[0].data_ptr(), sums[1].data_ptr(), sums[2].data_ptr(), stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("Pt_TraceSpp_Device: " + api.last_error())
        multigpu.reduce_sums(sums, dst=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    barrier()
    L.Pt_ResetCounters()
    L.Pt_SetKernelTiming(1)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    total_ms = 0.0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(args.steps):
        flush.fill_(i & 0xFF)                     # L2 flush between timed iterations, outside the events
        e0.record(stream)
        step(args.warmup + i)
        e1.record(stream)
        e1.synchronize()
        total_ms += e0.elapsed_time(e1)
    barrier()
    clocks = sampler.stop() if sampler else None
    counters = api.counters()
    kstats = api.kernel_stats()
    L.Pt_SetKernelTiming(0)
    total_ms = multigpu.max_over_ranks(total_ms, dev)
    samples_per_step = WIDTH * HEIGHT * spp * world
    value = samples_per_step * args.steps / (total_ms * 1e-3) / 1e6

    # end to end through the reference-facing call with host buffers (Pt_TraceSpp): the step's
    # inputs (camera + dof structs) go host->device, the three images come back device->host
    trace = api.Trace(WIDTH, HEIGHT)
    for i in range(2):
        trace.trace_spp(scene, cam, dof, i * spp, spp)
    barrier()
    t0 = time.perf_counter()
    e2e_steps = max(2, min(args.steps, 8))
    for i in range(e2e_steps):
        trace.trace_spp(scene, cam, dof, (2 + i) * spp, spp)
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_s = multigpu.max_over_ranks(e2e_s, dev)
    e2e_value = WIDTH * HEIGHT * spp * world * e2e_steps / e2e_s / 1e6
    mean_lum = float(trace.image("color").mean())
    trace.close()

    if rank == 0:
        peak, peak_src = measured_peaks()
        tc = traversal_counts()
        b_ray = 32 + 16 + tc["nodes_per_ray"] * 80 + tc["tris_per_ray"] * 48
        kern = {}
        for name in api.KERNEL_NAMES:
            k = kstats[name]
            kern[name] = {"ms": k["ms"], "launches": k["launches"],
                          "share": k["ms"] / max(sum(kstats[n]["ms"] for n in api.KERNEL_NAMES), 1e-9)}
        # the traversal kernels are the ones the bytes-per-ray roofline is about
        ext = kstats["extend"]
        rays_per_launch = kstats["extendRays"] / max(ext["launches"], 1)
        avg_s = ext["ms"] * 1e-3 / max(ext["launches"], 1)
        achieved = rays_per_launch * b_ray / max(avg_s, 1e-12) / 1e9
        total_rays = kstats["extendRays"] + kstats["connectRays"]
        # whole-step figure of SURVEY.md 8(d): B_sample = 72 + L*(R*B_ray + 192), L = bounces per
        # path, R = rays per bounce, 192 = one read + one write of the path state per bounce
        paths = max(counters["paths"], 1)
        bounces = kstats["extendRays"] / paths
        rays_per_bounce = total_rays / max(kstats["extendRays"], 1)
        b_sample = 72 + bounces * (rays_per_bounce * b_ray + 192)
        sample_gbs = value * 1e6 / world * b_sample / 1e9
        # measured DRAM bytes per ray of k_wfExtend from the committed ncu capture (path state in, hit out)
        traffic = None
        l2 = None
        tpath = os.path.join(ROOT, "profiles", "r01_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            traffic = tj["k_wfExtend"]["dram_bytes_per_ray"] * rays_per_launch
            if "l2_bytes_per_ray" in tj["k_wfExtend"] and "l2_copy_peak_gbs" in tj:
                # the tree is cache resident: the kernel's L2 traffic (ncu capture) at this run's ray rate
                # against the L2-resident copy rate measured with tools/l2peak.py
                l2_gbs = tj["k_wfExtend"]["l2_bytes_per_ray"] * rays_per_launch / avg_s / 1e9
                l2 = {"achieved": l2_gbs, "peak": tj["l2_copy_peak_gbs"], "unit": "GB/s", "frac": l2_gbs / tj["l2_copy_peak_gbs"],
                      "bytes_per_ray": tj["k_wfExtend"]["l2_bytes_per_ray"], "source": "profiles/r01_traffic.json"}
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "spp_per_step": spp, "width": WIDTH, "height": HEIGHT,
                       "sharding": "sample index per rank, scene replicated, NCCL reduce of the sums per step",
                       "l2": "256 MB buffer written between timed steps (L2 flush)",
                       "path_pool_slots": args.pool or (1 << 23)},
            "mrays_per_s": value * total_rays / max(counters["paths"], 1),      # all ranks, from rank 0's rays per sample
            "rays_per_sample": total_rays / max(counters["paths"], 1),
            "gpu_launches": int(counters["kernelLaunches"]),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": C.sizeof(api.Camera) + C.sizeof(api.PtDofInfo),
                    "d2h_bytes_per_step": 3 * n3 * 4, "steps": e2e_steps, "mean_luminance": mean_lum},
            "roofline": {"bound": "hbm", "kernel": "k_wfExtend", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_ray": b_ray, "rays_per_launch": rays_per_launch,
                         "avg_launch_ms": avg_s * 1e3, "l2": l2,
                         "note": "108-triangle BVH is L1/L2 resident: algorithmic bytes never reach HBM, "
                                 "so frac can exceed 1 (SURVEY.md 8(d))"},
            "roofline_sample": {"bound": "hbm", "bytes_per_sample": b_sample, "bounces_per_path": bounces,
                                "rays_per_bounce": rays_per_bounce, "achieved": sample_gbs, "peak": peak, "unit": "GB/s",
                                "frac": sample_gbs / peak, "note": "per GPU; whole step by SURVEY.md 8(d) B_sample"},
            "kernels": kern,
            "clocks": clocks,
        }
        if not args.no_cpu_baseline and world == 1:
            try:
                line["cpu_baseline"] = cpu_baseline(WIDTH, HEIGHT)
            except Exception as e:      # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"error": str(e)}
        print(json.dumps(line))
    scene.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
