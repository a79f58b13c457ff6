#!/usr/bin/env python
"""bench.py — Msamples/s of the path-tracing hot path on cornell_box at 1920x1080.

A step is one progressive pass of the hot path: `--spp-per-step` samples of every
pixel of the 1080p frame (BASELINE.json configs[1]; the full config is 1024 spp =
16 steps of 64 spp). `value` times Pt_TraceSpp_Device with the scene and the
accumulation images resident in HBM; `e2e` times the reference-facing call with host
PtTrace buffers (Pt_TraceSpp: device->host copies of the three images inside). With
N > 1 every rank renders its own sample indices of the full frame (scene replicated)
into per-step sums, one NCCL reduce per step adds them on rank 0, and `e2e` is
render + reduce + one device->host copy of the three images on rank 0.

The other BASELINE configs ride along in `extra` (C1 320x180x64 spp, C3 1 M-triangle
interior 1080p x 256 spp, C4 lightmap bake 16 texels/m x 64 spp, C5 10 M triangles x
100 M rays closest / any hit), each with its own roofline and CPU baseline; `verify`
holds the N-rank image against the 1-rank image over the same (pixel, sample) set.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--extras LIST|none]
"""
import argparse
import ctypes as C
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
C3_EYE, C3_AT = [3.0, 1.5, 3.0], [30.0, 1.2, 27.0]
METRIC, UNIT = "Msamples/s (1080p cornell_box)", "Msamples/s"
WORKLOAD = "cornell_box boxes 1920x1080 progressive accumulation (BASELINE configs[1])"
TRAVERSAL_COUNTS = os.path.join(ROOT, "baseline", "traversal_counts.json")
TRAFFIC = os.path.join(ROOT, "profiles", "r02_traffic.json")
DEFAULT_EXTRAS = "c1,c3,c4,c5"
BACKEND = ("reference path_tracer.c over the CPU BVH restatement (Embree 3.13.5 binary unavailable), "
           "work chunked as threading/task.c:35-64 does")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=64)
    ap.add_argument("--pool", type=int, default=0, help="path pool slots (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--extras", default=DEFAULT_EXTRAS,
                    help="comma list of c1,c3,c4,c5 or 'none' (N > 1 runs only c3 and c4: the configs BASELINE.md shards)")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def traversal_counts(key):
    """mean wide-node / triangle fetches per ray and rays per sample of the canonical CPU
    BVH on a workload; written by baseline/make_traversal_counts.py"""
    with open(TRAVERSAL_COUNTS) as f:
        return json.load(f)[key]


def bytes_per_ray(tc, prefix=""):
    return 32 + 16 + tc[prefix + "nodes_per_ray"] * 80 + tc[prefix + "tris_per_ray"] * 48


def host_cores():
    return min(os.cpu_count() or 1, 64)


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


def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _oracle as O
    return O


def _oracle_cornell(O, seed=1, lut_samples=1024):
    from pim_b200 import scenes
    O.set_threads(host_cores())
    O.set_task_mode(1)          # timing runs: work handed out as the reference's task.c does (chunks of N / T^2, no per-item reseed)
    O.set_seed(seed)
    lut = O.bake_brdf_lut(128, lut_samples)
    desc = scenes.cornell_box("boxes", brdf_lut=lut, seed=seed)
    O.load_scene_desc(desc)
    return desc


# ---------------------------------------------------------------- CPU baselines (reference code on the host cores)

def _cpu_frames(O, ref, eye, at, width, height, min_seconds, max_frames):
    cam, dof = O.camera_lookat(eye, at, FOVY), O.default_dof()
    dof.autoFocus = 0
    col = np.zeros((height, width, 3), np.float32); alb = col.copy(); nor = col.copy()
    ref.trace(col, alb, nor, 1.0, cam, dof, 0)          # warm-up frame (page faults, caches)
    O.lib().oracle_counters_reset()
    frames, t0 = 0, time.perf_counter()
    while True:
        ref.trace(col, alb, nor, 1.0 / (frames + 2), cam, dof, frames + 1)
        frames += 1
        dt = time.perf_counter() - t0
        if dt >= min_seconds or frames >= max_frames:
            break
    _, _, rays = O.counters()
    return frames, dt, rays, float(col.mean())


def cpu_baseline(width, height, min_seconds=10.0, max_frames=64):
    """The reference's own Pt_Trace (oracle/_ref) on the host cores, same scene and camera."""
    O = _oracle()
    if not O.available():
        return None
    _oracle_cornell(O)
    ref = O.Scene()
    frames, dt, rays, _ = _cpu_frames(O, ref, EYE, AT, width, height, min_seconds, max_frames)
    ref.close()
    samples = frames * width * height
    return {"value": samples / dt / 1e6, "unit": UNIT, "cores": host_cores(), "kind": "reference",
            "sample": "%d frames of 1 spp at %dx%d, %s" % (frames, width, height, BACKEND),
            "seconds": dt, "mrays_per_s": rays / dt / 1e6, "rays_per_sample": rays / samples}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path on the host cores, on the arm's own
    config: a step is 1 spp of the full 1920x1080 frame, same scene and camera (one step is a few tenths of a
    second on 16 threads, so K steps + W warm-ups stay within minutes)."""
    if rank != 0:
        return
    O = _oracle()
    if not O.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpim_ref.so not built"}))
        return
    w, h = WIDTH, HEIGHT
    cores = host_cores()
    _oracle_cornell(O)
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
    sample = "per step: 1 spp of the full %dx%d frame, same scene and camera" % (w, h)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "spp_per_step": 1, "width": w, "height": h, "sample": sample, "threads": cores,
                   "backend": BACKEND},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mean_luminance": float(col.mean()),
    }))


# ---------------------------------------------------------------- the GPU arm

class Ctx:
    """what every measurement of the GPU arm shares"""
    def __init__(self, args, rank, world, local_rank):
        import torch
        import torch.distributed as dist
        from pim_b200 import api, scenes, multigpu
        self.torch, self.dist, self.api, self.scenes, self.multigpu = torch, dist, api, scenes, multigpu
        self.args, self.rank, self.world = args, rank, world
        torch.cuda.set_device(local_rank)
        self.dev = torch.device("cuda", local_rank)
        if world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.L = api.load()
        self.stream = torch.cuda.Stream(device=self.dev)   # the library launches on this stream; events are recorded on it
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)     # > 126 MB L2
        torch.cuda.synchronize()
        torch.cuda.set_stream(self.stream)
        self.peak, self.peak_src = measured_peaks()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def images(self, w, h):
        return [self.torch.zeros(w * h * 3, dtype=self.torch.float32, device=self.dev) for _ in range(3)]

    def trace(self, scene, cam, dof, w, h, begin, spp, sums):
        rc = self.L.Pt_TraceSpp_Device(scene.h, C.byref(cam), C.byref(dof), w, h, begin, spp,
                                       sums[0].data_ptr(), sums[1].data_ptr(), sums[2].data_ptr(), self.stream.cuda_stream)
        if rc != 0:
            raise RuntimeError("Pt_TraceSpp_Device: " + self.api.last_error())

    def timed_steps(self, fn, steps, warmup):
        """W untimed + K timed calls of fn(i); each timed call between two events on the launching stream, an L2
        flush before it (outside the events); returns the max over ranks of the summed milliseconds"""
        torch = self.torch
        for i in range(warmup):
            fn(i)
        self.barrier()
        total_ms = 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for i in range(steps):
            self.flush.fill_(i & 0xFF)
            e0.record(self.stream)
            fn(warmup + i)
            e1.record(self.stream)
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
        self.barrier()
        return self.multigpu.max_over_ranks(total_ms, self.dev)


def sample_roofline(ctx, value_per_gpu, tc_key, kstats, counters):
    """whole-step figure of SURVEY.md 8(d): B_sample = 72 + L*(R*B_ray + 192), L = bounces per path and R = rays
    per bounce as this run traced them, B_ray from the canonical CPU BVH's committed counts"""
    tc = traversal_counts(tc_key)
    b_ray = bytes_per_ray(tc)
    paths = max(counters["paths"], 1)
    total_rays = kstats["extendRays"] + kstats["connectRays"]
    bounces = kstats["extendRays"] / paths
    rays_per_bounce = total_rays / max(kstats["extendRays"], 1)
    b_sample = 72 + bounces * (rays_per_bounce * b_ray + 192)
    gbs = value_per_gpu * 1e6 * b_sample / 1e9
    return {"bound": "hbm", "bytes_per_sample": b_sample, "bytes_per_ray": b_ray, "bounces_per_path": bounces,
            "rays_per_bounce": rays_per_bounce, "rays_per_sample": total_rays / paths, "achieved": gbs,
            "peak": ctx.peak, "unit": "GB/s", "frac": gbs / ctx.peak, "peak_source": ctx.peak_src,
            "counts": "baseline/traversal_counts.json:" + tc_key,
            "note": "per GPU; whole step by SURVEY.md 8(d) B_sample"}


def kernel_table(api, kstats):
    total = max(sum(kstats[n]["ms"] for n in api.KERNEL_NAMES), 1e-9)
    return {n: {"ms": kstats[n]["ms"], "launches": kstats[n]["launches"], "share": kstats[n]["ms"] / total,
                "us_per_launch": kstats[n]["ms"] / max(kstats[n]["launches"], 1) * 1e3} for n in api.KERNEL_NAMES}


def dominant_kernel_roofline(ctx, kstats, counters):
    """`roofline` of the line: the kernel with the largest share of the step (CUDA events on the launching stream
    inside the library). Traversal kernels by B_ray x rays per launch; shading kernels by the path state they are
    specified to move per slot (96 B state read + 96 B written per bounce, shared by the two
    shading stages, plus the two rays' 48 B in / out)."""
    api = ctx.api
    tab = kernel_table(api, kstats)
    name = max(api.KERNEL_NAMES, key=lambda n: tab[n]["ms"])
    launches = max(kstats[name]["launches"], 1)
    avg_s = kstats[name]["ms"] * 1e-3 / launches
    tc = traversal_counts("cornell_boxes_camera_paths")
    b_ray = bytes_per_ray(tc)
    slots = kstats["extendRays"] / max(kstats["extend"]["launches"], 1)        # live paths per iteration
    if name in ("extend", "connect"):
        units = kstats[name + "Rays"] / launches
        per_unit, what = b_ray, "B_ray = 32 + 16 + N_node*80 + N_tri*48 per ray (canonical CPU BVH counts)"
    else:
        units = slots
        per_unit = {"shadeA": 96.0 + 48.0, "shadeB": 96.0 + 48.0, "regen": 96.0}[name]
        what = ("path state per slot: 96 B (one read or one write of the state of SURVEY 8(d)) + 48 B (a ray in, a hit out)"
                if name != "regen" else "96 B of new path state per slot")
    achieved = units * per_unit / max(avg_s, 1e-12) / 1e9
    traffic = None
    if os.path.exists(TRAFFIC):
        with open(TRAFFIC) as f:
            tj = json.load(f)
        key = "k_wf" + name[0].upper() + name[1:]
        if key in tj and "dram_bytes_per_slot" in tj[key]:
            traffic = tj[key]["dram_bytes_per_slot"] * units
    return {"bound": "hbm", "kernel": "k_wf" + name[0].upper() + name[1:], "achieved": achieved, "peak": ctx.peak,
            "unit": "GB/s", "frac": achieved / ctx.peak, "traffic": traffic, "peak_source": ctx.peak_src,
            "algorithmic_bytes_per_unit": per_unit, "units_per_launch": units, "avg_launch_ms": avg_s * 1e3,
            "share_of_step": tab[name]["share"], "definition": what,
            "traffic_source": "profiles/r02_traffic.json (dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture, per slot)"
                              if traffic is not None else None}


def traversal_roofline(ctx, kstats):
    """k_wfExtend by the bytes-per-ray definition of SURVEY.md 8(d). The 19-node cornell tree is served from L1 / L2,
    so the HBM fraction is a normalised ray rate that may exceed 1; the honest bound for it is the L2: the kernel's
    measured L2 traffic per ray (lts__t_sectors.sum x 32 B from the committed ncu capture) at this run's ray rate,
    against the L2 read bandwidth measured live with a read-only sweep of a 32 MB buffer (Pt_MeasureL2ReadGBs)."""
    tc = traversal_counts("cornell_boxes_camera_paths")
    b_ray = bytes_per_ray(tc)
    ext = kstats["extend"]
    launches = max(ext["launches"], 1)
    rays = kstats["extendRays"] / launches
    avg_s = ext["ms"] * 1e-3 / launches
    achieved = rays * b_ray / max(avg_s, 1e-12) / 1e9
    out = {"bound": "hbm", "kernel": "k_wfExtend", "bytes_per_ray": b_ray, "rays_per_launch": rays, "avg_launch_ms": avg_s * 1e3,
           "achieved": achieved, "peak": ctx.peak, "unit": "GB/s", "frac": achieved / ctx.peak,
           "note": "108-triangle BVH is L1/L2 resident: algorithmic bytes never reach HBM, so frac can exceed 1 (SURVEY.md 8(d))"}
    l2_peak = float(ctx.L.Pt_MeasureL2ReadGBs(32 << 20, 64))
    if l2_peak > 0 and os.path.exists(TRAFFIC):
        with open(TRAFFIC) as f:
            tj = json.load(f)
        per_ray = tj.get("k_wfExtend", {}).get("l2_bytes_per_slot")
        if per_ray:
            l2_gbs = per_ray * rays / max(avg_s, 1e-12) / 1e9
            out["l2"] = {"achieved": l2_gbs, "peak": l2_peak, "unit": "GB/s", "frac": l2_gbs / l2_peak, "bytes_per_ray": per_ray,
                         "peak_source": "Pt_MeasureL2ReadGBs: 64 read-only sweeps of a 32 MB buffer, measured in this run",
                         "traffic_source": "profiles/r02_traffic.json"}
    return out


def verify_ranks(ctx, scene_desc_fn):
    """The N-rank image against the 1-rank image over the same (pixel, sample) set: 32 samples per pixel of the 1080p
    frame, split by sample index over the ranks (fresh replicas, so every rank samples the light grid as built),
    summed with the bench's reduce, against the same 32 samples rendered by rank 0 alone in one call. Differences
    are the order of the float atomics only. mean_luminance is over that fixed sample set, so it is the same
    number at every N."""
    torch, api, mg = ctx.torch, ctx.api, ctx.multigpu
    total_spp = 32
    per_rank = total_spp // ctx.world
    cam, dof = api.make_camera(EYE, AT, FOVY), api.default_dof(auto_focus=False)
    scene = api.Scene(scene_desc_fn())
    sums = ctx.images(WIDTH, HEIGHT)
    ctx.trace(scene, cam, dof, WIDTH, HEIGHT, ctx.rank * per_rank, per_rank, sums)
    torch.cuda.synchronize()
    mg.reduce_sums(sums, dst=0)
    ctx.barrier()
    scene.close()
    out = None
    if ctx.rank == 0:
        n_rank = sums[0] / total_spp
        out = {"samples_per_pixel": total_spp, "mean_luminance": float(n_rank.mean().item())}
        if ctx.world > 1:
            solo = api.Scene(scene_desc_fn())
            one = ctx.images(WIDTH, HEIGHT)
            ctx.trace(solo, cam, dof, WIDTH, HEIGHT, 0, total_spp, one)
            torch.cuda.synchronize()
            solo.close()
            ref = one[0] / total_spp
            diff = (n_rank - ref).abs()
            out.update({"mean_luminance_1_rank": float(ref.mean().item()),
                        "max_abs_diff": float(diff.max().item()),
                        "rel_diff_of_means": float(abs(n_rank.mean().item() - ref.mean().item()) / ref.mean().item()),
                        "pixels_beyond_1e-4_rel": int((diff > 1e-4 * (ref.abs() + 1e-3)).sum().item())})
    ctx.barrier()
    return out


def run_ours(args, rank, world, local_rank):
    ctx = Ctx(args, rank, world, local_rank)
    torch, api, scenes, mg, L = ctx.torch, ctx.api, ctx.scenes, ctx.multigpu, ctx.L
    spp = args.spp_per_step

    def cornell():
        # the same seed on every rank: a path's random stream is a function of (pixel, sample index, seed), and
        # the ranks render disjoint sample indices (multigpu.sample_range)
        d = scenes.cornell_box("boxes", brdf_lut=None, seed=1)      # BRDF table baked on the device
        d.path_pool_size = args.pool
        return d

    scene = api.Scene(cornell())
    cam = api.make_camera(EYE, AT, FOVY)
    dof = api.default_dof(auto_focus=False)
    n3 = WIDTH * HEIGHT * 3
    sums = ctx.images(WIDTH, HEIGHT)                 # this step's samples of this rank
    total = ctx.images(WIDTH, HEIGHT) if rank == 0 else None     # rank 0: everything rendered so far, all ranks
    counters_sync = mg.LightCounterSync(scene, ctx.dev, ctx.stream)   # N > 1: all replicas learn from all ranks' paths

    def step(i):
        begin, _ = mg.sample_range(i, rank, world, spp)
        if world == 1:
            ctx.trace(scene, cam, dof, WIDTH, HEIGHT, begin, spp, total)     # one rank: straight into the frame's total
            return
        for t in sums:
            t.zero_()
        counters_sync.before()
        ctx.trace(scene, cam, dof, WIDTH, HEIGHT, begin, spp, sums)
        counters_sync.after()                        # all-reduce of the light-learning counters the call added (16 KB)
        mg.reduce_sums(sums, dst=0)                  # per-step sums, not running totals
        if rank == 0:
            for a, b in zip(total, sums):
                a += b

    for i in range(args.warmup):
        step(i)
    ctx.barrier()
    L.Pt_ResetCounters()
    L.Pt_SetKernelTiming(1)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    total_ms = ctx.timed_steps(lambda i: step(args.warmup + i), args.steps, 0)
    clocks = sampler.stop() if sampler else None
    counters = api.counters()
    kstats = api.kernel_stats()
    L.Pt_SetKernelTiming(0)
    samples_per_step = WIDTH * HEIGHT * spp * world
    value = samples_per_step * args.steps / (total_ms * 1e-3) / 1e6
    steps_done = args.warmup + args.steps
    image_mean = float((total[0] / (steps_done * spp * world)).mean().item()) if rank == 0 else None

    # ---- end to end
    e2e_steps = max(2, min(args.steps, 8))
    if world == 1:
        # through the reference-facing call with host buffers (Pt_TraceSpp): the step's inputs (camera + dof
        # structs) go host->device, the three images come back device->host
        trace = api.Trace(WIDTH, HEIGHT)
        for i in range(2):
            trace.trace_spp(scene, cam, dof, i * spp, spp)
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            trace.trace_spp(scene, cam, dof, (2 + i) * spp, spp)
        ctx.barrier()
        e2e_s = time.perf_counter() - t0
        e2e_lum = float(trace.image("color").mean())
        trace.close()
        e2e_how = "Pt_TraceSpp with host PtTrace buffers"
    else:
        # one image on one host: every rank renders its samples, the sums are reduced to rank 0, rank 0 adds them
        # to the frame's total and copies the three images to page-locked host memory
        host = [torch.empty(n3, dtype=torch.float32, pin_memory=True) for _ in range(3)] if rank == 0 else None

        def e2e_step(i):
            step(steps_done + i)
            if rank == 0:
                for h, t in zip(host, total):
                    h.copy_(t, non_blocking=True)
            torch.cuda.synchronize()
        for i in range(2):
            e2e_step(i)
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            e2e_step(2 + i)
        ctx.barrier()
        e2e_s = time.perf_counter() - t0
        e2e_lum = float((host[0] / ((steps_done + 2 + e2e_steps) * spp * world)).mean().item()) if rank == 0 else None
        e2e_how = "Pt_TraceSpp_Device per rank + NCCL reduce to rank 0 + device->host copy of the three images on rank 0"
    e2e_s = mg.max_over_ranks(e2e_s, ctx.dev)
    e2e_value = WIDTH * HEIGHT * spp * world * e2e_steps / e2e_s / 1e6

    verify = verify_ranks(ctx, cornell)

    line = None
    if rank == 0:
        total_rays = kstats["extendRays"] + kstats["connectRays"]
        paths = max(counters["paths"], 1)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "spp_per_step": spp, "width": WIDTH, "height": HEIGHT,
                       "sharding": "sample index per rank, scene replicated, NCCL reduce of the per-step sums to rank 0, "
                                   "all-reduce of the light-learning counters each step adds",
                       "l2": "256 MB buffer written between timed steps (L2 flush)",
                       "path_pool_slots": args.pool or (1 << 23)},
            "mrays_per_s": value * total_rays / paths,      # all ranks, from rank 0's rays per sample
            "rays_per_sample": total_rays / paths,
            "gpu_launches": int(counters["kernelLaunches"]),
            "mean_luminance": image_mean,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": C.sizeof(api.Camera) + C.sizeof(api.PtDofInfo),
                    "d2h_bytes_per_step": 3 * n3 * 4, "steps": e2e_steps, "mean_luminance": e2e_lum, "how": e2e_how},
            "roofline": dominant_kernel_roofline(ctx, kstats, counters),
            "roofline_sample": sample_roofline(ctx, value / world, "cornell_boxes_camera_paths", kstats, counters),
            "roofline_traversal": traversal_roofline(ctx, kstats),
            "kernels": kernel_table(api, kstats),
            "clocks": clocks,
            "verify": verify,
        }
    scene.close()

    # ---- the other BASELINE configs
    wanted = [] if args.extras in ("none", "") else [x.strip() for x in args.extras.split(",") if x.strip()]
    if world > 1:
        wanted = [x for x in wanted if x in ("c3", "c4")]
    extra = {}
    for name in wanted:
        fn = {"c1": extra_c1, "c3": extra_c3, "c4": extra_c4, "c5": extra_c5}.get(name)
        if fn is None:
            continue
        try:
            r = fn(ctx)
        except Exception as e:      # an extra never takes the headline down with it
            r = {"error": "%s: %s" % (type(e).__name__, e)}
        if rank == 0:
            extra[name] = r
        ctx.barrier()

    if world > 1:
        # the in-process form: rank 0 alone drives all the GPUs; the others wait on a host-side (gloo) barrier so
        # that no NCCL kernel spins on their devices meanwhile
        cpu_group = ctx.dist.new_group(backend="gloo")
        if rank == 0:
            try:
                extra["inproc"] = extra_inproc(ctx)
            except Exception as e:
                extra["inproc"] = {"error": "%s: %s" % (type(e).__name__, e)}
        ctx.dist.barrier(group=cpu_group)
    if rank == 0:
        line["extra"] = extra
        if not args.no_cpu_baseline and world == 1:
            try:
                line["cpu_baseline"] = cpu_baseline(WIDTH, HEIGHT)
            except Exception as e:      # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"error": str(e)}
        print(json.dumps(line))
    if world > 1:
        ctx.dist.destroy_process_group()


# ---------------------------------------------------------------- the other BASELINE configs

def _timed_trace(ctx, scene, cam, dof, w, h, total_spp, chunk, warm_spp):
    """device-timed Pt_TraceSpp_Device over total_spp samples per pixel in calls of `chunk` (this rank's sample
    indices), after warm_spp untimed samples; per-call sums reduced to rank 0 like the main step"""
    torch, api, mg, L = ctx.torch, ctx.api, ctx.multigpu, ctx.L
    sums = ctx.images(w, h)
    total = ctx.images(w, h) if ctx.rank == 0 else None
    calls = total_spp // chunk

    def call(i):
        begin, _ = mg.sample_range(i, ctx.rank, ctx.world, chunk)
        for t in sums:
            t.zero_()
        ctx.trace(scene, cam, dof, w, h, begin, chunk, sums)
        mg.reduce_sums(sums, dst=0)
        if ctx.rank == 0:
            for a, b in zip(total, sums):
                a += b
    warm_calls = max(1, warm_spp // chunk)
    for i in range(warm_calls):
        call(i)
    ctx.barrier()
    L.Pt_ResetCounters(); L.Pt_SetKernelTiming(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.flush.fill_(1)
    e0.record(ctx.stream)
    for i in range(calls):
        call(warm_calls + i)
    e1.record(ctx.stream)
    e1.synchronize()
    ms = mg.max_over_ranks(e0.elapsed_time(e1), ctx.dev)
    ctx.barrier()
    counters, kstats = api.counters(), api.kernel_stats()
    L.Pt_SetKernelTiming(0)
    lum = float((total[0] / ((warm_calls + calls) * chunk * ctx.world)).mean().item()) if ctx.rank == 0 else None
    return ms, counters, kstats, lum


def extra_c1(ctx):
    """BASELINE configs[0]: cornell boxes via pt_test at 320x180, 64 spp (the reference's CPU-runnable case)"""
    api, scenes = ctx.api, ctx.scenes
    W, H, spp = 320, 180, 64
    scene = api.Scene(scenes.cornell_box("boxes", brdf_lut=None, seed=1))
    cam, dof = api.make_camera(EYE, AT, FOVY), api.default_dof(auto_focus=False)
    ms, counters, kstats, lum = _timed_trace(ctx, scene, cam, dof, W, H, spp, spp, spp)
    value = W * H * spp / ms / 1e3
    # end to end as pt_test runs it: 64 calls of Pt_Trace (1 spp each, host images refreshed every call)
    tr = api.Trace(W, H)
    tr.trace(scene, cam, dof, 1.0)
    t0 = time.perf_counter()
    for p in range(spp):
        tr.trace(scene, cam, dof, 1.0 / (p + 1))
    e2e_s = time.perf_counter() - t0
    tr.close(); scene.close()
    out = {"workload": "cornell_box boxes via pt_test 320x180, 64 spp (BASELINE configs[0])", "metric": "Msamples/s",
           "value": value, "ms": ms, "mean_luminance": lum,
           "e2e": {"value": W * H * spp / e2e_s / 1e6, "unit": "Msamples/s", "how": "64 calls of Pt_Trace with host PtTrace buffers",
                   "d2h_bytes_per_step": 3 * W * H * 12},
           "roofline": sample_roofline(ctx, value, "cornell_boxes_camera_paths", kstats, counters)}
    if not ctx.args.no_cpu_baseline:
        O = _oracle()
        if O.available():
            _oracle_cornell(O)
            ref = O.Scene()
            frames, dt, rays, clum = _cpu_frames(O, ref, EYE, AT, W, H, 1e9, spp)
            ref.close()
            out["cpu_baseline"] = {"value": frames * W * H / dt / 1e6, "unit": "Msamples/s", "cores": host_cores(), "kind": "reference",
                                   "sample": "the whole config: %d frames of 1 spp at %dx%d, %s" % (frames, W, H, BACKEND),
                                   "mrays_per_s": rays / dt / 1e6, "mean_luminance": clum}
    return out


def extra_c3(ctx):
    """BASELINE configs[2]: ~1 M-triangle procedural interior with emissive panels, 1080p, 256 spp"""
    api, scenes = ctx.api, ctx.scenes
    t0 = time.perf_counter()
    desc = scenes.interior(1_000_000, seed=7, emissive_fraction=0.001)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    scene = api.Scene(desc)
    build_s = time.perf_counter() - t0
    st = scene.stats
    cam, dof = api.make_camera(C3_EYE, C3_AT, FOVY), api.default_dof(auto_focus=False)
    spp, chunk = 256, 16
    ms, counters, kstats, lum = _timed_trace(ctx, scene, cam, dof, WIDTH, HEIGHT, spp, chunk, chunk)
    value = WIDTH * HEIGHT * spp * ctx.world / ms / 1e3
    total_rays = kstats["extendRays"] + kstats["connectRays"]
    out = {"workload": "procedural interior, %d triangles, %d emissive, 1920x1080, 256 spp per rank (BASELINE configs[2])"
                       % (desc.tri_count, st.emissiveCount),
           "metric": "Msamples/s", "value": value, "ms": ms, "n_gpus": ctx.world, "mean_luminance": lum,
           "mrays_per_s": value * total_rays / max(counters["paths"], 1),
           "scene": {"generate_s": gen_s, "build_s": build_s, "bvh_nodes": st.bvhNodeCount, "grid": list(st.gridSize),
                     "build_phase_ms": dict(zip(["upload_flatten", "bvh", "emissive", "grid", "lut", "pool"],
                                                [round(x, 1) for x in st.phaseMs]))},
           "roofline": sample_roofline(ctx, value / ctx.world, "interior_1m_camera_paths", kstats, counters),
           "kernels": kernel_table(api, kstats)}
    lut = scene.brdf_lut()
    scene.close()
    if ctx.world == 1 and not ctx.args.no_cpu_baseline:
        O = _oracle()
        if O.available():
            O.set_threads(host_cores()); O.set_task_mode(1); O.set_seed(1)
            desc.brdf_lut = lut
            t0 = time.perf_counter()
            O.load_scene_desc(desc)
            ref = O.Scene()
            cpu_build = time.perf_counter() - t0
            w, h = WIDTH // 4, HEIGHT // 4
            frames, dt, rays, clum = _cpu_frames(O, ref, C3_EYE, C3_AT, w, h, 4.0, 16)
            ref.close()
            out["cpu_baseline"] = {"value": frames * w * h / dt / 1e6, "unit": "Msamples/s", "cores": host_cores(), "kind": "reference",
                                   "sample": "%d frames of 1 spp at %dx%d (1/16 of the frame), same scene and camera, %s" % (frames, w, h, BACKEND),
                                   "scene_build_s": cpu_build, "mrays_per_s": rays / dt / 1e6, "mean_luminance": clum}
    return out


def extra_c4(ctx):
    """BASELINE configs[3]: lm_gen bake of cornell_box at 16 texels/m, 64 spp. LmPack_Pack(1024, 16.0, 0.1, 15 deg) runs
    as reference code on rank 0's host (the shared input); the pack then lives on the device(s). With N ranks the covered
    texels are split into N contiguous ranges (the fit is sequential per texel, so texels, never samples, are sharded),
    every rank bakes its range and the disjoint slabs are gathered on rank 0: no reduction."""
    torch, api, scenes, mg, L = ctx.torch, ctx.api, ctx.scenes, ctx.multigpu, ctx.L
    O = _oracle()
    if not O.available():
        return {"error": "oracle/_ref not built: no LmPack_Pack to make the shared input"}
    size, tpm, spp = 1024, 16.0, 64
    n = size * size
    pack_s = 0.0
    if ctx.rank == 0:
        desc = _oracle_cornell(O)
        t0 = time.perf_counter()
        pages = O.lm_pack(size, tpm, 0.1, 15.0)
        pack_s = time.perf_counter() - t0
        assert pages == 1
        probes, pos, nrm, cnt = O.lm_page(0)
        h_probes = [np.ascontiguousarray(p.copy()) for p in probes]
        h_pos, h_nrm, h_cnt = pos.copy(), nrm.copy(), cnt.copy()
        d_probes = torch.from_numpy(np.stack(h_probes)).to(ctx.dev)
        d_pos, d_nrm, d_cnt = torch.from_numpy(h_pos).to(ctx.dev), torch.from_numpy(h_nrm).to(ctx.dev), torch.from_numpy(h_cnt).to(ctx.dev)
    else:
        d_probes = torch.zeros((5, n, 4), dtype=torch.float32, device=ctx.dev)
        d_pos = torch.zeros((n, 3), dtype=torch.float32, device=ctx.dev); d_nrm = torch.zeros_like(d_pos)
        d_cnt = torch.zeros(n, dtype=torch.float32, device=ctx.dev)
    if ctx.world > 1:
        for t in (d_probes, d_pos, d_nrm, d_cnt):
            ctx.dist.broadcast(t, src=0)
    covered_idx = torch.nonzero(d_cnt > 0).flatten()
    covered = int(covered_idx.numel())
    begin, count = mg.texel_range(covered, ctx.rank, ctx.world)
    mine = covered_idx[begin:begin + count]
    start_cnt = d_cnt.clone()
    scene = api.Scene(scenes.cornell_box("boxes", brdf_lut=None, seed=1))

    def bake(pass_index):
        # this rank's range: every other texel is switched off through its sample count (BakeFn skips those)
        work = torch.zeros_like(d_cnt)
        work[mine] = d_cnt[mine]
        ctx.stream.synchronize()                        # the bake runs on the scene's own stream and blocks
        scene.bake_lightmap_device([d_probes[k].data_ptr() for k in range(5)], d_pos.data_ptr(), d_nrm.data_ptr(), work.data_ptr(),
                                   size, tpm, 1.0, spp, pass_index)
        d_cnt[mine] = work[mine]
        if ctx.world > 1:
            # disjoint slabs to rank 0 (equal sizes up to one texel: pad to the largest)
            m = (covered + ctx.world - 1) // ctx.world
            slab = torch.zeros((m, 21), dtype=torch.float32, device=ctx.dev)
            slab[:count, :20] = d_probes[:, mine, :].permute(1, 0, 2).reshape(count, 20)
            slab[:count, 20] = d_cnt[mine]
            parts = [torch.zeros_like(slab) for _ in range(ctx.world)] if ctx.rank == 0 else None
            ctx.dist.gather(slab, parts, dst=0)
            if ctx.rank == 0:
                for r, part in enumerate(parts):
                    b, c = mg.texel_range(covered, r, ctx.world)
                    idx = covered_idx[b:b + c]
                    d_probes[:, idx, :] = part[:c, :20].reshape(c, 5, 4).permute(1, 0, 2)
                    d_cnt[idx] = part[:c, 20]
    bake(99)                                            # warm-up pass (also grows the library's scratch once)
    ctx.barrier()
    L.Pt_ResetCounters(); L.Pt_SetKernelTiming(1)
    reps = 3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ctx.stream)
    for p in range(reps):
        bake(p)
    e1.record(ctx.stream)
    e1.synchronize()
    ms = mg.max_over_ranks(e0.elapsed_time(e1), ctx.dev) / reps
    ctx.barrier()
    counters, kstats = api.counters(), api.kernel_stats()
    L.Pt_SetKernelTiming(0)
    value = covered * spp / ms / 1e3
    out = None
    if ctx.rank == 0:
        advanced = bool(((d_cnt - start_cnt)[covered_idx] == (reps + 1) * spp).all().item())
        tc = traversal_counts("cornell_lightmap_16tpm_bake_paths")
        b_ray = bytes_per_ray(tc)
        paths = max(counters["paths"], 1)
        total_rays = kstats["extendRays"] + kstats["connectRays"]
        bounces = kstats["extendRays"] / paths
        rpb = total_rays / max(kstats["extendRays"], 1)
        b_sample = 2 * 108.0 / spp + 32 + 12 + bounces * (rpb * b_ray + 192)     # page in/out amortised + ray in, colour out
        gbs = value / ctx.world * 1e6 * b_sample / 1e9
        out = {"workload": "lm_gen bake of cornell_box, LmPack_Pack(1024, 16.0, 0.1, 15), 64 spp, pack resident on the device (BASELINE configs[3])",
               "metric": "M texel-samples/s", "value": value, "ms_per_pass": ms, "n_gpus": ctx.world, "covered_texels": covered,
               "bake_mode": "PtBake_Wavefront (default; statistically, not texel for texel, equal to the seeded reference: tests/test_gpu_configs.py)",
               "sharding": "contiguous ranges of the covered texels per rank, slabs gathered on rank 0",
               "all_texels_advanced": advanced, "lm_pack_cpu_s": pack_s,
               "mrays_per_s": value * total_rays / paths,
               "roofline": {"bound": "hbm", "bytes_per_sample": b_sample, "bytes_per_ray": b_ray, "bounces_per_path": bounces,
                            "rays_per_bounce": rpb, "achieved": gbs, "peak": ctx.peak, "unit": "GB/s", "frac": gbs / ctx.peak,
                            "counts": "baseline/traversal_counts.json:cornell_lightmap_16tpm_bake_paths", "note": "per GPU"}}
        if ctx.world == 1:
            # end to end through the reference-facing form: LmPack_BakeDesc on the host page (covered texels cross the bus)
            g_probes = [p.copy() for p in h_probes]; g_cnt = h_cnt.copy()
            scene.bake_lightmap(g_probes, h_pos, h_nrm, g_cnt, tpm, 1.0, spp, 50)
            t0 = time.perf_counter()
            for p in range(reps):
                scene.bake_lightmap(g_probes, h_pos, h_nrm, g_cnt, tpm, 1.0, spp, 51 + p)
            dt = (time.perf_counter() - t0) / reps
            out["e2e"] = {"value": covered * spp / dt / 1e6, "unit": "M texel-samples/s", "how": "LmPack_BakeDesc with the host page",
                          "h2d_bytes_per_step": covered * 112, "d2h_bytes_per_step": covered * 112}
            if not ctx.args.no_cpu_baseline:
                ref = O.Scene()
                O.lm_bake(ref, 1.0, 1, 0)
                t0 = time.perf_counter()
                O.lm_bake(ref, 1.0, spp, 1)
                dt = time.perf_counter() - t0
                ref.close()
                out["cpu_baseline"] = {"value": covered * spp / dt / 1e6, "unit": "M texel-samples/s", "cores": host_cores(), "kind": "reference",
                                       "sample": "the whole config: one LmPack_Bake(scene, 1, 64) pass, reference lightmap.c + path_tracer.c over the CPU BVH restatement"}
    scene.close()
    return out


def extra_c5(ctx):
    """BASELINE configs[4]: incoherent-ray intersection microbench, 100 M random rays against a 10 M-triangle soup,
    closest hit (Pt_IntersectN_Device) and any hit (Pt_OccludedN_Device), rays resident on the device"""
    torch, api, scenes, L = ctx.torch, ctx.api, ctx.scenes, ctx.L
    tris, batch, batches = 10_000_000, 1 << 24, 6           # 6 x 16.8 M = 100.7 M rays
    t0 = time.perf_counter()
    desc = scenes.random_soup(tris, seed=11)
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    scene = api.Scene(desc)
    build_s = time.perf_counter() - t0
    st = scene.stats
    ro, rd = scenes.random_rays(batch, seed=13)
    _, rd_o = scenes.random_rays(batch, seed=13, occlusion=True)
    d_ro, d_rd, d_rdo = torch.from_numpy(ro).to(ctx.dev), torch.from_numpy(rd).to(ctx.dev), torch.from_numpy(rd_o).to(ctx.dev)
    d_hits = torch.empty((batch, 12), dtype=torch.float32, device=ctx.dev)
    d_occ = torch.empty(batch, dtype=torch.uint8, device=ctx.dev)
    torch.cuda.synchronize()

    def timed(fn):
        fn()
        ctx.stream.synchronize()
        best = None
        for _ in range(2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ctx.flush.fill_(3)
            e0.record(ctx.stream)
            for _ in range(batches):
                fn()
            e1.record(ctx.stream)
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return best
    sp = ctx.stream.cuda_stream
    ms_c = timed(lambda: L.Pt_IntersectN_Device(scene.h, d_ro.data_ptr(), d_rd.data_ptr(), 0.0, d_hits.data_ptr(), batch, sp))
    ms_o = timed(lambda: L.Pt_OccludedN_Device(scene.h, d_ro.data_ptr(), d_rdo.data_ptr(), 0.0, d_occ.data_ptr(), batch, sp))
    total = batch * batches
    closest, anyhit = total / ms_c / 1e3, total / ms_o / 1e3
    own_nodes, own_tris = scene.fetch_counts(ro[:1 << 20], rd[:1 << 20])
    hits = d_hits.cpu().numpy().view(np.uint32)
    tc = traversal_counts("soup_10m_random_rays")
    b_ray, b_any = bytes_per_ray(tc), 32 + 1 + tc["anyhit_nodes_per_ray"] * 80 + tc["anyhit_tris_per_ray"] * 48
    out = {"workload": "100.7 M random rays (6 batches of 2^24, seed 13) vs 10 M-triangle soup (seed 11) (BASELINE configs[4])",
           "metric": "Mrays/s", "closest_hit_mrays_s": closest, "any_hit_mrays_s": anyhit, "value": closest,
           "hit_fraction": float((hits[:, 8] != 0).mean()), "occluded_fraction": float(d_occ.float().mean().item()),
           "scene": {"generate_s": gen_s, "build_s": build_s, "bvh_build_ms": st.phaseMs[1], "bvh_nodes": st.bvhNodeCount},
           "tree": {"own_nodes_per_ray": own_nodes / (1 << 20), "own_tris_per_ray": own_tris / (1 << 20),
                    "canonical_nodes_per_ray": tc["nodes_per_ray"], "canonical_tris_per_ray": tc["tris_per_ray"]},
           "roofline": {"bound": "hbm", "kernel": "k_closestHitPersistent", "bytes_per_ray": b_ray,
                        "achieved": closest * 1e6 * b_ray / 1e9, "peak": ctx.peak, "unit": "GB/s",
                        "frac": closest * 1e6 * b_ray / 1e9 / ctx.peak, "counts": "baseline/traversal_counts.json:soup_10m_random_rays"},
           "roofline_any_hit": {"bound": "hbm", "kernel": "k_anyHitPersistent", "bytes_per_ray": b_any,
                                "achieved": anyhit * 1e6 * b_any / 1e9, "peak": ctx.peak, "unit": "GB/s",
                                "frac": anyhit * 1e6 * b_any / 1e9 / ctx.peak}}
    # end to end with host rays and host hits (Pt_IntersectN): 32 B in, 48 B out per ray over PCIe
    n_e2e = 1 << 22
    h_hits = np.zeros(n_e2e, api.HIT_DTYPE)           # the caller's result buffer, reused from call to call
    scene.intersect(ro[:n_e2e], rd[:n_e2e], out=h_hits)
    dt = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        scene.intersect(ro[:n_e2e], rd[:n_e2e], out=h_hits)
        dt = min(dt, time.perf_counter() - t0)
    out["e2e"] = {"value": n_e2e / dt / 1e6, "unit": "Mrays/s", "how": "Pt_IntersectN, 4.2 M host rays in, PtRayHit out",
                  "h2d_bytes_per_step": n_e2e * 32, "d2h_bytes_per_step": n_e2e * 48}
    scene.close()
    if not ctx.args.no_cpu_baseline:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import truth
        verts = desc.entities[0].positions
        t0 = time.perf_counter()
        bvh = truth.CpuBvh(verts)
        cpu_build = time.perf_counter() - t0
        n = 2_000_000
        t0 = time.perf_counter()
        prim, _, _ = bvh.intersect(ro[:n], rd[:n])
        dt = time.perf_counter() - t0
        t0 = time.perf_counter()
        bvh.occluded(ro[:n], rd_o[:n])
        dt_o = time.perf_counter() - t0
        bvh.close()
        gpu_prim = np.where(hits[:n, 8] != 0, hits[:n, 9].view(np.int32) // 3, -1)
        out["cpu_baseline"] = {"value": n / dt / 1e6, "unit": "Mrays/s", "any_hit_mrays_s": n / dt_o / 1e6, "cores": truth.threads(), "kind": "port",
                               "sample": "2 M of the same rays against the same mesh, CPU BVH restatement of Embree's rtcIntersect1 / rtcOccluded (binary unavailable)",
                               "bvh_build_s": cpu_build, "agreement_with_gpu": float((gpu_prim == prim).mean())}
    return out


def extra_inproc(ctx):
    """The same step through the library's own multi-GPU mode (PtSys_SetDevices): ONE process drives all N devices
    with a host thread each, the way a single-process engine would. Rank 0 runs it while the other ranks wait;
    end to end with host PtTrace buffers (Pt_TraceSpp), N x 64 samples per pixel per call."""
    api, scenes = ctx.api, ctx.scenes
    n, spp = ctx.world, ctx.args.spp_per_step
    api.set_devices(n)
    try:
        scene = api.Scene(scenes.cornell_box("boxes", brdf_lut=None, seed=1))
        cam, dof = api.make_camera(EYE, AT, FOVY), api.default_dof(auto_focus=False)
        trace = api.Trace(WIDTH, HEIGHT)
        calls = max(2, min(ctx.args.steps, 8))
        for i in range(2):
            trace.trace_spp(scene, cam, dof, i * spp * n, spp * n)
        t0 = time.perf_counter()
        for i in range(calls):
            trace.trace_spp(scene, cam, dof, (2 + i) * spp * n, spp * n)
        dt = time.perf_counter() - t0
        lum = float(trace.image("color").mean())
        trace.close(); scene.close()
    finally:
        api.set_devices(1)
    return {"workload": WORKLOAD + ", one process, PtSys_SetDevices(%d)" % n, "metric": "Msamples/s", "n_gpus": n,
            "value": WIDTH * HEIGHT * spp * n * calls / dt / 1e6, "ms_per_call": dt / calls * 1e3, "mean_luminance": lum,
            "how": "Pt_TraceSpp with host PtTrace buffers; samples sharded over the devices inside the library, partial sums "
                   "to device 0 by peer copy, three images to the host"}


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
