#!/usr/bin/env python
"""bench.py — Mrays/s of the Whitted hot path (BASELINE.json metric) on N B200s, one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--res WxH]

A "step" is one Render() of the frame. `value` = rays of all timed steps / their device time with the scene, the
camera and the output image resident in HBM (snl_render_dev, CUDA events on the scene's stream, L2 flushed between
steps, max over ranks). `e2e` = the same through snl_render with a pinned HOST image (camera in, device->host copy of
the frame inside the timed region). `--impl reference` times the reference's own SSE packet tracer (oracle/_ref) on
the host cores. One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "Mrays/s on sponza 1080p Whitted (1/2/4/8 B200) vs host-CPU tracer"
UNIT = "Mrays/s"
DEFAULT_WORKLOAD = "barracks_whitted"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--res", default="1920x1080")
    ap.add_argument("--compose", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    return ap.parse_args()


def camera_for_step(sb, bmin, bmax, inside, step):
    """Bench camera of SURVEY 8(d), nudged sideways a little every step so that no per-origin state (the
    pre-transformed triangle records) survives from one frame to the next."""
    cam, light = sb.bench_camera(bmin, bmax, inside=inside)
    size = np.asarray(bmax, np.float32) - np.asarray(bmin, np.float32)
    s = float(np.sqrt((size * size).sum()))
    cam.pos[0] = float(np.float32(cam.pos[0]) + np.float32(0.0005 * s * step))
    return cam, light


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 7:
                    self.samples.append(f)
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[3 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm), "power_w_max": max(float(s[2]) for s in self.samples)}


def host_threads():
    return min(32, os.cpu_count() or 1)  # thread_pool.cpp:6 caps the reference at 32 threads


def reference_frames(w, W, H, frames, threads, first_step=0):
    """Times the reference's own Render() (SSE packet tracer, oracle/_ref) on the host cores. Returns
    (total_rays, total_seconds, frames)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import snailref as R
    import snail_b200 as sb
    variant = "sse" if R.available("sse") else "scalar"
    ref = R.RefScene(w["obj_path"], w["mtl_path"], w["tex_dir"], variant=variant)
    bmin, bmax = ref.bbox()
    ref.set_gvals(w["gvals"])
    rays, secs = 0, 0.0
    for i in range(frames):
        cam, light = camera_for_step(sb, bmin, bmax, w["inside"], first_step + i)
        ref.set_lights([[*light.pos, *light.color, light.radius]] if w["lit"] else [])
        _, st, ms = ref.render(cam.as_array(), W, H, threads=threads)
        rays += st["rays"]
        secs += ms * 1e-3
    ref.close()
    return rays, secs, variant


def run_reference(args, w, W, H):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    reference_frames(w, W, H, max(1, args.warmup), threads)  # warm-up (page in, thread pool spin-up)
    t0 = time.time()
    rays, secs, variant = reference_frames(w, W, H, args.steps, threads, first_step=args.warmup)
    wall = time.time() - t0
    val = rays / secs / 1e6
    sample = f"{args.steps} full {W}x{H} frames of the workload, Render() wall time only"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(secs / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "reference scene assets",
        "config": {"workload": f"{w['name']} {W}x{H}: {w['desc']}", "backend": f"oracle/_ref veclib {variant}"},
        "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": round(wall, 2),
    }
    print(json.dumps(line), flush=True)


def algorithmic_bytes(sb, w, W, H, cam, light):
    """B = 64 B x inner nodes fetched + 48 B x per-origin triangle records tested + 64 B x general
    triangle records tested + 16 B hit record per ray + 64 B ShTriangle per shaded hit + 3 B per pixel, counted by
    the SAME kernels built with -DSNL_COUNTERS (single-ray, near-child-first walk). One untimed frame."""
    if not os.path.exists(sb.api.LIB_COUNTERS_PATH):
        return None, None
    sc = sb.Scene(0, counters=True)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"])
    sc.set_options(w["gvals"])
    sc.set_lights([light] if w["lit"] else [])
    _, st = sc.render(cam, W, H)
    sc.close()
    shaded = st["reflection"] + st["reflection"]  # upper bound on shaded hits: primary hits + reflected hits
    b = (64 * st["node_visits"] + 48 * st["tri_tests"] + 64 * st["gtri_tests"] + 16 * st["rays"] + 64 * shaded + 3 * W * H)
    return b, st


def tiles_for_rank(W, H, rank, n):
    """Image-space split of the farm (src/server.cpp:233-265): blockWidth x blockHeight = 16 x 64 tiles
    (src/rtbase_math.h:31-34), interleaved diagonally over the ranks."""
    tiles = []
    for ty, y in enumerate(range(0, H, 64)):
        for tx, x in enumerate(range(0, W, 16)):
            if (tx + ty) % n == rank:
                tiles.append((x, y, min(16, W - x), min(64, H - y)))
    return tiles


def run_ours(args, w, W, H):
    import torch
    import snail_b200 as sb
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)

    sc = sb.Scene(local_rank)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"])
    sc.set_options(w["gvals"])
    bmin, bmax = sc.bbox()
    stream = torch.cuda.ExternalStream(sc.stream(), device=dev)
    tiles = tiles_for_rank(W, H, rank, world) if world > 1 else None
    pitch = W * 3
    image = torch.zeros((H, W, 3), dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
    host_np = host_img.numpy()

    def compose():
        if world > 1:  # NCCL compose: tiles are disjoint, every other pixel is zero -> a SUM reduce is a gather
            with torch.cuda.stream(stream):
                dist.reduce(image, dst=0, op=dist.ReduceOp.SUM)

    def step_resident(i):
        cam, light = camera_for_step(sb, bmin, bmax, w["inside"], i)
        sc.set_lights([light] if w["lit"] else [])
        flush.fill_(i & 255)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        sc.render_dev(cam, W, H, image.data_ptr(), pitch, tiles=tiles, rank=rank)
        compose()
        e1.record(stream)
        st = sc.sync()
        torch.cuda.synchronize(dev)
        return e0.elapsed_time(e1), st["rays"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for i in range(args.warmup):
        step_resident(i)
    barrier()
    launches0 = sc.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    t_wall = time.time()
    ms, rays = [], []
    for i in range(args.steps):
        m, r = step_resident(args.warmup + i)
        ms.append(m)
        rays.append(r)
    barrier()
    wall = time.time() - t_wall
    launches = sc.launch_count() - launches0

    # end to end through the host-facing call: camera in, frame out to pinned host memory, every step
    e2e_ms, e2e_rays = [], []
    if world == 1:
        for i in range(args.warmup + args.steps):
            cam, light = camera_for_step(sb, bmin, bmax, w["inside"], i)
            sc.set_lights([light] if w["lit"] else [])
            flush.fill_(i & 255)
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            _, st = sc.render(cam, W, H, out=host_np)
            dt = (time.perf_counter() - t0) * 1e3
            if i >= args.warmup:
                e2e_ms.append(dt)
                e2e_rays.append(st["rays"])
    clocks = sampler.summary()

    ms_t = torch.tensor(ms, dtype=torch.float64, device=dev)
    rays_t = torch.tensor(rays, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays_t, op=dist.ReduceOp.SUM)
    total_ms, total_rays = float(ms_t.sum()), float(rays_t.sum())
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return
    value = total_rays / (total_ms * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(total_ms / args.steps, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "reference scene assets (assets/), synthetic camera path",
        "config": {"workload": f"{w['name']} {W}x{H}: {w['desc']}", "rays_per_step": round(total_rays / args.steps),
                   "l2": "256 MiB buffer written between timed steps (L2 flush); scene itself is L2-sized",
                   "timing": "CUDA events on the scene stream around each step, summed; max over ranks",
                   "compose": ("nccl reduce(sum) of disjoint tiles to rank 0" if world > 1 else "none (1 GPU)"),
                   "tiles": "16x64 interleaved" if world > 1 else "full frame"},
        "gpu_launches": int(launches), "clocks": clocks, "wall_s": round(wall, 3),
    }
    if e2e_ms:
        ev = sum(e2e_rays) / (sum(e2e_ms) * 1e-3) / 1e6
        line["e2e"] = {"value": round(ev, 2), "unit": UNIT, "h2d_bytes_per_step": 52 + 36 * (1 if w["lit"] else 0),
                       "d2h_bytes_per_step": W * H * 3 + 64, "ms_per_step": round(sum(e2e_ms) / len(e2e_ms), 4),
                       "timing": "host wall clock around snl_render (camera in, BGR8 frame out to pinned host memory)"}
    # roofline of the dominant kernel (k_render): algorithmic bytes / measured launch time
    cam, light = camera_for_step(sb, bmin, bmax, w["inside"], args.warmup)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    if world == 1:
        abytes, cst = algorithmic_bytes(sb, w, W, H, cam, light)
        if abytes:
            ach = abytes / (ms[0] * 1e-3) / 1e9
            line["roofline"] = {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                                "frac": round(ach / peak, 4), "traffic": None,
                                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                                "note": "algorithmic bytes are served by L1/L2 (scene is L2-resident): frac is vs the HBM "
                                        "copy peak as the contract asks; the real bound is latency/issue, see DESIGN.md",
                                "algorithmic_bytes_per_launch": int(abytes),
                                "node_visits": cst["node_visits"], "tri_tests": cst["tri_tests"] + cst["gtri_tests"]}
    if world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        reference_frames(w, W, H, 1, threads)
        r1, s1, variant = reference_frames(w, W, H, 1, threads)
        frames = int(min(60, max(3, 12.0 / max(s1, 1e-3))))
        rr, ss, _ = reference_frames(w, W, H, frames, threads, first_step=args.warmup)
        line["cpu_baseline"] = {"value": round(rr / ss / 1e6, 3), "unit": UNIT, "cores": threads, "kind": "reference",
                                "sample": f"{frames} full {W}x{H} frames of the same workload through the reference's "
                                          f"Render() (oracle/_ref, veclib {variant}), {threads} threads"}
    print(json.dumps(line), flush=True)
    if dist:
        dist.destroy_process_group()


def main():
    args = parse_args()
    from snail_b200 import workloads
    w = workloads.resolve(args.workload)
    W, H = [int(x) for x in args.res.lower().split("x")]
    if args.impl == "reference":
        run_reference(args, w, W, H)
    else:
        run_ours(args, w, W, H)


if __name__ == "__main__":
    main()
