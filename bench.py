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
DEFAULT_WORKLOAD = "sponza_like_whitted"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--res", default="1920x1080")
    ap.add_argument("--compose", default="p2p", choices=["p2p", "nccl"])
    ap.add_argument("--sustain", type=float, default=1.0, help="seconds of back-to-back replay for clock sampling")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--refl-depth", type=int, default=1, choices=[1, 2],
                    help="reflection bounces (2 = BASELINE config 5; beyond the reference, whose arm then runs its depth-2 patched copy)")
    ap.add_argument("--tree", default="host", choices=["host", "lbvh", "sah"],
                    help="which tree the GPU arm walks: host = the reference's SAH sweep tree (bit-identical port, default), "
                         "lbvh / sah = built on the device (Morton LBVH / binned SAH, snl_build*.cu)")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--ncu-child", action="store_true", help="internal: the frames ncu profiles for the roofline")
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


_REF_CACHE = {}


REFL_DEPTH = 1


def reference_frames(w, W, H, frames, threads, first_step=0):
    """Times the reference's own Render() (SSE packet tracer, oracle/_ref) on the host cores. Returns
    (total_rays, total_seconds, variant, cpu_utilisation = process CPU time / wall time).

    The reference's pthread pool (src/thread_pool.cpp) can lose the wake-up of freshly created workers, after which
    a frame is traced by fewer threads than asked for; the pool is therefore warmed with a 1,2,4,..,threads ramp and
    the achieved CPU utilisation is reported next to the number."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import snailref as R
    import snail_b200 as sb
    variant = "sse" if R.available("sse") else "scalar"
    if REFL_DEPTH == 2:  # oracle/_ref/libsnailref_depth2.so: the scalar build with `cache.reflections < 2` (oracle/build_ref.sh)
        variant = "depth2"
    key = (w["name"], variant)
    if key not in _REF_CACHE:
        ref = R.RefScene(w["obj_path"], w["mtl_path"], w["tex_dir"], variant=variant)
        ref.set_gvals(w["gvals"])
        bmin, bmax = ref.bbox()
        cam, light = camera_for_step(sb, bmin, bmax, w["inside"], 0)
        ref.set_lights([[*light.pos, *light.color, light.radius]] if w["lit"] else [])
        t = 1
        while t < threads:
            ref.render(cam.as_array(), W, H, threads=t)
            t *= 2
        _REF_CACHE[key] = ref
    ref = _REF_CACHE[key]
    bmin, bmax = ref.bbox()
    rays, secs = 0, 0.0
    c0, t0 = os.times(), time.time()
    for i in range(frames):
        cam, light = camera_for_step(sb, bmin, bmax, w["inside"], first_step + i)
        ref.set_lights([[*light.pos, *light.color, light.radius]] if w["lit"] else [])
        _, st, ms = ref.render(cam.as_array(), W, H, threads=threads)
        rays += st["rays"]
        secs += ms * 1e-3
    c1, t1 = os.times(), time.time()
    util = ((c1.user - c0.user) + (c1.system - c0.system)) / max(t1 - t0, 1e-9)
    return rays, secs, variant, util


def run_reference(args, w, W, H):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    reference_frames(w, W, H, max(1, args.warmup), threads)  # warm-up (page in, thread pool spin-up)
    t0 = time.time()
    rays, secs, variant, util = reference_frames(w, W, H, args.steps, threads, first_step=args.warmup)
    wall = time.time() - t0
    val = rays / secs / 1e6
    sample = f"{args.steps} full {W}x{H} frames of the workload, Render() wall time only"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(val, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(secs / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "reference scene assets",
        "config": {"workload": f"{w['name']} {W}x{H}: {w['desc']}", "backend": f"oracle/_ref veclib {variant}"},
        "cpu_baseline": {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample,
                         "cpu_utilisation": round(util, 2)},
        "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "wall_s": round(wall, 2),
    }
    print(json.dumps(line), file=JSON_OUT, flush=True)


NCU_METRICS = ["smsp__inst_executed.sum", "smsp__thread_inst_executed.sum", "l1tex__t_bytes.sum",
               "l1tex__data_pipe_lsu_wavefronts.sum", "lts__t_bytes.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
               "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
               "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "gpu__time_duration.sum"]


def walker_counts(name, W, H):
    """Nodes visited / triangles tested by a SINGLE-RAY walk of the reference's tree (near child first, early out on
    lmin > t_best) for this workload's frame: the oracle's instrumented walker, run once where the oracle lives and
    committed next to the goldens (tools/make_walker_counts.py -> tests/golden/walker_counts.json). Data, not code: the
    bench never executes the oracle on this path."""
    try:
        t = json.load(open(os.path.join(ROOT, "tests", "golden", "walker_counts.json")))
        return t.get(f"{name}@{W}x{H}")
    except Exception:
        return None


def counters_build_counts(sb, w, W, H, cam, light, args=None):
    """Fallback for workloads without committed walker counts: the same kernels built with -DSNL_COUNTERS."""
    if not os.path.exists(sb.api.LIB_COUNTERS_PATH):
        return None
    sc = sb.Scene(0, counters=True)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"], build_flags=tree_flags(sb, args.tree if args else "host"))
    sc.set_options(w["gvals"])
    sc.set_reflection_depth(args.refl_depth if args else 1)
    sc.set_lights([light] if w["lit"] else [])
    _, st = sc.render(cam, W, H)
    sc.close()
    return {"box_tests": 2 * st["node_visits"], "tri_tests": st["tri_tests"] + st["gtri_tests"], "rays": st["rays"],
            "shaded_hits": st["reflection"] + st["reflection"], "source": "snail_b200 kernels built with -DSNL_COUNTERS"}


def tree_flags(sb, tree):
    """--tree: the reference's SAH sweep tree from the host port (default) or one of the two device builders."""
    return {"host": sb.USE_SAH, "lbvh": sb.BUILD_DEVICE, "sah": sb.BUILD_DEVICE | sb.BUILD_SAH_BINNED}[tree]


def ncu_child(args, w, W, H):
    """`bench.py --ncu-child` (run under ncu by ncu_pass): three device-resident frames of the workload at the bench
    camera with the kept register-budget variant pinned; ncu profiles the third k_render launch."""
    import snail_b200 as sb
    sc = sb.Scene(0)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"], build_flags=tree_flags(sb, args.tree))
    sc.set_options(w["gvals"])
    sc.set_reflection_depth(args.refl_depth)
    bmin, bmax = sc.bbox()
    frame, _ = sc.frame_create(W, H)
    for i in range(3):
        cam, light = camera_for_step(sb, bmin, bmax, w["inside"], args.warmup)
        sc.set_lights([light] if w["lit"] else [])
        sc.render_dev(cam, W, H, frame, W * 3)
        sc.sync()
    sc.close()


def ncu_pass(args, W, H, variant):
    """Counters of ONE k_render launch of this workload, collected by ncu in a child process after the timed region
    (nothing here is a timing: instruction / byte / wavefront counts and utilisation ratios only)."""
    import csv
    import shutil
    ncu = shutil.which("ncu") or "/usr/local/cuda/bin/ncu"
    if not os.path.exists(ncu):
        return None, "ncu not found"
    env = dict(os.environ, SNL_RENDER_VARIANT=str(max(0, variant)), CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "0"))
    cmd = [ncu, "--metrics", ",".join(NCU_METRICS), "--clock-control", "none", "-k", "regex:k_render", "--launch-skip", "2",
           "--launch-count", "1", "--csv", sys.executable, os.path.abspath(__file__), "--ncu-child", "--workload", args.workload,
           "--res", args.res, "--warmup", str(args.warmup), "--refl-depth", str(args.refl_depth), "--tree", args.tree]
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env)
    except Exception as e:  # noqa: BLE001
        return None, f"ncu failed: {e}"
    rows = [row for row in csv.reader(r.stdout.splitlines()) if len(row) > 5]
    hdr = next((row for row in rows if "Metric Name" in row), None)
    if hdr is None:
        return None, "ncu gave no metrics: " + (r.stderr or r.stdout)[-300:].replace("\n", " ")
    iN, iV, iK = hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Kernel Name")
    out = {"kernel": None}
    for row in rows:
        if row is hdr or len(row) <= max(iN, iV):
            continue
        try:
            out[row[iN]] = float(row[iV].replace(",", ""))
            out["kernel"] = row[iK]
        except ValueError:
            pass
    return (out, None) if "smsp__inst_executed.sum" in out else (None, "ncu output lacks smsp__inst_executed.sum")


def build_roofline(sb, args, w, W, H, sc, ms_per_launch, clocks, rank):
    """What bounds k_render (the scene is L2-resident, the walk is served by L1): instruction issue first, then L1
    wavefronts; HBM is idle. Peaks are measured in this run (snl_measure_peaks), counts come from an ncu pass of the same
    frame, the algorithmic figure from the oracle's single-ray walker."""
    bmin, bmax = sc.bbox()
    cam, light = camera_for_step(sb, bmin, bmax, w["inside"], args.warmup)
    t = ms_per_launch * 1e-3
    peaks = sb.measure_peaks(rank)
    driver = {}
    try:
        driver = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(driver.get("hbm_gbs", 6650.0))
    n, err = ncu_pass(args, W, H, sc.render_variant())
    wc = walker_counts(w["name"], W, H) if args.tree == "host" and args.refl_depth == 1 else None  # committed counts: the reference's tree, one bounce
    wc = wc or counters_build_counts(sb, w, W, H, cam, light, args)
    rf = {"bound": "issue", "unit": "Gwarp-inst/s", "kernel": "k_render",
          "peak": round(peaks["issue_peak_gwarpinst_s"], 1),
          "peak_source": f"{peaks['sms']} SMs x 4 schedulers x {peaks['sm_clock_mhz_max']:.0f} MHz; an FFMA / integer-ALU interleave "
                         f"measured {peaks['issue_gwarpinst_s']:.1f} in this run (snl_measure_peaks, builder-measured)",
          "launch_ms": round(ms_per_launch, 4)}
    if n:
        inst, tinst = n["smsp__inst_executed.sum"], n.get("smsp__thread_inst_executed.sum", 0.0)
        dram = n.get("dram__bytes_read.sum", 0.0) + n.get("dram__bytes_write.sum", 0.0)
        lanes = tinst / inst if inst else 0.0
        ach = inst / t / 1e9
        sm_clk = (clocks or {}).get("sm_mhz") or peaks["sm_clock_mhz_max"]
        wave = n.get("l1tex__data_pipe_lsu_wavefronts.sum", 0.0)
        rf.update({
            "achieved": round(ach, 1), "frac": round(ach / peaks["issue_peak_gwarpinst_s"], 4),
            "warp_instructions": int(inst), "active_lanes": round(lanes, 2),
            "useful_frac": round(ach / peaks["issue_peak_gwarpinst_s"] * lanes / 32.0, 4),
            "traffic": int(dram), "kernel_profiled": n.get("kernel"), "registers": int(n.get("launch__registers_per_thread", 0)),
            "l1": {"wavefronts": int(wave), "frac": round(wave / (peaks["sms"] * sm_clk * 1e6 * t), 4),
                   "frac_is": "LSU data-pipe wavefronts / (SMs x SM clock x launch time): one wavefront per SM and cycle",
                   "bytes": int(n.get("l1tex__t_bytes.sum", 0)), "achieved_gbs": round(n.get("l1tex__t_bytes.sum", 0) / t / 1e9, 1),
                   "peak_gbs": round(peaks["l1_gbs"], 1), "ncu_throughput_pct": n.get("l1tex__throughput.avg.pct_of_peak_sustained_elapsed")},
            "l2": {"bytes": int(n.get("lts__t_bytes.sum", 0)), "achieved_gbs": round(n.get("lts__t_bytes.sum", 0) / t / 1e9, 1),
                   "peak_gbs": round(peaks["l2_gbs"], 1), "frac": round(n.get("lts__t_bytes.sum", 0) / t / 1e9 / peaks["l2_gbs"], 4),
                   "ncu_throughput_pct": n.get("lts__throughput.avg.pct_of_peak_sustained_elapsed")},
            "hbm": {"bytes": int(dram), "achieved_gbs": round(dram / t / 1e9, 1), "peak_gbs": hbm_peak,
                    "dram_frac": round(dram / t / 1e9 / hbm_peak, 4),
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (driver-measured copy)" if driver else "6650 GB/s fallback"},
            "ncu_issue_active_pct": n.get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "counts_source": "ncu pass of this bench run: same workload, camera and kept k_render variant, one launch, child process, "
                             "after the timed region (counts and ratios only; launch_ms is the un-profiled CUDA-event time)"})
    else:
        rf.update({"achieved": None, "frac": None, "traffic": None, "counts_source": err})
    # the roofline headline is the resource that binds: whichever of instruction issue and L1 wavefronts runs closer to
    # its ceiling (both are reported); HBM and L2 are far from theirs for these L2-resident scenes
    if n and rf.get("l1", {}).get("frac") and rf["l1"]["frac"] > rf["frac"]:
        rf["issue"] = {"achieved": rf["achieved"], "peak": rf["peak"], "unit": rf["unit"], "frac": rf["frac"], "useful_frac": rf["useful_frac"]}
        wps = rf["l1"]["wavefronts"] / t / 1e9
        rf.update({"bound": "l1", "unit": "Gwavefront/s", "achieved": round(wps, 1), "frac": rf["l1"]["frac"],
                   "peak": round(wps / rf["l1"]["frac"], 1),
                   "peak_source": "one L1 (LSU data pipe) wavefront per SM and cycle: SMs x SM clock under load; coalesced loads reach "
                                  f"{peaks['l1_gbs']:.0f} GB/s = {peaks['l1_gbs'] / (peaks['sms'] * sm_clk * 1e6 * 128 / 1e9):.2f} of it in "
                                  "snl_measure_peaks (builder-measured)"})
    if wc:
        ab = 32 * wc["box_tests"] + 64 * wc["tri_tests"] + 16 * wc["rays"] + 64 * wc["shaded_hits"] + 3 * W * H
        rf["algorithmic"] = {"bytes_per_launch": int(ab), "gbs": round(ab / t / 1e9, 1), "box_tests": wc["box_tests"],
                             "tri_tests": wc["tri_tests"], "rays": wc["rays"], "source": wc.get("source"),
                             "formula": "32 B x boxes tested (BVH::Node) + 64 B x triangles tested + 16 B hit record per ray + 64 B "
                                        "ShTriangle per shaded hit + 3 B per pixel (SURVEY 8d)",
                             "vs_l1_peak": round(ab / t / 1e9 / peaks["l1_gbs"], 4), "vs_hbm_peak": round(ab / t / 1e9 / hbm_peak, 4)}
    rf["peaks_builder_measured"] = {k: round(v, 1) if isinstance(v, float) else v for k, v in peaks.items()}
    return rf


def run_ours(args, w, W, H):
    import torch
    import snail_b200 as sb
    from snail_b200 import farm
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    dist = None
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        import datetime
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))

    sc = sb.Scene(local_rank)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"], build_flags=tree_flags(sb, args.tree))
    sc.set_options(w["gvals"])
    sc.set_reflection_depth(args.refl_depth)
    bmin, bmax = sc.bbox()
    stream = torch.cuda.ExternalStream(sc.stream(), device=dev)
    pitch = W * 3
    # N > 1: farm.Farm = bootstrap of snl_comm_* / snl_render_multi ("p2p": static deal, peer stores into rank 0's
    # frame, in-stream completion, banded device->host copy) or the static tile split + NCCL gather ("nccl")
    fm = farm.Farm(sc, dist, W, H, compose="p2p" if args.compose == "p2p" else "gather") if world > 1 else None
    frame = sc.frame_create(W, H)[0] if world == 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(1)  # > 126 MB L2
    flush64 = flush.view(torch.int64)
    host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
    host_np = host_img.numpy()

    def l2_flush():
        # READ 256 MiB: the scene leaves L2 and only clean lines stay behind (a written buffer would be evicted as
        # dirty lines during the timed kernel and show up in its DRAM traffic)
        return flush64.sum()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    steps_cam = {}

    def set_step(i):  # cameras / lights of all steps are made once: no numpy work inside the timed loops
        if i not in steps_cam:
            steps_cam[i] = camera_for_step(sb, bmin, bmax, w["inside"], i)
        cam, light = steps_cam[i]
        sc.set_lights([light] if w["lit"] else [])
        return cam

    for i in range(args.warmup + args.steps):
        set_step(i)

    def step_resident(i):
        cam = set_step(i)
        with torch.cuda.stream(stream):
            l2_flush()  # in-stream, before the step's first event: not part of the timed region
        if fm is None:
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            sc.render_dev(cam, W, H, frame, pitch, rank=rank)
            e1.record(stream)
            st = sc.sync()
            torch.cuda.synchronize(dev)
            return e0.elapsed_time(e1), st["rays"]
        # collective, no host barrier per step: the ranks meet on the device (rank 0 opens the frame slot, every rank's
        # packets go into its frame). The library times every rank on the device with CUDA events on its stream, from the
        # moment the frame slot is open to the rank's last kernel (per-origin records, its packets, its publish): a rank
        # that is early for the slot does not count the wait. The K steps as a whole are bracketed by barrier + synchronize.
        if fm.comm is None:  # "nccl" compose: static split + torch.distributed gather, timed around the whole call
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            with torch.cuda.stream(stream):
                _, st = fm.render(cam, host=False)
            e1.record(stream)
            torch.cuda.synchronize(dev)
            return e0.elapsed_time(e1), st["rays"]
        _, st = fm.render(cam, host=False)
        return st["kernel_ms"], st["rays"]

    def step_e2e(i):
        """The host-facing call: camera in, BGR8 frame out to pinned host memory on rank 0."""
        cam = set_step(i)
        with torch.cuda.stream(stream):
            l2_flush()
        if fm is None:
            barrier()
        else:
            torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        if fm is None:
            _, st = sc.render(cam, W, H, out=host_np)
        else:
            _, st = fm.render(cam, out=host_np)
        return (time.perf_counter() - t0) * 1e3, st["rays"]

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

    e2e_ms, e2e_rays = [], []
    for i in range(args.warmup + args.steps):
        if i == args.warmup:
            barrier()
        m, r = step_e2e(i)
        if i >= args.warmup:
            e2e_ms.append(m)
            e2e_rays.append(r)
    barrier()

    # back-to-back replay of the same step (no flush, no per-step sync) for ~1 s: keeps the GPU under load long
    # enough for nvidia-smi to sample clocks / throttle reasons, and gives the streaming (pipelined) frame rate
    sus_frames, sus_rays1 = 0, 0
    cam = set_step(args.warmup)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t0 = time.time()
    mean_ms = torch.tensor([sum(ms) / len(ms)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mean_ms, op=dist.ReduceOp.MAX)  # the SAME frame count on every rank: the render call is collective
    n_sus = max(1, int(args.sustain / max(1e-4, float(mean_ms[0]) * 1e-3)))
    for k in range(n_sus):
        if fm is None:
            sc.render_dev(cam, W, H, frame, pitch, rank=rank)
            if k % 50 == 49 or k == n_sus - 1:
                sus_rays1 = sc.sync()["rays"]
        else:
            sus_rays1 = fm.render(cam, host=False)[1]["rays"]
        sus_frames += 1
    e1.record(stream)
    torch.cuda.synchronize(dev)
    sus_ms = e0.elapsed_time(e1)
    clocks = sampler.summary()
    clocks["window"] = "timed steps + e2e steps + %.1f s back-to-back replay of the step" % args.sustain
    # rank 0 holds the frame's summed TreeStats (snl_render_multi sums them in-stream); the gather compose all-reduces them
    # the K steps of a rank are timed back to back (sum of its per-step device times); the job's time is the slowest rank's
    # (the host-facing number is rank 0's: its call takes the camera and returns the frame)
    ms_t = torch.tensor([sum(ms), sum(e2e_ms) if rank == 0 else 0.0, sus_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    K = args.steps
    total_ms, total_rays = float(ms_t[0]), float(sum(rays))
    e2e_total_ms, e2e_total_rays = float(ms_t[1]), float(sum(e2e_rays))
    sus_value = float(sus_rays1) * sus_frames / (float(ms_t[2]) * 1e-3) / 1e6
    if rank != 0:
        if fm:
            fm.close()
        sc.close()
        if dist:
            dist.destroy_process_group()
        return
    value = total_rays / (total_ms * 1e-3) / 1e6
    compose = {1: "none (1 GPU)"}.get(world, "p2p (snl_render_multi): packets dealt to the ranks like the reference's taskMap; every rank's k_render stores its share "
                                             "into rank 0's frame through CUDA-IPC peer memory (NVLink), one launch + one system-scope signal per row band, "
                                             "in-stream completion"
                                      if args.compose == "p2p" else "static 16x64 tile split + NCCL gather of packed tiles to rank 0")
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": args.warmup, "ms_per_step": round(total_ms / K, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "scene assets under assets/ (reference scenes; sponza_like is procedurally generated), synthetic camera path",
        "config": {"workload": f"{w['name']} {W}x{H}: {w['desc']}", "rays_per_step": round(total_rays / K),
                   "l2": "256 MiB buffer READ between timed steps (L2 flush that leaves no dirty lines)",
                   "timing": ("CUDA events on the scene stream around each step (barrier + synchronize on both sides of every step)" if world == 1 else
                              "per rank: CUDA events on its stream from 'frame slot open' to its last kernel of the step (snl_stats.kernel_ms), "
                              "summed over the K steps, max over ranks; the K steps are bracketed by barrier + synchronize, no host barrier in between"),
                   "compose": compose},
        "gpu_launches": int(launches), "clocks": clocks, "wall_s": round(wall, 3),
        "sustained": {"value": round(sus_value, 2), "unit": UNIT, "frames": sus_frames,
                      "note": "same step replayed back to back, no L2 flush, camera fixed"},
    }
    ev = e2e_total_rays / (e2e_total_ms * 1e-3) / 1e6
    line["e2e"] = {"value": round(ev, 2), "unit": UNIT, "h2d_bytes_per_step": (52 + 36 * (1 if w["lit"] else 0)) * world,
                   "d2h_bytes_per_step": W * H * 3 + 72, "ms_per_step": round(e2e_total_ms / K, 4),
                   "timing": "host wall clock around the host-facing render call (camera + light in, BGR8 frame out to "
                             "pinned host memory on rank 0" + ("; rank 0's snl_render_multi call, frames back to back, no host collective inside)" if world > 1 else ")")}
    if world == 1 and not args.no_extras:
        try:
            line["roofline"] = build_roofline(sb, args, w, W, H, sc, total_ms / K, clocks, local_rank)
        except Exception as e:  # noqa: BLE001
            line["roofline"] = {"bound": "issue", "achieved": None, "frac": None, "traffic": None, "error": str(e)[:300]}
    if world == 1 and not args.no_cpu_baseline:
        threads = host_threads()
        reference_frames(w, W, H, 1, threads)
        r1, s1, variant, _ = reference_frames(w, W, H, 1, threads)
        frames = int(min(60, max(3, 12.0 / max(s1, 1e-3))))
        rr, ss, _, util = reference_frames(w, W, H, frames, threads, first_step=args.warmup)
        line["cpu_baseline"] = {"value": round(rr / ss / 1e6, 3), "unit": UNIT, "cores": threads, "kind": "reference",
                                "cpu_utilisation": round(util, 2),
                                "sample": f"{frames} full {W}x{H} frames of the same workload through the reference's "
                                          f"Render() (oracle/_ref, veclib {variant}), {threads} threads"}
    print(json.dumps(line), file=JSON_OUT, flush=True)
    if fm:
        fm.close()
    sc.close()
    if dist:
        dist.destroy_process_group()


JSON_OUT = sys.stdout


def main():
    # stdout carries ONE JSON line: anything native code prints there (the reference's loader reports "Done loading"
    # through C stdio, flushed whenever) is sent to stderr instead
    global JSON_OUT
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    args = parse_args()
    from snail_b200 import workloads
    w = workloads.resolve(args.workload)
    W, H = [int(x) for x in args.res.lower().split("x")]
    global REFL_DEPTH
    REFL_DEPTH = args.refl_depth
    if args.refl_depth == 2:
        w["desc"] += "; TWO reflection bounces (beyond the reference's cap of one)"
    if args.tree != "host":
        w["desc"] += f"; tree built on the device ({'Morton LBVH' if args.tree == 'lbvh' else 'binned SAH'}) instead of the reference's"
    if args.ncu_child:
        ncu_child(args, w, W, H)
    elif args.impl == "reference":
        run_reference(args, w, W, H)
    else:
        run_ours(args, w, W, H)


if __name__ == "__main__":
    main()
