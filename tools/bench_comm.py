#!/usr/bin/env python
"""GPU box, 1 GPU: snl_render_multi (world = 1: frame protocol, per-band launches + signals, banded copy) against the
single launch of snl_render_dev on the same frames - what the multi-GPU machinery costs by itself. (Written for the
cursor-fed persistent k_render of profiles/r2f / r2g, which lost and was removed.)
    python tools/bench_comm.py [workload res]..."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import snail_b200 as sb  # noqa: E402
from snail_b200 import workloads  # noqa: E402

cases = sys.argv[1:] or ["sponza_like_whitted", "1920x1080", "barracks_shadow", "3840x2160", "lancia_whitted", "3840x2160"]
for name, res in zip(cases[0::2], cases[1::2]):
    W, H = [int(x) for x in res.split("x")]
    w = workloads.resolve(name)
    sc = sb.Scene(0)
    sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"])
    sc.set_options(w["gvals"])
    cam, light = sb.bench_camera(*sc.bbox(), inside=w["inside"])
    sc.set_lights([light] if w["lit"] else [])
    frame, _ = sc.frame_create(W, H)
    host = np.zeros((H, W, 3), np.uint8)
    for _ in range(4):
        sc.render_dev(cam, W, H, frame, W * 3)
        st = sc.sync()
    ms = []
    for _ in range(20):
        sc.render_dev(cam, W, H, frame, W * 3)
        ms.append(sc.sync()["kernel_ms"])
    t0 = time.perf_counter()
    for _ in range(20):
        sc.render(cam, W, H, out=host)
    e2e_static = (time.perf_counter() - t0) / 20 * 1e3
    comm = sb.Comm.local([sc], W, H)
    for _ in range(4):
        comm.render(cam, host=False)
    md = []
    t0 = time.perf_counter()
    for _ in range(20):
        md.append(comm.render(cam, host=False)[1]["kernel_ms"])
    wall_dyn = (time.perf_counter() - t0) / 20 * 1e3
    t0 = time.perf_counter()
    for _ in range(20):
        comm.render(cam, out=host)
    e2e_dyn = (time.perf_counter() - t0) / 20 * 1e3
    print(f"{name} {res} batch={os.environ.get('SNL_DYN_BATCH', 'auto')}: static kernel {np.median(ms):.4f} ms, e2e {e2e_static:.4f} ms | "
          f"dyn kernel {np.median(md):.4f} ms, device-resident call {wall_dyn:.4f} ms, e2e {e2e_dyn:.4f} ms", flush=True)
    comm.close()
    sc.close()
