#!/usr/bin/env python
"""Device BVH build vs host SAH build on the bench scenes: build time, tree depth, and what each tree costs at render
time. Prints one JSON line per scene (run on the GPU box; results are copied into profiles/)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import snail_b200 as sb  # noqa: E402
from snail_b200 import workloads  # noqa: E402

W, H = 1920, 1080
for name in sys.argv[1:] or ["sponza_like_whitted", "feline_whitted", "barracks_whitted"]:
    w = workloads.resolve(name)
    row = {"workload": name}
    for label, flags in (("host_sah", sb.USE_SAH), ("device_lbvh", sb.BUILD_DEVICE), ("device_binned_sah", sb.BUILD_DEVICE | sb.BUILD_SAH_BINNED)):
        sc = sb.Scene(0)
        t0 = time.perf_counter()
        sc.load_obj(w["obj_path"], w["mtl_path"], w["tex_dir"], build_flags=flags)
        load_s = time.perf_counter() - t0
        cam, light = sb.bench_camera(*sc.bbox(), inside=w["inside"])
        sc.set_options(w["gvals"])
        sc.set_lights([light] if w["lit"] else [])
        for _ in range(3):
            sc.render(cam, W, H)
        ms = []
        for _ in range(10):
            _, st = sc.render(cam, W, H)
            ms.append(st["kernel_ms"])
        c = sc.counts()
        row[label] = {"load_s": round(load_s, 3), "device_build_ms": round(sc.build_ms(), 3), "tris": c["tris"],
                      "nodes": c["nodes"], "depth": c["depth"], "render_ms": round(sorted(ms)[len(ms) // 2], 4),
                      "rays": int(st["rays"])}
        sc.close()
    t = row["device_lbvh"]
    # HBM traffic of the build as accounted in snl_build.cu's header (bytes per triangle, sort = 8 passes r+w of 12 B)
    bytes_per_tri = 96 + 44 + 8 * 24 + 320 + 200
    t["algorithmic_GBps"] = round(bytes_per_tri * t["tris"] / (t["device_build_ms"] * 1e-3) / 1e9, 1) if t["device_build_ms"] else None
    for k in ("device_lbvh", "device_binned_sah"):
        row[k]["render_vs_host_tree"] = round(row[k]["render_ms"] / row["host_sah"]["render_ms"], 4)
    print(json.dumps(row), flush=True)
