#!/usr/bin/env python
"""TEST INFRASTRUCTURE (GPU box): where does the tail of snl_raytrace vs the reference's Scene::RayTrace come from?
Runs the packets of tests/test_gpu_parity.py::test_raytrace_packets_vs_oracle stage by stage (shading mode, lights,
bounce) and classifies the rays that differ by more than 1/255."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import snail_b200 as sb  # noqa: E402
import snailref as R  # noqa: E402
from test_gpu_parity import _random_rays  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "feline"
path = os.path.join(ROOT, "oracle", "_ref", "scenes", name + ".obj")
ref = R.RefScene(path, variant="scalar")
sc = sb.Scene(0)
sc.set_geometry(ref.nodes(), ref.tris(), ref.shtris() if ref.n_shtris else None, ref.matids())
bmin, bmax = ref.bbox()
cam, light = sb.bench_camera(bmin, bmax)
n = 256 * 9 + 50
o, d = _random_rays(n, bmin, bmax, 21)
mask = (np.arange(n) % 3 != 0).astype(np.uint8)
tri_r, t_r, b_r = ref.trace_rays(o, d, None)
tri_g, t_g, b_g = sc.trace_rays(o, d, None)
print("primary hits: ids equal", (tri_r == tri_g).mean(), "hit fraction", (tri_r >= 0).mean())
for gv in ({}, {6: 1}, {7: 1}, {6: 1, 7: 1}):
    for lit in (False, True):
        ref.set_gvals(gv)
        ref.set_lights([[*light.pos, *light.color, light.radius]] if lit else [])
        sc.set_options(gv)
        sc.set_lights([light] if lit else [])
        for m, shared in ((None, False), (mask, False), (None, True)):
            oo = o[0] if shared else o
            rgb_r, rays_r = ref.raytrace(oo, d, m, shared_origin=shared)
            rgb_g, st = sc.raytrace(oo, d, m, shared_origin=shared)
            err = np.abs(rgb_g - rgb_r).max(axis=1)
            bad = np.nonzero(err > 1.0 / 255)[0]
            print(f"gv={gv} lit={lit} mask={m is not None} shared={shared}: within 1/255 {100*(err <= 1/255).mean():.3f}% "
                  f"exact {100*(err == 0).mean():.3f}% rays gpu {st['rays']} ref {rays_r} bad {len(bad)}")
            if len(bad) and gv == {6: 1, 7: 1} and lit:
                q = bad // 4
                lane = bad % 4
                blk = bad // 16
                print("   bad rays by lane-in-quad:", np.bincount(lane, minlength=4).tolist(),
                      " distinct quads", len(np.unique(q)), " distinct 16-blocks", len(np.unique(blk)),
                      " distinct packets", len(np.unique(bad // 256)))
                for i in bad[:12]:
                    print(f"   ray {i}: gpu {rgb_g[i]} ref {rgb_r[i]} primary tri g/r {tri_g[i] if not shared else '-'} {tri_r[i] if not shared else '-'}")
