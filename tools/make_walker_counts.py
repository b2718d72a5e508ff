#!/usr/bin/env python
"""TEST INFRASTRUCTURE — runs the oracle's instrumented single-ray walker (oracle/snail_oracle.c: near child first, early
out on lmin > t_best, one 32-byte BVH::Node per box test) over the bench workloads' frames and commits the counts next to
the goldens: tests/golden/walker_counts.json. bench.py reads that file for the ALGORITHMIC bytes of its roofline
(SURVEY.md 8(d)); it never executes the oracle on that path.

    python tools/make_walker_counts.py [workload@WxH ...]
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import snail_b200 as sb  # noqa: E402  (host ingest + builder only)
import snail_oracle as O  # noqa: E402
from snail_b200 import workloads  # noqa: E402
from bench import camera_for_step  # noqa: E402

DEFAULT = ["sponza_like_whitted@1920x1080", "sponza_like_shadow@1920x1080", "feline_whitted@1920x1080",
           "barracks_shadow@3840x2160", "lancia_whitted@3840x2160", "box_primary@800x600", "box_light@800x600"]
STEP = 3  # bench.py's default --warmup: the first timed step's camera
OUT = os.path.join(ROOT, "tests", "golden", "walker_counts.json")


def main():
    cases = sys.argv[1:] or DEFAULT
    table = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for case in cases:
        name, res = case.split("@")
        W, H = [int(x) for x in res.split("x")]
        w = workloads.resolve(name)
        hs = sb.HostScene(w["obj_path"], w["mtl_path"], w["tex_dir"])
        hn, ht, hsh, hm = hs.bvh()
        orc = O.OracleScene(hn, ht, hsh, hm, hs.materials(), hs.textures())
        cam, light = camera_for_step(sb, hn[0]["bmin"], hn[0]["bmax"], w["inside"], STEP)
        orc.set_gvals(w["gvals"])
        orc.set_lights([[*light.pos, *light.color, light.radius]] if w["lit"] else [])
        t0 = time.time()
        _, c = orc.render(cam.as_array(), W, H)
        # shaded hits: one per reflection ray cast (= primary hits) plus at most as many at the bounce; without the bounce
        # the shadow rays cast bound the primary hits from below
        shaded = 2 * c["reflection"] if w["gvals"].get(7) else c["shadow"]
        table[case] = {"box_tests": c["node_visits"], "tri_tests": c["tri_tests"], "rays": c["rays"], "primary": c["primary"],
                       "shadow": c["shadow"], "reflection": c["reflection"], "reflection_shadow": c["reflection_shadow"],
                       "shaded_hits": shaded, "camera_step": STEP,
                       "source": "oracle single-ray walker (oracle/snail_oracle.c via tools/make_walker_counts.py), committed as "
                                 "tests/golden/walker_counts.json"}
        print(case, table[case], f"{time.time() - t0:.1f} s", flush=True)
        orc.close()
    with open(OUT, "w") as f:
        json.dump(table, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
