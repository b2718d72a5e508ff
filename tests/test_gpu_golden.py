"""GPU parity against the COMMITTED golden fixtures (tests/golden/, generated from the reference by
tools/make_golden.py): needs neither /root/reference nor oracle/_ref at run time. Scenes are read by OUR host ingest
+ SAH builder, so this covers the whole product path: OBJ -> BVH -> flatten -> sm_100a kernels -> C ABI."""
import json
import os

import numpy as np
import pytest

from conftest import DATA, GOLDEN, ROOT

pytestmark = pytest.mark.gpu

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))
STAGES = {"simple_nolight": ({}, False), "depth": ({1: 1}, False), "simple_light": ({}, True),
          "full_light": ({6: 1}, True), "whitted": ({6: 1, 7: 1}, True),
          "whitted_ss": ({6: 1, 7: 1, 9: 1}, True), "whitted_3l": ({6: 1, 7: 1}, 3)}
T_REL = 1e-5


def find(sub, name):
    if name.startswith("@"):  # authored scene committed under data/ (tools/make_panes.py)
        return os.path.join(DATA, name[1:])
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        p = os.path.join(d, sub, name)
        if os.path.exists(p):
            return p
    pytest.skip(f"{sub}/{name}: scene assets not present (run __graft_entry__.build() where the reference is mounted)")


def load(sb, name):
    c = MANIFEST["cases"][name]
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    sc = sb.Scene(0)
    mtl = os.path.join(DATA, c["mtl"]) if c["mtl"] else None
    texdir = (os.path.dirname(find("data", "347.png")) + "/") if c["mtl"] else None
    sc.load_obj(find("scenes", c["obj"]), mtl, texdir)
    return c, g, sc


@pytest.mark.parametrize("name", list(MANIFEST["cases"]))
def test_primary_hits_vs_golden(sb, name):
    c, g, sc = load(sb, name)
    W, H = int(g["size"][0]), int(g["size"][1])
    cam = sb.Camera.from_array(g["camera"])
    tri, t, bary = sc.trace_primary(cam, W, H)
    same = tri == g["tri"]
    # ids: bit-exact except t ties / silhouette edges (a handful of pixels at most)
    diff = np.argwhere(~same)
    for y, x in diff:
        tg, tr = t[y, x], g["t"][y, x]
        tie = np.isfinite(tg) and np.isfinite(tr) and abs(tg - tr) <= 10 * T_REL * abs(tr)
        edge = any(min(abs(b[0]), abs(b[1]), abs(1 - b[0] - b[1])) < 1e-4
                   for i, b in ((tri[y, x], bary[y, x]), (g["tri"][y, x], g["bary"][y, x])) if i >= 0)
        assert tie or edge, f"pixel {x},{y}: gpu tri {tri[y, x]} t {tg}, golden tri {g['tri'][y, x]} t {tr}"
    assert len(diff) <= max(2, 1e-3 * W * H)
    hit = same & (g["tri"] >= 0)
    assert (np.abs(t[hit] - g["t"][hit]) <= T_REL * np.abs(g["t"][hit])).all()
    assert float((t[hit] == g["t"][hit]).mean()) >= 0.9999  # bit-identical with the scalar-veclib reference
    assert float((bary[hit] == g["bary"][hit]).mean()) >= 0.9999


@pytest.mark.parametrize("name", list(MANIFEST["cases"]))
def test_render_stages_vs_golden(sb, name):
    c, g, sc = load(sb, name)
    W, H = int(g["size"][0]), int(g["size"][1])
    cam = sb.Camera.from_array(g["camera"])
    lg = g["light"]
    light = sb.Light(lg[0:3], lg[3:6], lg[6])
    for stage, (gv, lit) in STAGES.items():
        sc.set_options(gv)
        sc.set_lights([sb.Light(l[0:3], l[3:6], l[6]) for l in g["lights3"]] if lit == 3 else [light] if lit else [])
        img, st = sc.render(cam, W, H)
        err = np.abs(img.astype(np.int16) - g["img_" + stage].astype(np.int16)).max(axis=2)
        undef = g["undef_" + stage] if stage != "depth" else g["undef_simple"]
        if stage != "depth":
            err[undef] = 0  # pixels the reference shades from uninitialised memory
        frac = float((err <= 1).mean())
        print(f"{name}/{stage}: within 1/255 on {frac*100:.3f}% (max err {int(err.max())}), rays {st['rays']} golden {c['rays'][stage]}")
        assert frac >= 0.999, stage
        assert abs(st["rays"] - c["rays"][stage]) <= max(4 * int(undef.sum()), 0.002 * c["rays"][stage])


@pytest.mark.parametrize("name", list(MANIFEST["cases"]))
def test_rays_and_shadow_vs_golden(sb, name):
    c, g, sc = load(sb, name)
    for mask, key in ((None, "ray"), (g["ray_mask"], "rayM")):
        tri, t, bary = sc.trace_rays(g["ray_o"], g["ray_d"], mask)
        same = tri == g[key + "_tri"]
        assert (~same).sum() <= 2
        hit = same & (tri >= 0)
        assert (np.abs(t[hit] - g[key + "_t"][hit]) <= T_REL * np.abs(g[key + "_t"][hit])).all()
    out = sc.trace_shadow(g["sh_origin"], g["ray_d"], g["sh_tmax"])
    assert (np.isneginf(out) != np.isneginf(g["sh_out"])).sum() <= 1
    lg = g["light"]
    sc.set_options({6: 1, 7: 1})
    sc.set_lights([sb.Light(lg[0:3], lg[3:6], lg[6])])
    rgb, st = sc.raytrace(g["ray_o"], g["ray_d"], None)
    err = np.abs(rgb - g["rt_rgb"]).max(axis=1)
    # Rays the REFERENCE shades from uninitialised memory are not comparable: a quad whose lane 0 missed never initialises
    # its Sample, and lanes of that quad hitting triangle 0 are skipped by the per-lane fix-up (src/scene_trace.cpp:240-290,
    # see tools/make_golden.py); in the golden that shows up as garbage or NaN (lancia ray 513, one NaN ray in panes).
    tri0 = g["ray_tri"]
    lane0 = np.repeat(tri0[: len(tri0) // 4 * 4].reshape(-1, 4)[:, 0], 4)
    undef = np.zeros(len(err), bool)
    undef[: len(lane0)] = (tri0[: len(lane0)] == 0) & (lane0 < 0)
    undef |= ~np.isfinite(g["rt_rgb"]).all(axis=1)
    bad = (~(err <= 1.0 / 255)) & ~undef
    print(f"{name}: snl_raytrace vs golden: {len(err)} rays, {int(undef.sum())} reference-undefined, {int(bad.sum())} of the rest differ by "
          f"more than 1/255 (at rays {np.nonzero(bad)[0][:8].tolist()}, worst {np.sort(np.nan_to_num(err[~undef]))[-3:]})")
    # the same UB one bounce deeper (a masked lane 0 in the reflection packet) cannot be told from the hit records: one ray
    # of 845 (0.12 %) is allowed for it; the live-oracle test on feline holds 100 % (tests/test_gpu_parity.py)
    assert int(bad.sum()) <= max(1, int(0.001 * len(err)))
