"""CPU tests pinning oracle/snail_oracle.c (plain-C per-ray restatement) against the committed goldens, which are
outputs of the reference itself (tools/make_golden.py). The scene is fed in the reference layouts produced by the host
ingest + builder (whose bit-identity with the reference's builder is pinned by tests/test_golden_cpu.py)."""
import json
import os

import numpy as np
import pytest

from conftest import DATA, GOLDEN, ROOT, need_oracle, scene_path

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))
STAGES = {"simple_nolight": ({}, False), "depth": ({1: 1}, False), "simple_light": ({}, True),
          "full_light": ({6: 1}, True), "whitted": ({6: 1, 7: 1}, True),
          "whitted_ss": ({6: 1, 7: 1, 9: 1}, True), "whitted_3l": ({6: 1, 7: 1}, 3)}


def find(sub, name):
    if name.startswith("@"):  # authored scene committed under data/ (tools/make_panes.py)
        return os.path.join(DATA, name[1:])
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        p = os.path.join(d, sub, name)
        if os.path.exists(p):
            return p
    pytest.skip(f"{sub}/{name}: scene assets not present")


def load(sb, name):
    import snail_oracle as O
    c = MANIFEST["cases"][name]
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    mtl = os.path.join(DATA, c["mtl"]) if c["mtl"] else None
    texdir = (os.path.dirname(find("data", "347.png")) + "/") if c["mtl"] else None
    hs = sb.HostScene(find("scenes", c["obj"]), mtl, texdir)
    nodes, tris, shtris, matids = hs.bvh()
    return c, g, O.OracleScene(nodes, tris, shtris, matids, hs.materials(), hs.textures())


@pytest.mark.parametrize("name", list(MANIFEST["cases"]))
def test_c_oracle_hits_vs_golden(sb, name):
    c, g, orc = load(sb, name)
    W, H = int(g["size"][0]), int(g["size"][1])
    tri, t, bary = orc.trace_primary(g["camera"], W, H)
    same = tri == g["tri"]
    # per-ray walk vs the reference's packet walk: ids may differ only on exact-t ties / silhouette edges
    assert (~same).sum() <= max(2, 1e-3 * W * H)
    for y, x in np.argwhere(~same):
        tie = np.isfinite(t[y, x]) and np.isfinite(g["t"][y, x]) and abs(t[y, x] - g["t"][y, x]) <= 1e-4 * abs(g["t"][y, x])
        edge = any(min(abs(b[0]), abs(b[1]), abs(1 - b[0] - b[1])) < 1e-4
                   for i, b in ((tri[y, x], bary[y, x]), (g["tri"][y, x], g["bary"][y, x])) if i >= 0)
        assert tie or edge
    hit = same & (tri >= 0)
    assert (t[hit] == g["t"][hit]).all() and (bary[hit] == g["bary"][hit]).all()  # bit-identical hit records
    for mask, key in ((None, "ray"), (g["ray_mask"], "rayM")):
        tri, t, bary = orc.trace_rays(g["ray_o"], g["ray_d"], mask)
        same = tri == g[key + "_tri"]
        assert (~same).sum() <= 2
        hit = same & (tri >= 0)
        assert (t[hit] == g[key + "_t"][hit]).all()
    out = orc.trace_shadow(g["sh_origin"], g["ray_d"], g["sh_tmax"])
    assert (np.isneginf(out) != np.isneginf(g["sh_out"])).sum() <= 1
    orc.close()


@pytest.mark.parametrize("name", list(MANIFEST["cases"]))
def test_c_oracle_frames_vs_golden(sb, name):
    c, g, orc = load(sb, name)
    W, H = int(g["size"][0]), int(g["size"][1])
    for stage, (gv, lit) in STAGES.items():
        orc.set_gvals(gv)
        orc.set_lights(list(g["lights3"]) if lit == 3 else [g["light"]] if lit else [])
        img, st = orc.render(g["camera"], W, H)
        err = np.abs(img.astype(np.int16) - g["img_" + stage].astype(np.int16)).max(axis=2)
        undef = g["undef_" + stage] if stage != "depth" else g["undef_simple"]
        if stage != "depth":
            err[undef] = 0
        frac = float((err <= 1).mean())
        assert frac >= 0.999, (stage, frac)
        assert abs(st["rays"] - c["rays"][stage]) <= max(4 * int(undef.sum()), 0.002 * c["rays"][stage]), stage
    orc.close()


@pytest.mark.parametrize("name,inside,W,H", [("box", True, 160, 96), ("lancia", False, 192, 128), ("feline", False, 160, 96)])
def test_two_bounces_restatement_vs_patched_reference(name, inside, W, H):
    """Reflection depth 2 (BASELINE config 5) is beyond the reference's cap; its comparison is the reference with
    `cache.reflections < 1` relaxed to `< 2` (oracle/_ref/libsnailref_depth2.so, oracle/build_ref.sh). The plain-C
    restatement with the same relaxation must reproduce that library's frames and ray counts, which pins the depth-2
    semantics (second bounce shaded and lit like the first, 0.3 blend at both levels) twice."""
    R = need_oracle("depth2")
    import snail_oracle as O
    ref = R.RefScene(scene_path(name), variant="depth2")
    orc = O.OracleScene(ref.nodes(), ref.tris(), ref.shtris() if ref.n_shtris else None, ref.matids())
    cam, light = R.bench_camera(*ref.bbox(), inside=inside)
    gv = {6: 1, 7: 1}
    ref.set_gvals(gv)
    ref.set_lights([light])
    img_r, st_r, _ = ref.render(cam, W, H, threads=1)
    orc.set_gvals(gv)
    orc.set_lights([light])
    orc.set_reflection_depth(1)
    img_1, c1 = orc.render(cam, W, H)
    orc.set_reflection_depth(2)
    undef = np.zeros((H, W), np.uint8)
    img_2, c2 = orc.render(cam, W, H, undef=undef)
    err = np.abs(img_2.astype(np.int16) - img_r.astype(np.int16)).max(axis=2)
    err[undef != 0] = 0
    print(f"{name}: two bounces, restatement vs patched reference: within 1/255 on {(err <= 1).mean()*100:.3f} %; rays {c2['rays']} vs "
          f"{st_r['rays']} (one bounce {c1['rays']})")
    assert (err <= 1).mean() >= 0.999
    # (the per-ray restatement casts a few shadow rays less than the packet code where a packet straddles the light's reach -
    # at either depth: 0.12 % on this lancia frame with ONE bounce too)
    assert abs(c2["rays"] - st_r["rays"]) <= max(4 * int((undef != 0).sum()), 0.005 * st_r["rays"])
    assert c2["reflection"] > c1["reflection"] and c2["primary"] == c1["primary"]
    orc.close()
    ref.close()
