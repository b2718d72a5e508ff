"""CPU tests: (1) the oracle (the reference compiled in place, oracle/_ref) reproduces the committed golden
fixtures; (2) OUR host ingest + SAH builder (snl_host_scene_load_obj, no device needed) produces the reference
builder's BVH bit for bit — pinned both by sha256 digests in tests/golden/manifest.json and, when the oracle is
built, by a direct comparison."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import DATA, GOLDEN, TEXDIR, need_oracle, scene_path

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))
CASES = list(MANIFEST["cases"])
STAGES = {"simple_nolight": ({}, False), "depth": ({1: 1}, False), "simple_light": ({}, True),
          "full_light": ({6: 1}, True), "whitted": ({6: 1, 7: 1}, True),
          "whitted_ss": ({6: 1, 7: 1, 9: 1}, True), "whitted_3l": ({6: 1, 7: 1}, 3)}


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def case_paths(name):
    c = MANIFEST["cases"][name]
    obj = os.path.join(DATA, c["obj"][1:]) if c["obj"].startswith("@") else scene_path(c["obj"][:-4])
    mtl = os.path.join(DATA, c["mtl"]) if c["mtl"] else None
    return c, obj, mtl, (TEXDIR + "/") if c["mtl"] else ""


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_goldens(name):
    R = need_oracle("scalar")
    c, obj, mtl, texdir = case_paths(name)
    g = golden(name)
    ref = R.RefScene(obj, mtl, texdir, variant="scalar")
    W, H = int(g["size"][0]), int(g["size"][1])
    ref.set_gvals({})
    ref.set_lights([])
    tri, t, bary = ref.trace_primary(g["camera"], W, H)
    assert (tri == g["tri"]).all() and (t == g["t"]).all() and (bary == g["bary"]).all()
    assert (ref.primary_dirs(g["camera"], W, H) == g["dirs"]).all()
    for stage, (gv, lit) in STAGES.items():
        ref.set_gvals(gv)
        ref.set_lights(list(g["lights3"]) if lit == 3 else [g["light"]] if lit else [])
        img, st, _ = ref.render(g["camera"], W, H, threads=2)
        ok = (img == g["img_" + stage]).all(axis=2)
        undef = g["undef_" + stage] if stage != "depth" else g["undef_simple"]
        if stage != "depth":
            ok |= undef  # uninitialised-normal pixels of the reference, see tools/make_golden.py
        assert ok.all(), stage
        slack = 4 * int(undef.sum()) if stage != "depth" else 0  # their shadow rays are garbage too
        assert abs(st["rays"] - c["rays"][stage]) <= slack
    tri, t, bary = ref.trace_rays(g["ray_o"], g["ray_d"], g["ray_mask"])
    assert (tri == g["rayM_tri"]).all() and (t == g["rayM_t"]).all()
    out = ref.trace_shadow(g["sh_origin"], g["ray_d"], g["sh_tmax"])
    assert (out == g["sh_out"]).all()
    ref.close()


def test_sse_oracle_agrees_with_scalar_goldens_within_tolerance():
    """The timed baseline (rcpps/rsqrtps + 1 Newton step) against the canonical scalar goldens: same ids except at
    ties/edges, |dt|/t <= 1e-4."""
    R = need_oracle("sse")
    c, obj, mtl, texdir = case_paths("feline")
    g = golden("feline")
    ref = R.RefScene(obj, mtl, texdir, variant="sse")
    W, H = int(g["size"][0]), int(g["size"][1])
    tri, t, _ = ref.trace_primary(g["camera"], W, H)
    same = tri == g["tri"]
    assert same.mean() >= 0.999
    hit = same & (tri >= 0)
    assert (np.abs(t[hit] - g["t"][hit]) / g["t"][hit]).max() <= 1e-4
    ref.close()


@pytest.mark.parametrize("name", CASES)
def test_host_builder_matches_reference_digests(sb, name):
    c, obj, mtl, texdir = case_paths(name)
    hs = sb.HostScene(obj, mtl, texdir)
    nodes, tris, shtris, matids = hs.bvh()
    assert (hs.n_nodes, hs.n_tris, hs.depth) == (c["n_nodes"], c["n_tris"], c["depth"])
    assert sha(nodes) == c["sha256_nodes"]
    assert sha(tris) == c["sha256_tris"]
    assert sha(shtris) == c["sha256_shtris"]
    assert [int(x) for x in matids] == c["matids"]
    hs.close()


@pytest.mark.parametrize("flags", [4, 4 | 1, 2, 0])
def test_host_builder_build_flags_vs_oracle(sb, flags):
    """BVH::Construct flags (src/bvh/tree.h:21-25): useSah sweep, noShadingData, fastBuild (16-bin), median."""
    R = need_oracle("scalar")
    obj = scene_path("lancia")
    ref = R.RefScene(obj, build_flags=flags, variant="scalar")
    hs = sb.HostScene(obj, build_flags=flags)
    nodes, tris, shtris, matids = hs.bvh()
    assert nodes.tobytes() == ref.nodes().tobytes()
    assert tris.tobytes() == ref.tris().tobytes()
    if not flags & 1:
        assert shtris.tobytes() == ref.shtris().tobytes()
    ref.close()
    hs.close()


def test_host_materials_and_textures_vs_oracle(sb):
    """LoadMaterialDescs / LoadTextures / MakeMaterials (src/shading/material.cpp:59-192) on the authored
    data/barracks.mtl: same flags, same RGB8 mip chain bytes as the reference's MipmapTexture."""
    R = need_oracle("scalar")
    c, obj, mtl, texdir = case_paths("barracks_tex")
    ref = R.RefScene(obj, mtl, texdir, variant="scalar")
    hs = sb.HostScene(obj, mtl, texdir)
    rm, hm, ht = ref.materials(), hs.materials(), hs.textures()
    assert len(rm) == len(hm) and len(hm) >= 1
    for a, b in zip(rm, hm):
        assert a["flags"] == b["flags"]
        if a["tex"] is not None:
            t = ht[b["texture"]]
            assert (t["width"], t["height"], t["mips"]) == (a["tex_w"], a["tex_h"], a["tex_mips"])
            assert t["data"].tobytes() == a["tex"].tobytes()
    ref.close()
    hs.close()


def test_host_loader_errors(sb):
    with pytest.raises(sb.SnlError):
        sb.HostScene("/nonexistent/file.obj")
