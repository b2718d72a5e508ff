"""Named bench / parity workloads: scene + resolution + gVals + light, resolved against the asset
directories that exist on this machine. Product-side: never touches oracle/."""
from __future__ import annotations

import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ASSET_DIRS = [os.path.join(ROOT, "assets"), "/root/reference"]
DATA = os.path.join(ROOT, "data")

# name -> (obj, mtl (authored, in data/) or None, inside-camera?, gVals, lit?, description)
WORKLOADS = {
    # BASELINE.json metric config ("sponza 1080p Whitted"): scenes/sponza.obj and its textures are absent from the
    # reference checkout, so the default workload is a procedurally generated atrium of the same shape (268 K
    # triangles, 12 textured + 1 Uber + 1 default material, interior camera) - tools/make_sponza_like.py.
    "sponza_like_whitted": dict(obj="sponza_like.obj", mtl="@sponza_like.mtl", inside=True, gvals={6: 1, 7: 1}, lit=True,
                                desc="sponza_like.obj (procedural stand-in for the absent sponza.obj: 267 936 tris, textured, "
                                     "camera inside), Whitted: primary + shadow + 1 reflection, 1 point light"),
    "sponza_like_shadow": dict(obj="sponza_like.obj", mtl="@sponza_like.mtl", inside=True, gvals={6: 1}, lit=True,
                               desc="sponza_like.obj textured, primary + shadow"),
    # BASELINE.json configs[1] / metric: sponza.obj is absent from the reference checkout (.MISSING_LARGE_BLOBS);
    # SURVEY.md §8(d) names barracks.obj + data/347.png as the textured substitute.
    "barracks_whitted": dict(obj="barracks.obj", mtl="barracks.mtl", inside=False, gvals={6: 1, 7: 1}, lit=True,
                             desc="barracks.obj (sponza.obj absent from the checkout; SURVEY 8d substitute), textured "
                                  "(data/barracks.mtl + 347.png), Whitted: primary + shadow + 1 reflection, 1 point light"),
    "barracks_shadow": dict(obj="barracks.obj", mtl="barracks.mtl", inside=False, gvals={6: 1}, lit=True,
                            desc="barracks.obj textured, primary + shadow"),
    "feline_whitted": dict(obj="feline.obj", mtl=None, inside=False, gvals={6: 1, 7: 1}, lit=True,
                           desc="feline.obj (99 732 tris) Whitted: primary + shadow + 1 reflection, 1 point light"),
    "feline_primary": dict(obj="feline.obj", mtl=None, inside=False, gvals={}, lit=False,
                           desc="feline.obj primary rays only"),
    "lancia_whitted": dict(obj="lancia.obj", mtl=None, inside=False, gvals={6: 1, 7: 1}, lit=True,
                           desc="lancia.obj (30 327 tris) Whitted, 1 point light"),
    "box_primary": dict(obj="box.obj", mtl=None, inside=True, gvals={}, lit=False, desc="box.obj primary rays only"),
    "box_light": dict(obj="box.obj", mtl=None, inside=True, gvals={}, lit=True, desc="box.obj primary + shadow"),
}


def find_asset(sub: str, name: str) -> str:
    for d in ASSET_DIRS:
        p = os.path.join(d, sub, name)
        if os.path.exists(p):
            return p
    raise FileNotFoundError(f"{sub}/{name} not found under {ASSET_DIRS}: run __graft_entry__.build() where "
                            f"/root/reference is present to copy the reference's scene assets into assets/")


def resolve(name: str) -> dict:
    w = dict(WORKLOADS[name])
    w["name"] = name
    w["obj_path"] = find_asset("scenes", w["obj"])
    if w["mtl"] and w["mtl"].startswith("@"):  # generated next to its textures, under assets/data
        w["mtl_path"] = find_asset("data", w["mtl"][1:])
        w["tex_dir"] = os.path.dirname(w["mtl_path"]) + "/"
    else:
        w["mtl_path"] = os.path.join(DATA, w["mtl"]) if w["mtl"] else None
        w["tex_dir"] = os.path.dirname(find_asset("data", "347.png")) + "/" if w["mtl"] else ""
    return w
