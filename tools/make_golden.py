#!/usr/bin/env python
"""TEST INFRASTRUCTURE — generates tests/golden/*.npz + tests/golden/manifest.json from the reference itself.

The reference (nadult/Snail) ships no golden vectors or known-answer tests for its tracing path, so
the goldens are OUTPUTS OF THE REFERENCE RUN HERE: oracle/_ref/libsnailref_scalar.so is the unmodified reference
hot path compiled in place from /root/reference by oracle/build_ref.sh on veclib's scalar back-end (exact 1/x and
1/sqrt, so the numbers do not depend on the host's rcpps/rsqrtps tables). Re-run with

    bash oracle/build_ref.sh && python tools/make_golden.py

Each case stores: the camera, the light, W, H, the per-pixel nearest hit {triId, t, bary} of
BVH::TraversePrimary, the BGR8 frame of Render() for every gVals stage, TreeStats ray counts, a set of incoherent
rays with their hits, a set of shadow rays with their verdicts, and sha256 digests of the builder's output
(BVH::nodes / tris / shTris) so that the host ingest + SAH builder can be pinned without the oracle.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import snailref as R  # noqa: E402
import snail_oracle as O  # noqa: E402
sys.path.insert(0, ROOT)
import snail_b200 as sb  # noqa: E402  (host ingest + builder only: no device needed)

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(ROOT, "oracle", "_ref", "scenes")
TEXDIR = os.path.join(ROOT, "oracle", "_ref", "data") + "/"
DATA = os.path.join(ROOT, "data")

STAGES = {
    "simple_nolight": ({}, False),
    "depth": ({1: 1}, False),
    "simple_light": ({}, True),
    "full_light": ({6: 1}, True),
    "whitted": ({6: 1, 7: 1}, True),
    "whitted_ss": ({6: 1, 7: 1, 9: 1}, True),  # + 2x2 supersampling (src/render.cpp:71-110)
    "whitted_3l": ({6: 1, 7: 1}, 3),  # three lights (out["lights3"]): colours, a radius that leaves part of the scene dark
}


def three_lights(light, bmin, bmax):
    """The case's light plus two coloured ones offset from it (they stay inside the closed test rooms); the second has a
    radius shorter than the scene, so TraceLight's attenuation reaches zero inside the frame (src/scene_trace.cpp:575-585)."""
    S = np.float32(np.linalg.norm(np.asarray(bmax, np.float32) - np.asarray(bmin, np.float32)))
    p = np.asarray(light[:3], np.float32)
    l1 = np.array([*(p + np.float32([0.15, 0.10, 0.05]) * S), 1.0, 0.25, 0.1, 0.45 * S], np.float32)
    l2 = np.array([*(p + np.float32([-0.20, -0.05, 0.10]) * S), 0.1, 0.3, 1.0, float(light[6])], np.float32)
    return np.stack([np.asarray(light, np.float32), l1, l2])

# name -> (obj, mtl, inside camera, W, H)
CASES = {
    "box": ("box.obj", None, True, 96, 80),
    "lancia": ("lancia.obj", None, False, 128, 80),
    "feline": ("feline.obj", None, False, 128, 80),
    "barracks_tex": ("barracks.obj", "barracks.mtl", False, 128, 80),
    # authored multi-material scene (tools/make_panes.py): Uber d < 1, TransparentMaterial, transparency chains
    "panes": ("@panes.obj", "panes.mtl", False, 128, 80),
}
# cases with a fixed camera / light instead of bench_camera: name -> (camera pos, light [pos, color, radius])
# (camera slightly off the axes of symmetry: with it at x = y = 0 reflected rays run exactly into the room's corner edges)
VIEWS = {"panes": ((0.0137, 0.0211, -2.0), [0.5, 1.0, -2.0, 1.0, 1.0, 1.0, 12.0])}


def random_rays(n, bmin, bmax, seed):
    rng = np.random.default_rng(seed)
    c = (bmin + bmax) * 0.5
    ext = bmax - bmin
    o = (c + (rng.random((n, 3)) - 0.5) * ext * 2.5).astype(np.float32)
    target = (c + (rng.random((n, 3)) - 0.5) * ext).astype(np.float32)
    d = target - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return o, d.astype(np.float32)


def make_instances(n, scale, seed=1):
    """Instances in the spirit of MakeDBVH (src/rtracer.cpp:146-178): n random rotations + translations on a grid of
    `scale`, plus the identity instance. Returns (n + 1, 12) float32 {rotation rows, translation}."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        a = rng.random(3) * 2 * np.pi
        (cx, cy, cz), (sx, sy, sz) = np.cos(a), np.sin(a)
        rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
        ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
        rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
        rot = (rz @ ry @ rx).astype(np.float32)
        tr = ((rng.integers(0, 1000, 3) - 500) * scale).astype(np.float32)
        out.append(np.concatenate([rot.reshape(-1), tr]))
    out.append(np.float32([1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0, 0]))
    return np.array(out, np.float32)


INST_STAGES = {"simple_nolight": ({}, False), "depth": ({1: 1}, False), "simple_light": ({}, True),
               "refl_nolight": ({7: 1}, False),
               "whitted": ({7: 1}, True)}  # Scene<DBVH> has no shading data: gVals[6] changes nothing
# name -> (obj, instances besides the identity, W, H)
INST_CASES = {"lancia_inst": ("lancia.obj", 60, 160, 96), "box_inst": ("box.obj", 24, 128, 80)}


def instanced_cases(manifest):
    """Scene<DBVH> (src/dbvh/tree.h): top-level tree + instance order, per-pixel {instance, triangle, t, bary} of
    DBVH::TraversePrimary, frames per stage."""
    for name, (obj, n, W, H) in INST_CASES.items():
        ref = R.RefScene(os.path.join(SCENES, obj), variant="scalar")
        bmin, bmax = ref.bbox()
        S = float(np.linalg.norm(bmax - bmin))
        inst = make_instances(n, S * 0.004, seed=len(name))
        ref.set_instances(inst)
        nodes, elems, boxes = ref.dbvh()
        c = (bmin + bmax) * 0.5
        cam = R.camera_array(c - np.float32([0.07, 0.11, 1]) * np.float32(3.5 * S))
        light = np.array([*cam[:3], 1, 1, 1, 12 * S], np.float32)
        out = {"instances": inst, "camera": cam, "light": light, "size": np.array([W, H], np.int32),
               "dbvh_elements": elems, "dbvh_boxes": boxes}
        ref.set_gvals({})
        ref.set_lights([])
        out["inst"], out["tri"], out["t"], out["bary"] = ref.trace_primary_inst(cam, W, H)
        # same uninitialised-normal quirk as in BVH scenes, here for (instance 0, triangle 0) (src/scene_trace.cpp:410-423)
        lane0 = np.repeat(out["inst"][:, 0::4], 4, axis=1)[:, :W]
        out["undef_simple"] = (out["inst"] == 0) & (out["tri"] == 0) & (lane0 < 0)
        rays = {}
        for stage, (gv, lit) in INST_STAGES.items():
            ref.set_gvals(gv)
            ref.set_lights([light] if lit else [])
            img, st, _ = ref.render(cam, W, H, threads=1)
            out["img_" + stage] = img
            rays[stage] = st["rays"]
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **out)
        manifest["instanced"][name] = {"obj": obj, "W": W, "H": H, "n_instances": int(len(inst)), "n_nodes": int(len(nodes)),
                                       "sha256_dbvh_nodes": sha(nodes), "rays": rays,
                                       "hits": int((out["inst"] >= 0).sum()),
                                       "instances_hit": int(len(np.unique(out["inst"][out["inst"] >= 0])))}
        print(name, manifest["instanced"][name])
        ref.close()


FARM_CASES = {"box": 1, "barracks_tex": 1, "panes": 1}  # CASES entries also rendered through the farm overload
FARM_W, FARM_H = 128, 128  # 8 x 2 tiles of blockWidth x blockHeight = 16 x 64 (src/rtbase_math.h:31-34)


def farm_cases(manifest):
    """Render(scene, camera, resx, resy, data, coords, offsets, options, rank, threads) (src/render.cpp:243-265): the
    planar delta-coded 16x64 tiles the farm nodes produce. Two tile sets per scene: (A) every tile, back to back in a
    shuffled order, gVals[8] rank tint with rank 5; (B) every other tile with gaps between them, no tint."""
    from snail_b200 import farm
    W, H = FARM_W, FARM_H
    for name in FARM_CASES:
        obj, mtl, inside, _, _ = CASES[name]
        mtl_path = os.path.join(DATA, mtl) if mtl else None
        obj_path = os.path.join(DATA, obj[1:]) if obj.startswith("@") else os.path.join(SCENES, obj)
        ref = R.RefScene(obj_path, mtl_path, TEXDIR if mtl else "", variant="scalar")
        bmin, bmax = ref.bbox()
        cam, light = R.bench_camera(bmin, bmax, inside=inside)
        if name in VIEWS:
            cam = R.camera_array(VIEWS[name][0])
            light = np.array(VIEWS[name][1], np.float32)
        hs = sb.HostScene(obj_path, mtl_path, TEXDIR if mtl else "")
        hn, ht, hsh, hm = hs.bvh()
        orc = O.OracleScene(hn, ht, hsh, hm, hs.materials(), hs.textures())
        tiles = farm.divide_image(W, H)
        rng = np.random.default_rng(11)
        out = {"camera": cam, "light": light, "size": np.array([W, H], np.int32)}
        ref.set_lights([light])
        orc.set_lights([light])
        tri, _, _ = ref.trace_primary(cam, W, H)
        lane0 = np.repeat(tri[:, 0::4], 4, axis=1)[:, :W]
        for tag, gv, rank, sel, stride in (("A", {6: 1, 7: 1, 8: 1}, 5, rng.permutation(len(tiles)), 3072),
                                           ("B", {6: 1, 7: 1}, 0, np.arange(0, len(tiles), 2), 4096)):
            coords = tiles[sel]
            offsets = (np.arange(len(coords)) * stride).astype(np.int32)
            ref.set_gvals(gv)
            data, st = ref.render_tiles(cam, W, H, coords, offsets, int(offsets.max()) + 3072, rank=rank)
            # the same tiles through the image overload must hold the same bytes (sanity of the generator itself)
            img, _, _ = ref.render(cam, W, H, threads=1)
            dec = farm.decode_farm_tiles(data, coords, offsets, W, H)
            orc.set_gvals({k: v for k, v in gv.items() if k != 8})
            u = np.zeros((H, W), np.uint8)
            orc.render(cam, W, H, undef=u)
            undef = (u != 0) | ((tri == 0) & (lane0 < 0))
            for x, y, w, h in (coords if not gv.get(8) else []):  # the image overload tints with rank 0
                ok = (dec[y:y + h, x:x + w] == img[y:y + h, x:x + w]).all(axis=2) | undef[y:y + h, x:x + w]
                assert ok.all(), (name, tag, x, y)
            out["coords_" + tag], out["offsets_" + tag], out["data_" + tag] = coords, offsets, data
            out["rank_" + tag] = np.int32(rank)
            out["gvals_" + tag] = np.array([[k, v] for k, v in gv.items()], np.int32)
            out["undef_" + tag] = undef
            out["rays_" + tag] = np.int64(st["rays"])
        np.savez_compressed(os.path.join(GOLDEN, "farm_" + name + ".npz"), **out)
        manifest["farm"][name] = {"W": W, "H": H, "tiles_A": int(len(out["coords_A"])), "tiles_B": int(len(out["coords_B"])),
                                  "rays_A": int(out["rays_A"]), "rays_B": int(out["rays_B"]),
                                  "undef_A": int(out["undef_A"].sum()), "sha256_data_A": sha(out["data_A"])}
        print("farm", name, manifest["farm"][name])
        orc.close()
        ref.close()


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    os.makedirs(GOLDEN, exist_ok=True)
    manifest = {"generator": "tools/make_golden.py", "oracle": "oracle/_ref/libsnailref_scalar.so (reference compiled in place)",
                "backend": R.lib("scalar").ref_backend().decode(), "cases": {}}
    for name, (obj, mtl, inside, W, H) in CASES.items():
        mtl_path = os.path.join(DATA, mtl) if mtl else None
        obj_path = os.path.join(DATA, obj[1:]) if obj.startswith("@") else os.path.join(SCENES, obj)
        ref = R.RefScene(obj_path, mtl_path, TEXDIR if mtl else "", variant="scalar")
        bmin, bmax = ref.bbox()
        cam, light = R.bench_camera(bmin, bmax, inside=inside)
        if name in VIEWS:
            cam = R.camera_array(VIEWS[name][0])
            light = np.array(VIEWS[name][1], np.float32)
        out = {"camera": cam, "light": light, "size": np.array([W, H], np.int32)}
        ref.set_gvals({})
        ref.set_lights([])
        tri, t, bary = ref.trace_primary(cam, W, H)
        out["tri"], out["t"], out["bary"] = tri, t, bary
        out["dirs"] = ref.primary_dirs(cam, W, H)
        # Reference quirk (src/scene_trace.cpp:410-411 simple shading, :316-330 full shading): a quad whose lane 0 missed never gets
        # its Sample::normal initialised, and lanes of that quad that hit triangle 0 (== the "no object" value the
        # missed lane was given, :374) are skipped by the per-lane fix-up (:423) -> they are shaded with whatever the
        # stack held. Those pixels are undefined behaviour in the reference and are excluded from image goldens.
        lane0 = np.repeat(tri[:, 0::4], 4, axis=1)[:, :W]
        out["undef_simple"] = (tri == 0) & (lane0 < 0)
        # the same on the 2W x 2H grid the supersampled frame is traced on, reduced to output pixels
        tri2, _, _ = ref.trace_primary(cam, 2 * W, 2 * H)
        lane02 = np.repeat(tri2[:, 0::4], 4, axis=1)[:, :2 * W]
        u2 = (tri2 == 0) & (lane02 < 0)
        out["undef_ss"] = u2.reshape(H, 2, W, 2).any(axis=(1, 3))
        # The same quirk at the deeper levels (reflection / transparency packets are masked, and a masked lane 0 reads
        # object 0 as well): which pixels it reaches is worked out by the plain-C restatement, which follows the
        # reference's packet / quad structure and flags those lanes (oracle/snail_oracle.c shade_block).
        hs = sb.HostScene(obj_path, mtl_path, TEXDIR if mtl else "")
        hn, ht, hsh, hm = hs.bvh()
        orc = O.OracleScene(hn, ht, hsh, hm, hs.materials(), hs.textures())
        rays = {}
        out["lights3"] = three_lights(light, bmin, bmax)
        for stage, (gv, lit) in STAGES.items():
            lights = list(out["lights3"]) if lit == 3 else [light] if lit else []
            ref.set_gvals(gv)
            ref.set_lights(lights)
            img, st, _ = ref.render(cam, W, H, threads=1)
            out["img_" + stage] = img
            rays[stage] = st["rays"]
            if stage != "depth":
                orc.set_gvals(gv)
                orc.set_lights(lights)
                u = np.zeros((H, W), np.uint8)
                orc.render(cam, W, H, undef=u)
                out["undef_" + stage] = (u != 0) | (out["undef_ss"] if stage.endswith("_ss") else out["undef_simple"])
        orc.close()
        # incoherent rays with their own origins (Context<0,0> / <0,1>)
        n = 256 * 3 + 77
        o, d = random_rays(n, bmin, bmax, 7)
        mask = (np.arange(n) % 5 != 0).astype(np.uint8)
        out["ray_o"], out["ray_d"], out["ray_mask"] = o, d, mask
        out["ray_tri"], out["ray_t"], out["ray_bary"] = ref.trace_rays(o, d, None)
        out["rayM_tri"], out["rayM_t"], out["rayM_bary"] = ref.trace_rays(o, d, mask)
        # shadow rays from one origin (ShadowContext)
        origin = ((bmin + bmax) * 0.5).astype(np.float32) if inside else \
            ((bmin + bmax) * 0.5 + np.float32([0.3, 0.5, -1]) * np.linalg.norm(bmax - bmin)).astype(np.float32)
        rng = np.random.default_rng(5)
        tmax = (rng.random(n) * 3 * np.linalg.norm(bmax - bmin)).astype(np.float32)
        tmax[::7] = -1.0
        out["sh_origin"], out["sh_tmax"] = origin, tmax
        out["sh_out"] = ref.trace_shadow(origin, d, tmax)
        # Scene::RayTrace on caller packets (RayGroup<0,0>), Whitted
        ref.set_gvals({6: 1, 7: 1})
        ref.set_lights([light])
        out["rt_rgb"], rt_rays = ref.raytrace(o, d, None)
        nodes, tris, shtris = ref.nodes(), ref.tris(), ref.shtris()
        np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **out)
        manifest["cases"][name] = {
            "obj": obj, "mtl": mtl, "inside": inside, "W": W, "H": H, "rays": rays, "raytrace_rays": rt_rays,
            "n_nodes": int(ref.n_nodes), "n_tris": int(ref.n_tris), "depth": int(ref.depth),
            "sha256_nodes": sha(nodes), "sha256_tris": sha(tris), "sha256_shtris": sha(shtris),
            "matids": [int(x) for x in ref.matids()],
            "hits": int((tri >= 0).sum()),
        }
        print(name, manifest["cases"][name]["hits"], "hits of", W * H, "rays:", rays)
        ref.close()
    manifest["instanced"] = {}
    instanced_cases(manifest)
    manifest["farm"] = {}
    farm_cases(manifest)
    with open(os.path.join(GOLDEN, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
