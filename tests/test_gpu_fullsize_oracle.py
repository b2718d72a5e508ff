"""BASELINE.json configs 3-5 at their FULL sizes, GPU frame against the LIVE reference (oracle/_ref, the reference's
own Render() compiled in place, veclib scalar back-end = the canonical oracle) on min(32, nproc) host threads, judged by
tools/compare_img (north_star: max channel error <= 1/255 on >= 99.9 % of pixels) plus TreeStats ray counts within
0.2 %. Also the two authored / instanced scenes whose committed goldens are only 128 x 80 (`panes`: transparency +
TransparentMaterial; `lancia_inst`: Scene<DBVH>) at 960 x 540.

The reference renders a 4K Whitted frame in about a second on 16-32 threads, so these run in seconds each."""
import os

import numpy as np
import pytest

from conftest import DATA, TEXDIR, compare_img, need_oracle, scene_path

pytestmark = pytest.mark.gpu

THREADS = min(32, os.cpu_count() or 1)  # src/thread_pool.cpp:6 caps the reference's pool at 32

# name, mtl, W, H, gVals  (configs[2], [3], [4]; abrams.obj is absent from the checkout, barracks is config 4's mesh)
CASES = [
    ("feline", None, 1920, 1080, {6: 1, 7: 1}),
    ("barracks", "barracks.mtl", 3840, 2160, {6: 1}),
    ("lancia", None, 3840, 2160, {6: 1, 7: 1}),
]


def _pair(sb, R, name, mtl):
    path = scene_path(name)
    mtl_path = os.path.join(DATA, mtl) if mtl else None
    texdir = TEXDIR + "/" if mtl else ""
    ref = R.RefScene(path, mtl_path, texdir, variant="scalar")
    sc = sb.Scene(0)
    sc.load_obj(path, mtl_path, texdir if mtl else None)
    return ref, sc


@pytest.mark.parametrize("name,mtl,W,H,gv", CASES)
def test_full_size_frame_vs_live_reference(sb, tmp_path, name, mtl, W, H, gv):
    R = need_oracle("scalar")
    ref, sc = _pair(sb, R, name, mtl)
    cam, light = sb.bench_camera(*ref.bbox())
    ref.set_gvals(gv)
    ref.set_lights([[*light.pos, *light.color, light.radius]])
    img_r, st_r, ms = ref.render(cam.as_array(), W, H, threads=THREADS)
    sc.set_options(gv)
    sc.set_lights([light])
    img_g, st_g = sc.render(cam, W, H)
    ok, line = compare_img(img_g, img_r, tmp_path)
    print(f"{name} {W}x{H} {gv}: {line}; rays gpu {st_g['rays']} ref {st_r['rays']}; reference {ms:.0f} ms on {THREADS} threads, "
          f"gpu {st_g['kernel_ms']:.3f} ms")
    assert ok, line
    assert abs(st_g["rays"] - st_r["rays"]) <= 0.002 * st_r["rays"]
    # nearest-hit records of the same frame: ids exact where unambiguous, t within 1e-5 (bit-identical in practice)
    tri_r, t_r, _ = ref.trace_primary(cam.as_array(), W, H)
    tri_g, t_g, _ = sc.trace_primary(cam, W, H)
    same = tri_g == tri_r
    hit = same & (tri_r >= 0)
    assert (~same).sum() <= 1e-4 * W * H
    assert (np.abs(t_g[hit] - t_r[hit]) <= 1e-5 * np.abs(t_r[hit])).all()
    assert float((t_g[hit] == t_r[hit]).mean()) >= 0.9999
    ref.close()
    sc.close()


def test_panes_transparency_960x540_vs_live_reference(sb, tmp_path):
    """a18 + TransparentMaterial on the authored panes scene (tools/make_panes.py) at 960 x 540. Pixels the reference
    shades from an uninitialised Sample (src/scene_trace.cpp:240-290: a masked lane 0 in the transparency / reflection
    packets) are worked out by the plain-C restatement and excluded; their share is printed."""
    import snail_oracle as O
    R = need_oracle("scalar")
    obj, mtl = os.path.join(DATA, "panes.obj"), os.path.join(DATA, "panes.mtl")
    ref = R.RefScene(obj, mtl, TEXDIR + "/", variant="scalar")
    sc = sb.Scene(0)
    sc.load_obj(obj, mtl, TEXDIR + "/")
    cam = sb.Camera((0.0137, 0.0211, -2.0))
    light = sb.Light((0.5, 1.0, -2.0), (1.0, 1.0, 1.0), 12.0)
    W, H = 960, 540
    hs = sb.HostScene(obj, mtl, TEXDIR + "/")
    hn, ht, hsh, hm = hs.bvh()
    orc = O.OracleScene(hn, ht, hsh, hm, hs.materials(), hs.textures())
    for gv in ({6: 1}, {6: 1, 7: 1}):
        ref.set_gvals(gv)
        ref.set_lights([[*light.pos, *light.color, light.radius]])
        img_r, st_r, _ = ref.render(cam.as_array(), W, H, threads=THREADS)
        sc.set_options(gv)
        sc.set_lights([light])
        img_g, st_g = sc.render(cam, W, H)
        orc.set_gvals(gv)
        orc.set_lights([[*light.pos, *light.color, light.radius]])
        undef = np.zeros((H, W), np.uint8)
        orc.render(cam.as_array(), W, H, undef=undef)
        u = undef != 0
        a, b = img_g.copy(), img_r.copy()
        a[u] = 0
        b[u] = 0
        ok, line = compare_img(a, b, tmp_path)
        print(f"panes {W}x{H} {gv}: {line}; reference-undefined pixels {u.mean()*100:.2f} %; transparency rays gpu "
              f"{st_g['transparency']}; rays gpu {st_g['rays']} ref {st_r['rays']}")
        assert st_g["transparency"] > 0
        assert ok, line
        assert u.mean() < 0.08
        assert abs(st_g["rays"] - st_r["rays"]) <= max(4 * int(u.sum()), 0.002 * st_r["rays"])
    orc.close()
    ref.close()
    sc.close()


def test_lancia_instanced_960x540_vs_live_reference(sb, tmp_path):
    """Scene<DBVH> (61 instances of lancia, the goldens' set) at 960 x 540 through the live reference: hits and every
    well-defined stage; the lit bounce keeps the documented reference defect (tests/test_gpu_dbvh.py)."""
    from conftest import GOLDEN
    R = need_oracle("scalar")
    g = np.load(os.path.join(GOLDEN, "lancia_inst.npz"))
    path = scene_path("lancia")
    ref = R.RefScene(path, variant="scalar")
    ref.set_instances(g["instances"])
    sc = sb.Scene(0)
    sc.load_obj(path)
    sc.set_instances(g["instances"])
    cam = sb.Camera.from_array(g["camera"])
    lg = g["light"]
    light = sb.Light(lg[0:3], lg[3:6], lg[6])
    W, H = 960, 540
    inst_r, tri_r, t_r, _ = ref.trace_primary_inst(cam.as_array(), W, H)
    inst_g, tri_g, t_g, _ = sc.trace_primary_inst(cam, W, H)
    same = (inst_g == inst_r) & (tri_g == tri_r)
    hit = same & (inst_r >= 0)
    print(f"lancia_inst {W}x{H}: (instance, triangle) identical on {same.mean()*100:.4f} %, hits {int(hit.sum())}")
    assert (~same).sum() <= 1e-3 * W * H
    assert float((t_g[hit] == t_r[hit]).mean()) >= 0.9999
    lane0 = np.repeat(inst_r[:, 0::4], 4, axis=1)[:, :W]
    undef = (inst_r == 0) & (tri_r == 0) & (lane0 < 0)  # src/scene_trace.cpp:410-423, see tools/make_golden.py
    for stage, gv, lit, need in (("simple_nolight", {}, False, 0.999), ("simple_light", {}, True, 0.999),
                                 ("refl_nolight", {7: 1}, False, 0.999), ("whitted", {7: 1}, True, 0.98)):
        ref.set_gvals(gv)
        ref.set_lights([[*light.pos, *light.color, light.radius]] if lit else [])
        img_r, st_r, _ = ref.render(cam.as_array(), W, H, threads=THREADS)
        sc.set_options(gv)
        sc.set_lights([light] if lit else [])
        img_g, st_g = sc.render(cam, W, H)
        a, b = img_g.copy(), img_r.copy()
        a[undef] = 0
        b[undef] = 0
        ok, line = compare_img(a, b, tmp_path, frac=need)
        print(f"lancia_inst/{stage}: {line}; rays gpu {st_g['rays']} ref {st_r['rays']}")
        assert ok, (stage, line)
        assert abs(st_g["rays"] - st_r["rays"]) <= max(8, 0.003 * st_r["rays"])
    ref.close()
    sc.close()
