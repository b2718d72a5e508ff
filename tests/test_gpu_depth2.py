"""Two reflection bounces (BASELINE config 5: "reflection depth 2"), snl_set_reflection_depth(2). The reference itself
stops after one (`cache.reflections < 1`, src/scene_trace.cpp:454-460), so the comparison is the reference with exactly
that condition relaxed to `< 2`: oracle/_ref/libsnailref_depth2.so (oracle/build_ref.sh compiles a generated copy of
scene_trace.cpp). Depth 1 stays the default and is what every other test pins."""
import os

import numpy as np
import pytest

from conftest import DATA, TEXDIR, compare_img, need_oracle, scene_path

pytestmark = pytest.mark.gpu

THREADS = min(32, os.cpu_count() or 1)


@pytest.mark.parametrize("name,mtl,inside,W,H", [("box", None, True, 640, 400), ("lancia", None, False, 1920, 1080),
                                                 ("feline", None, False, 960, 540), ("barracks", "barracks.mtl", False, 960, 540)])
def test_two_bounces_vs_patched_reference(sb, tmp_path, name, mtl, inside, W, H):
    R = need_oracle("depth2")
    path = scene_path(name)
    mtl_path = os.path.join(DATA, mtl) if mtl else None
    ref = R.RefScene(path, mtl_path, TEXDIR + "/" if mtl else "", variant="depth2")
    sc = sb.Scene(0)
    sc.load_obj(path, mtl_path, TEXDIR + "/" if mtl else None)
    cam, light = sb.bench_camera(*ref.bbox(), inside=inside)
    light2 = sb.Light((light.pos[0] + 0.1 * light.radius, light.pos[1], light.pos[2]), (0.3, 0.6, 1.0), 0.5 * light.radius)
    gv = {6: 1, 7: 1}
    ref.set_gvals(gv)
    ref.set_lights([[*l.pos, *l.color, l.radius] for l in (light, light2)])
    img_r, st_r, _ = ref.render(cam.as_array(), W, H, threads=THREADS)
    sc.set_options(gv)
    sc.set_lights([light, light2])
    img_1, st_1 = sc.render(cam, W, H)
    sc.set_reflection_depth(2)
    img_g, st_g = sc.render(cam, W, H)
    ok, line = compare_img(img_g, img_r, tmp_path)
    print(f"{name} {W}x{H} two bounces: {line}; rays gpu {st_g['rays']} patched reference {st_r['rays']} (one bounce: {st_1['rays']})")
    assert ok, line
    assert abs(st_g["rays"] - st_r["rays"]) <= 0.002 * st_r["rays"]
    assert st_g["reflection"] > st_1["reflection"] and st_g["primary"] == st_1["primary"] and st_g["shadow"] == st_1["shadow"]
    # the tile path and the multi-GPU entry take the same kernel
    tiles = [(x, y, 16, min(64, H - y)) for y in range(0, H, 64) for x in range(0, W - 15, 16)]
    part = np.zeros_like(img_g)
    sc.render(cam, W, H, tiles=tiles, out=part)
    assert (part[:, : W // 16 * 16] == img_g[:, : W // 16 * 16]).all()
    comm = sb.Comm.local([sc], W, H)
    img_m, st_m = comm.render(cam)
    assert (img_m == img_g).all() and st_m["rays"] == st_g["rays"]
    comm.close()
    sc.set_reflection_depth(1)
    img_b, _ = sc.render(cam, W, H)
    assert (img_b == img_1).all()
    sc.close()
    ref.close()


def test_depth2_limits(sb):
    sc = sb.Scene(0)
    with pytest.raises(sb.SnlError):
        sc.set_reflection_depth(3)
    sc.load_obj(os.path.join(DATA, "panes.obj"), os.path.join(DATA, "panes.mtl"), TEXDIR + "/")
    sc.set_options({6: 1, 7: 1})
    sc.set_reflection_depth(2)
    with pytest.raises(sb.SnlError):  # transparent materials: the unrolled transparency loop has no second bounce
        sc.render(sb.Camera((0.0137, 0.0211, -2.0)), 64, 48)
    sc.close()
