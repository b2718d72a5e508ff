"""The reference-side adapter include/scene_b200.h (`Scene<B200Acc>`: the reference's Scene template with RayTrace
answered by libsnail_b200.so), compiled INSIDE the reference by oracle/build_ref.sh (oracle/ref_adapter.cpp) into
oracle/_ref/libsnailref_b200adapter.so. The frame under test is produced by the reference's OWN Render() / RenderTask::Work
packet loop, ray generator, thread pool and ConvColor (src/render.cpp:47-280) with `scene->RayTrace(...)`
(src/render.cpp:81,115; src/scene.h:44-46) swapped for SceneB200::RayTrace; it is compared with the pure-reference frame
of the same library, through tools/compare_img."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import DATA, TEXDIR, compare_img, need_oracle, scene_path

pytestmark = pytest.mark.gpu

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def adapter_lib():
    R = need_oracle("b200adapter")
    L = R.lib("b200adapter")
    L.ref_adapter_create.restype = C.c_void_p
    L.ref_adapter_create.argtypes = [C.c_void_p, C.c_int]
    L.ref_adapter_free.argtypes = [C.c_void_p]
    L.ref_adapter_set_lights.argtypes = [C.c_void_p, _f32p, C.c_int]
    L.ref_adapter_render.restype = C.c_double
    L.ref_adapter_render.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, _u8p, C.POINTER(C.c_uint)]
    return R, L


CASES = [  # name, obj, mtl, camera / light override, W, H, gVals, threads
    ("box", None, None, True, 320, 200, {6: 1, 7: 1}, 4),
    ("barracks", None, "barracks.mtl", False, 256, 160, {6: 1, 7: 1}, 4),  # Tex + Uber + default materials
    ("feline", None, None, False, 320, 192, {7: 1}, 2),                     # simple shading + bounce
    ("panes", os.path.join(DATA, "panes.obj"), "panes.mtl", None, 128, 80, {6: 1, 7: 1}, 1),  # transparency chains
    ("box", None, None, True, 96, 64, {6: 1, 9: 1}, 1),                      # 2x2 supersampling in RenderTask::Work
]


@pytest.mark.parametrize("name,obj,mtl,inside,W,H,gv,threads", CASES)
def test_reference_render_loop_over_scene_b200(sb, tmp_path, name, obj, mtl, inside, W, H, gv, threads):
    R, L = adapter_lib()
    path = obj or scene_path(name)
    ref = R.RefScene(path, os.path.join(DATA, mtl) if mtl else None, TEXDIR + "/" if mtl else "", variant="b200adapter")
    if inside is None:
        cam = R.camera_array((0.0137, 0.0211, -2.0))
        light = np.array([0.5, 1.0, -2.0, 1.0, 1.0, 1.0, 12.0], np.float32)
    else:
        cam, light = R.bench_camera(*ref.bbox(), inside=inside)
    ref.set_gvals(gv)
    ref.set_lights([light])
    img_r, st_r, _ = ref.render(cam, W, H, threads=threads)
    ad = L.ref_adapter_create(ref.h, 0)
    assert ad, "SceneB200 construction failed (see stderr)"
    L.ref_adapter_set_lights(ad, np.ascontiguousarray(light, np.float32), 1)
    img_a = np.zeros((H, W, 3), np.uint8)
    rays = C.c_uint()
    ms = L.ref_adapter_render(ad, cam, W, H, threads, 0, img_a.reshape(-1), C.byref(rays))
    assert ms >= 0
    ok, line = compare_img(img_a, img_r, tmp_path, frac=0.999 if name != "panes" else 0.95)  # panes 128x80: 3.5 % reference-UB pixels
    print(f"{name} {W}x{H} {gv}: reference Render() over SceneB200::RayTrace, {threads} threads, {ms:.0f} ms: {line}; "
          f"rays {rays.value} ref {st_r['rays']}")
    assert ok, line
    assert abs(int(rays.value) - st_r["rays"]) <= 0.01 * st_r["rays"]
    if not gv.get(9):
        # the batch entry point of the same adapter (one snl_render) gives the same frame as its per-packet RayTrace
        img_b = np.zeros((H, W, 3), np.uint8)
        assert L.ref_adapter_render(ad, cam, W, H, threads, 1, img_b.reshape(-1), C.byref(rays)) >= 0
        assert (img_b == img_a).all()
    L.ref_adapter_free(ad)
    ref.close()
