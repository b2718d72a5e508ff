"""CPU checks of the reference-side adapter include/scene_b200.h (`template <> class Scene<B200Acc>`): the library in
which oracle/build_ref.sh compiles it INSIDE the reference (with the reference's own render.cpp instantiated over it) is
there, loads, exports its entry points, and - like every product path - refuses to work without a CUDA device instead of
falling back. The frames it produces are compared on the GPU box (tests/test_gpu_adapter.py)."""
import ctypes as C
import os
import re

from conftest import ROOT, need_oracle, scene_path


def test_adapter_header_specialises_the_reference_template():
    src = open(os.path.join(ROOT, "include", "scene_b200.h")).read()
    assert re.search(r"template\s*<>\s*class\s+Scene<B200Acc>", src)
    for entry in ("RayTrace", "RenderFrame", "snl_raytrace", "snl_render"):
        assert entry in src, entry
    # the adapter is a caller of the C ABI only: nothing of the product's internals, nothing of the oracle
    code = re.sub(r"//.*", "", src)
    assert '#include "snail_b200.h"' in code and "snl_scene.h" not in code and "oracle" not in code


def test_adapter_compiled_inside_the_reference_loads_and_needs_a_device(sb, capfd):
    R = need_oracle("b200adapter")
    L = R.lib("b200adapter")
    for sym in ("ref_adapter_create", "ref_adapter_free", "ref_adapter_set_lights", "ref_adapter_render"):
        assert hasattr(L, sym), sym
    if sb.load_library().snl_device_count() > 0:
        return  # construction and frames: tests/test_gpu_adapter.py
    L.ref_adapter_create.restype = C.c_void_p
    L.ref_adapter_create.argtypes = [C.c_void_p, C.c_int]
    ref = R.RefScene(scene_path("box"), None, "", variant="b200adapter")
    try:
        assert L.ref_adapter_create(ref.h, 0) is None
        assert "no CUDA device" in capfd.readouterr().err
    finally:
        ref.close()
