"""ctypes binding of libsnail_b200.so (include/snail_b200.h) + a thin host-side mirror of the reference's
Scene / Render interface (src/scene.h:13-57, src/render.h:16-23) so tests read like the reference's own call sites.

There is no CPU path here: if the library cannot be loaded, or no CUDA device is visible, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsnail_b200.so")
LIB_COUNTERS_PATH = os.path.join(HERE, "libsnail_b200_counters.so")

# names every include/snail_b200.h declaration must resolve to (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "snl_last_error", "snl_abi_version", "snl_device_count", "snl_scene_create", "snl_scene_destroy",
    "snl_scene_set_geometry", "snl_scene_set_materials", "snl_scene_set_lights", "snl_set_options", "snl_set_reflection_depth",
    "snl_scene_load_obj", "snl_scene_counts", "snl_scene_copy_host_bvh", "snl_scene_bbox", "snl_trace_primary",
    "snl_trace_rays", "snl_trace_shadow", "snl_raytrace", "snl_render", "snl_render_dev", "snl_render_farm", "snl_scene_sync",
    "snl_scene_launch_count", "snl_scene_render_variant", "snl_scene_stream", "snl_host_scene_load_obj", "snl_host_scene_free",
    "snl_host_scene_counts", "snl_host_scene_copy", "snl_host_scene_material", "snl_host_scene_texture",
    "snl_scene_set_host_scene", "snl_scene_build_device", "snl_scene_build_ms", "snl_scene_set_device_builder", "snl_host_build_binned_sah", "snl_scene_set_instances",
    "snl_trace_primary_inst", "snl_scene_copy_dbvh", "snl_host_dbvh_build", "snl_frame_create", "snl_frame_open", "snl_frame_close", "snl_frame_read",
    "snl_measure_peaks", "snl_comm_create_root", "snl_comm_attach", "snl_comm_create_local", "snl_comm_info", "snl_render_multi",
    "snl_comm_frame", "snl_comm_destroy", "snl_comm_plan",
]

# reference record layouts (src/bvh/tree.h:60-72, src/triangle.h:133-135,181-230)
NODE_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("sub", "<i4"), ("axis", "<i2"), ("first_node", "<i2")])
TRI_DTYPE = np.dtype([("a", "<f4", 3), ("ba", "<f4", 3), ("ca", "<f4", 3), ("t0", "<f4"), ("it0", "<f4"),
                      ("pad", "<i4"), ("plane", "<f4", 4)])
SHTRI_DTYPE = np.dtype([("uv", "<f4", (3, 2)), ("nrm", "<f4", (3, 3)), ("mat", "<i4")])

NO_SHADING_DATA, FAST_BUILD, USE_SAH = 1, 2, 4
BUILD_DEVICE = 8  # ours: tree built on the GPU (snl_scene_build_device)
BUILD_SAH_BINNED = 16  # with BUILD_DEVICE: the binned-SAH device builder instead of the LBVH
BUILDER_LBVH, BUILDER_SAH = 0, 1
GPUNODE_DTYPE = np.dtype([("b0", "<f4", 4), ("b1", "<f4", 4), ("bz", "<f4", 4), ("link", "<i4", 4)])
MAT_SIMPLE, MAT_UBER, MAT_TEX, MAT_TRANSPARENT = 0, 1, 2, 3


class SnlError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"snail_b200 error {code}: {msg}")
        self.code = code


class Camera(C.Structure):
    """src/camera.h:7-14"""
    _fields_ = [("pos", C.c_float * 3), ("right", C.c_float * 3), ("up", C.c_float * 3), ("front", C.c_float * 3),
                ("plane_dist", C.c_float)]

    def __init__(self, pos, right=(1, 0, 0), up=(0, 1, 0), front=(0, 0, 1), plane_dist=1.0):
        super().__init__()
        for name, v in (("pos", pos), ("right", right), ("up", up), ("front", front)):
            setattr(self, name, (C.c_float * 3)(*[float(x) for x in v]))
        self.plane_dist = float(plane_dist)

    @classmethod
    def from_array(cls, a):
        a = np.asarray(a, np.float32)
        return cls(a[0:3], a[3:6], a[6:9], a[9:12], a[12])

    def as_array(self):
        return np.array([*self.pos, *self.right, *self.up, *self.front, self.plane_dist], np.float32)


class Light(C.Structure):
    """src/light.h:5-17"""
    _fields_ = [("pos", C.c_float * 3), ("color", C.c_float * 3), ("radius", C.c_float), ("radSq", C.c_float),
                ("iRadius", C.c_float)]

    def __init__(self, pos, color=(1, 1, 1), radius=1.0):
        super().__init__()
        self.pos = (C.c_float * 3)(*[float(x) for x in pos])
        self.color = (C.c_float * 3)(*[float(x) for x in color])
        r = np.float32(radius)
        self.radius = r
        self.radSq = np.float32(r * r)
        self.iRadius = np.float32(np.float32(1.0) / r)


class Material(C.Structure):
    _fields_ = [("kind", C.c_int), ("flags", C.c_int), ("color", C.c_float * 3), ("specular", C.c_float * 3),
                ("opacity", C.c_float), ("texture", C.c_int), ("mask_texture", C.c_int)]


class Texture(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t), ("width", C.c_int), ("height", C.c_int), ("mips", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("primary", C.c_uint64), ("shadow", C.c_uint64), ("reflection", C.c_uint64),
                ("reflection_shadow", C.c_uint64), ("transparency", C.c_uint64), ("node_visits", C.c_uint64), ("tri_tests", C.c_uint64), ("gtri_tests", C.c_uint64),
                ("transparency_truncated", C.c_uint64), ("kernel_ms", C.c_float)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Peaks(C.Structure):
    _fields_ = [("l1_gbs", C.c_float), ("l2_gbs", C.c_float), ("hbm_read_gbs", C.c_float), ("issue_gwarpinst_s", C.c_float),
                ("issue_peak_gwarpinst_s", C.c_float), ("sm_clock_mhz_max", C.c_float), ("sms", C.c_int)]


_libs: dict[str, C.CDLL] = {}


def load_library(counters: bool = False) -> C.CDLL:
    path = LIB_COUNTERS_PATH if counters else os.environ.get("SNL_LIB", LIB_PATH)  # SNL_LIB: tuning variants
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise SnlError(-5, f"{path} not built: run `python -m snail_b200.build` (there is no CPU fallback)")
    L = C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    vp, ip, fp, i = C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_int
    L.snl_last_error.restype = C.c_char_p
    L.snl_scene_create.restype = vp
    L.snl_scene_create.argtypes = [i]
    L.snl_scene_destroy.argtypes = [vp]
    L.snl_scene_destroy.restype = None
    L.snl_scene_set_geometry.argtypes = [vp, vp, i, vp, vp, i, vp, i]
    L.snl_scene_set_instances.argtypes = [vp, vp, i]
    L.snl_trace_primary_inst.argtypes = [vp, vp, i, i, vp, vp, vp, vp]
    L.snl_scene_copy_dbvh.argtypes = [vp, ip, vp, ip, vp, vp]
    L.snl_host_dbvh_build.argtypes = [vp, i, vp, vp, vp, ip, vp, vp]
    L.snl_scene_build_device.argtypes = [vp, vp, vp, i, vp, i]
    L.snl_scene_set_device_builder.argtypes = [vp, i]
    L.snl_host_build_binned_sah.argtypes = [vp, i, vp, ip, vp, ip]
    L.snl_scene_build_ms.argtypes = [vp]
    L.snl_scene_build_ms.restype = C.c_float
    L.snl_scene_set_materials.argtypes = [vp, vp, i, vp, i]
    L.snl_scene_set_lights.argtypes = [vp, vp, i, vp]
    L.snl_set_options.argtypes = [vp, vp]
    L.snl_set_reflection_depth.argtypes = [vp, i]
    L.snl_scene_load_obj.argtypes = [vp, C.c_char_p, C.c_char_p, C.c_char_p, i, i]
    L.snl_scene_counts.argtypes = [vp, ip, ip, ip, ip, ip]
    L.snl_scene_copy_host_bvh.argtypes = [vp, vp, vp, vp, vp]
    L.snl_scene_bbox.argtypes = [vp, vp, vp]
    L.snl_trace_primary.argtypes = [vp, C.POINTER(Camera), i, i, vp, fp, fp]
    L.snl_trace_rays.argtypes = [vp, fp, fp, vp, i, i, vp, fp, fp]
    L.snl_trace_shadow.argtypes = [vp, fp, fp, fp, i]
    L.snl_raytrace.argtypes = [vp, fp, fp, vp, i, i, fp, C.POINTER(Stats)]
    L.snl_render.argtypes = [vp, C.POINTER(Camera), i, i, vp, i, vp, C.c_size_t, i, C.POINTER(Stats)]
    L.snl_render_dev.argtypes = [vp, C.POINTER(Camera), i, i, vp, i, vp, C.c_size_t, i]
    L.snl_render_farm.argtypes = [vp, C.POINTER(Camera), i, i, vp, vp, vp, i, i, C.POINTER(Stats)]
    L.snl_scene_sync.argtypes = [vp, C.POINTER(Stats)]
    L.snl_scene_launch_count.restype = C.c_uint64
    L.snl_scene_launch_count.argtypes = [vp]
    L.snl_scene_render_variant.argtypes = [vp]
    L.snl_scene_stream.restype = C.c_uint64
    L.snl_scene_stream.argtypes = [vp]
    L.snl_host_scene_load_obj.restype = vp
    L.snl_host_scene_load_obj.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, i, i]
    L.snl_host_scene_free.argtypes = [vp]
    L.snl_host_scene_free.restype = None
    L.snl_host_scene_counts.argtypes = [vp, ip, ip, ip, ip, ip, ip]
    L.snl_host_scene_copy.argtypes = [vp, vp, vp, vp, vp]
    L.snl_host_scene_material.argtypes = [vp, i, C.POINTER(Material)]
    L.snl_host_scene_texture.argtypes = [vp, i, C.POINTER(Texture)]
    L.snl_scene_set_host_scene.argtypes = [vp, vp]
    L.snl_frame_create.argtypes = [vp, i, i, C.POINTER(C.c_void_p), C.c_char_p]
    L.snl_frame_open.argtypes = [vp, C.c_char_p, C.POINTER(C.c_void_p)]
    L.snl_frame_close.argtypes = [vp, vp]
    L.snl_frame_read.argtypes = [vp, vp, i, i, vp, C.c_size_t]
    L.snl_measure_peaks.argtypes = [i, C.POINTER(Peaks)]
    L.snl_comm_create_root.argtypes = [vp, i, i, i, C.c_char_p, C.POINTER(vp)]
    L.snl_comm_attach.argtypes = [vp, i, C.c_char_p, C.POINTER(vp)]
    L.snl_comm_create_local.argtypes = [C.POINTER(vp), i, i, i, C.POINTER(vp)]
    L.snl_comm_info.argtypes = [vp, ip, ip, ip, ip]
    L.snl_render_multi.argtypes = [vp, C.POINTER(Camera), vp, C.c_size_t, C.POINTER(Stats)]
    L.snl_comm_frame.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]
    L.snl_comm_plan.argtypes = [i, i, i, i, vp, ip, ip, vp]
    L.snl_comm_destroy.argtypes = [vp]
    L.snl_comm_destroy.restype = None
    _libs[path] = L
    return L


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def measure_peaks(device: int = 0) -> dict:
    """Builder-measured L1 / L2 / HBM-read bandwidth and issue rate of `device` (snl_measure_peaks)."""
    L = load_library()
    p = Peaks()
    r = L.snl_measure_peaks(device, C.byref(p))
    if r:
        raise SnlError(r, L.snl_last_error().decode())
    return {n: getattr(p, n) for n, _ in p._fields_}


def host_build_binned_sah(tris):
    """The binned-SAH builder of snl_build_sah.cu run on the host: (64-byte traversal nodes, triangle order, height)."""
    L = load_library()
    tris = np.ascontiguousarray(tris)
    assert tris.dtype.itemsize == 64
    n = len(tris)
    nodes = np.zeros(max(n - 1, 1), GPUNODE_DTYPE)
    order = np.zeros(n, np.int32)
    nn, h = C.c_int(), C.c_int()
    r = L.snl_host_build_binned_sah(_ptr(tris), n, _ptr(nodes), C.byref(nn), _ptr(order), C.byref(h))
    if r:
        raise SnlError(r, L.snl_last_error().decode())
    return nodes[: nn.value], order, h.value


def host_dbvh_build(instances, tree_min, tree_max):
    """ObjectInstance::ComputeBBox + DBVH::Construct on the host (no device): (nodes, elements, boxes)."""
    L = load_library()
    a = np.ascontiguousarray(np.asarray(instances, np.float32).reshape(-1, 12))
    lo, hi = np.ascontiguousarray(tree_min, np.float32), np.ascontiguousarray(tree_max, np.float32)
    nodes = np.zeros(2 * len(a), NODE_DTYPE)
    el, bb = np.zeros((len(a), 12), np.float32), np.zeros((len(a), 6), np.float32)
    nn = C.c_int()
    r = L.snl_host_dbvh_build(_ptr(a), len(a), _ptr(lo), _ptr(hi), _ptr(nodes), C.byref(nn), _ptr(el), _ptr(bb))
    if r:
        raise SnlError(r, L.snl_last_error().decode())
    return nodes[: nn.value], el, bb


class HostScene:
    """Host ingest + SAH builder output (no device needed): BaseScene::LoadWavefrontObj + BVH::Construct +
    MakeMaterials, src/rtracer.cpp:505-586."""

    def __init__(self, obj_path, mtl_path=None, tex_dir=None, flip_normals=0, build_flags=USE_SAH):
        self.L = load_library()
        self.h = self.L.snl_host_scene_load_obj(obj_path.encode(), (mtl_path or "").encode(), (tex_dir or "").encode(),
                                                flip_normals, build_flags)
        if not self.h:
            raise SnlError(-5, self.L.snl_last_error().decode())
        v = [C.c_int() for _ in range(6)]
        self.L.snl_host_scene_counts(self.h, *[C.byref(x) for x in v])
        self.n_nodes, self.n_tris, self.n_matids, self.depth, self.n_materials, self.n_textures = [x.value for x in v]
        self.has_shading = not (build_flags & NO_SHADING_DATA)

    def close(self):
        if getattr(self, "h", None):
            self.L.snl_host_scene_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bvh(self):
        nodes = np.zeros(self.n_nodes, NODE_DTYPE)
        tris = np.zeros(self.n_tris, TRI_DTYPE)
        shtris = np.zeros(self.n_tris if self.has_shading else 0, SHTRI_DTYPE)
        matids = np.zeros(max(self.n_matids, 1), np.int32)
        self.L.snl_host_scene_copy(self.h, _ptr(nodes), _ptr(tris), _ptr(shtris) if self.has_shading else None,
                                   _ptr(matids))
        return nodes, tris, shtris, matids[: self.n_matids]

    def bbox(self):
        nodes = self.bvh()[0]
        return nodes[0]["bmin"].copy(), nodes[0]["bmax"].copy()

    def materials(self):
        out = []
        for n in range(self.n_materials):
            m = Material()
            self.L.snl_host_scene_material(self.h, n, C.byref(m))
            out.append(dict(kind=m.kind, flags=m.flags, color=tuple(m.color), specular=tuple(m.specular),
                            opacity=m.opacity, texture=m.texture, mask_texture=m.mask_texture))
        return out

    def textures(self):
        out = []
        for n in range(self.n_textures):
            t = Texture()
            self.L.snl_host_scene_texture(self.h, n, C.byref(t))
            data = np.frombuffer((C.c_ubyte * t.size).from_address(t.data), np.uint8).copy()
            out.append(dict(data=data, width=t.width, height=t.height, mips=t.mips))
        return out


class Scene:
    """Scene<BVH> on one B200 (src/scene.h:13-57)."""

    def __init__(self, device: int = 0, counters: bool = False):
        self.L = load_library(counters)
        self.h = self.L.snl_scene_create(device)
        if not self.h:
            raise SnlError(-2, self.L.snl_last_error().decode())
        self.device = device
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.L.snl_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, code):
        if code != 0:
            raise SnlError(code, self.L.snl_last_error().decode())

    # ---- set-up ------------------------------------------------------------------------------------------------
    def set_geometry(self, nodes, tris, shtris=None, matids=None):
        nodes = np.ascontiguousarray(nodes)
        tris = np.ascontiguousarray(tris)
        assert nodes.dtype.itemsize == 32 and tris.dtype.itemsize == 64
        if shtris is not None:
            shtris = np.ascontiguousarray(shtris)
            assert shtris.dtype.itemsize == 64 and len(shtris) == len(tris)
        matids = np.ascontiguousarray(matids if matids is not None else [], np.int32)
        self._check(self.L.snl_scene_set_geometry(self.h, _ptr(nodes), len(nodes), _ptr(tris), _ptr(shtris), len(tris),
                                                  _ptr(matids) if len(matids) else None, len(matids)))

    def build_device(self, tris, shtris=None, matids=None):
        """Triangle soup (reference layouts, any order) -> BVH built on the GPU (snl_scene_build_device)."""
        tris = np.ascontiguousarray(tris)
        assert tris.dtype.itemsize == 64
        if shtris is not None:
            shtris = np.ascontiguousarray(shtris)
            assert shtris.dtype.itemsize == 64 and len(shtris) == len(tris)
        matids = np.ascontiguousarray(matids if matids is not None else [], np.int32)
        self._check(self.L.snl_scene_build_device(self.h, _ptr(tris), _ptr(shtris), len(tris),
                                                  _ptr(matids) if len(matids) else None, len(matids)))

    def set_device_builder(self, builder: int):
        """BUILDER_LBVH (fast) or BUILDER_SAH (binned SAH, the reference's tree quality) for build_device / BUILD_DEVICE."""
        self._check(self.L.snl_scene_set_device_builder(self.h, builder))

    def set_instances(self, instances):
        """instances: (n, 12) float32 = ObjectInstance {rotation rows (9), translation (3)}; n = 0: plain BVH again."""
        a = np.ascontiguousarray(np.asarray(instances, np.float32).reshape(-1, 12))
        self._check(self.L.snl_scene_set_instances(self.h, _ptr(a) if len(a) else None, len(a)))

    def dbvh(self):
        """(DBVH::nodes as NODE_DTYPE records, elements (n, 12) in the builder's order, boxes (n, 6))."""
        nn, ni = C.c_int(), C.c_int()
        self._check(self.L.snl_scene_copy_dbvh(self.h, C.byref(nn), None, C.byref(ni), None, None))
        nodes = np.zeros(nn.value, NODE_DTYPE)
        el, bb = np.zeros((ni.value, 12), np.float32), np.zeros((ni.value, 6), np.float32)
        self._check(self.L.snl_scene_copy_dbvh(self.h, None, _ptr(nodes), None, _ptr(el), _ptr(bb)))
        return nodes, el, bb

    def trace_primary_inst(self, cam: Camera, W, H):
        inst, tri = np.zeros((H, W), np.int32), np.zeros((H, W), np.int32)
        t, bary = np.zeros((H, W), np.float32), np.zeros((H, W, 2), np.float32)
        self._check(self.L.snl_trace_primary_inst(self.h, C.byref(cam), W, H, _ptr(inst), _ptr(tri), _ptr(t), _ptr(bary)))
        return inst, tri, t, bary

    def build_ms(self) -> float:
        return float(self.L.snl_scene_build_ms(self.h))

    def set_materials(self, materials, textures=()):
        """materials: list of dict(kind, flags, color, specular, opacity, texture); textures: list of
        dict(data=np.uint8 chain, width, height, mips)."""
        mats = (Material * max(len(materials), 1))()
        for n, m in enumerate(materials):
            mats[n].kind, mats[n].flags = int(m.get("kind", 0)), int(m.get("flags", 0))
            mats[n].color = (C.c_float * 3)(*m.get("color", (1, 1, 1)))
            mats[n].specular = (C.c_float * 3)(*m.get("specular", (0, 0, 0)))
            mats[n].opacity = float(m.get("opacity", 0.0))
            mats[n].texture = int(m.get("texture", -1))
            mats[n].mask_texture = int(m.get("mask_texture", -1))
        texs = (Texture * max(len(textures), 1))()
        keep = []
        for n, t in enumerate(textures):
            data = np.ascontiguousarray(t["data"], np.uint8)
            keep.append(data)
            texs[n].data, texs[n].size = data.ctypes.data, data.size
            texs[n].width, texs[n].height, texs[n].mips = int(t["width"]), int(t["height"]), int(t["mips"])
        self._check(self.L.snl_scene_set_materials(self.h, mats, len(materials), texs, len(textures)))

    def set_lights(self, lights, ambient=None):
        arr = (Light * max(len(lights), 1))(*lights)
        amb = None if ambient is None else (C.c_float * 3)(*ambient)
        self._check(self.L.snl_scene_set_lights(self.h, arr, len(lights), amb))

    def set_options(self, gvals):
        g = np.zeros(16, np.int32)
        for k, v in (gvals.items() if isinstance(gvals, dict) else enumerate(gvals)):
            g[k] = v
        self._check(self.L.snl_set_options(self.h, _ptr(g)))

    def set_reflection_depth(self, depth: int):
        """1 = the reference's cap (src/scene_trace.cpp:454), 2 = two bounces (BASELINE config 5; beyond the reference)."""
        self._check(self.L.snl_set_reflection_depth(self.h, depth))

    def load_obj(self, obj_path, mtl_path=None, tex_dir=None, flip_normals=0, build_flags=USE_SAH):
        self._check(self.L.snl_scene_load_obj(self.h, obj_path.encode(), (mtl_path or "").encode(),
                                              (tex_dir or "").encode(), flip_normals, build_flags))

    def set_host_scene(self, hs: HostScene):
        self._check(self.L.snl_scene_set_host_scene(self.h, hs.h))

    def counts(self):
        v = [C.c_int() for _ in range(5)]
        self._check(self.L.snl_scene_counts(self.h, *[C.byref(x) for x in v]))
        return dict(zip(("nodes", "tris", "matids", "depth", "materials"), [x.value for x in v]))

    def host_bvh(self):
        c = self.counts()
        nodes = np.zeros(c["nodes"], NODE_DTYPE)
        tris = np.zeros(c["tris"], TRI_DTYPE)
        shtris = np.zeros(c["tris"], SHTRI_DTYPE)
        matids = np.zeros(max(c["matids"], 1), np.int32)
        self._check(self.L.snl_scene_copy_host_bvh(self.h, _ptr(nodes), _ptr(tris), _ptr(shtris), _ptr(matids)))
        return nodes, tris, shtris, matids[: c["matids"]]

    def bbox(self):
        a, b = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self._check(self.L.snl_scene_bbox(self.h, _ptr(a), _ptr(b)))
        return a, b

    # ---- hot path ------------------------------------------------------------------------------------------------
    def trace_primary(self, cam: Camera, W, H):
        tri = np.zeros((H, W), np.int32)
        t = np.zeros((H, W), np.float32)
        bary = np.zeros((H, W, 2), np.float32)
        self._check(self.L.snl_trace_primary(self.h, C.byref(cam), W, H, _ptr(tri), _ptr(t), _ptr(bary)))
        return tri, t, bary

    def trace_rays(self, origins, dirs, mask=None, shared_origin=False):
        dirs = np.ascontiguousarray(dirs, np.float32)
        origins = np.ascontiguousarray(origins, np.float32)
        n = dirs.shape[0] if dirs.ndim == 2 else dirs.size // 3
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        tri = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        bary = np.zeros((n, 2), np.float32)
        self._check(self.L.snl_trace_rays(self.h, _ptr(origins), _ptr(dirs), _ptr(m), n, int(shared_origin), _ptr(tri),
                                          _ptr(t), _ptr(bary)))
        return tri, t, bary

    def trace_shadow(self, origin, dirs, tmax):
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.ascontiguousarray(tmax, np.float32).copy()
        o = np.ascontiguousarray(origin, np.float32)
        self._check(self.L.snl_trace_shadow(self.h, _ptr(o), _ptr(dirs), _ptr(out), out.size))
        return out

    def raytrace(self, origins, dirs, mask=None, shared_origin=False):
        dirs = np.ascontiguousarray(dirs, np.float32)
        origins = np.ascontiguousarray(origins, np.float32)
        n = dirs.shape[0] if dirs.ndim == 2 else dirs.size // 3
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        rgb = np.zeros((n, 3), np.float32)
        st = Stats()
        self._check(self.L.snl_raytrace(self.h, _ptr(origins), _ptr(dirs), _ptr(m), n, int(shared_origin), _ptr(rgb),
                                        C.byref(st)))
        return rgb, st.as_dict()

    def render(self, cam: Camera, W, H, tiles=None, out=None, rank=0):
        """Render() into a host image (H, W, 3) uint8 in ConvColor byte order. Returns (image, stats)."""
        if out is None:
            out = np.zeros((H, W, 3), np.uint8)
        t = None if tiles is None else np.ascontiguousarray(tiles, np.int32).reshape(-1, 4)
        st = Stats()
        self._check(self.L.snl_render(self.h, C.byref(cam), W, H, _ptr(t), 0 if t is None else len(t), _ptr(out),
                                      out.strides[0], rank, C.byref(st)))
        return out, st.as_dict()

    def render_farm(self, cam: Camera, W, H, coords, offsets=None, data=None, rank=0):
        """The farm overload of Render() (src/render.h:20-23): tiles `coords` (x, y, w, h) written at data + offsets[i]
        as planar x, y - x, z - x byte planes (src/render.cpp:140-166). offsets default to back-to-back tiles.
        Returns (data, offsets, stats)."""
        c = np.ascontiguousarray(coords, np.int32).reshape(-1, 4)
        sizes = c[:, 2].astype(np.int64) * c[:, 3] * 3
        if offsets is None:
            offsets = np.concatenate([[0], np.cumsum(sizes)[:-1]]) if len(c) else np.zeros(0)
        o = np.ascontiguousarray(offsets, np.int32)
        if data is None:
            data = np.zeros(int((o + sizes).max()) if len(c) else 1, np.uint8)
        st = Stats()
        self._check(self.L.snl_render_farm(self.h, C.byref(cam), W, H, _ptr(data), _ptr(c), _ptr(o), len(c), rank,
                                           C.byref(st)))
        return data, o, st.as_dict()

    def render_dev(self, cam: Camera, W, H, dev_ptr: int, pitch: int, tiles=None, rank=0):
        """Render() into device memory (e.g. a torch uint8 tensor's data_ptr()); asynchronous until sync()."""
        t = None if tiles is None else np.ascontiguousarray(tiles, np.int32).reshape(-1, 4)
        self._check(self.L.snl_render_dev(self.h, C.byref(cam), W, H, _ptr(t), 0 if t is None else len(t),
                                          C.c_void_p(dev_ptr), pitch, rank))

    def sync(self):
        st = Stats()
        self._check(self.L.snl_scene_sync(self.h, C.byref(st)))
        return st.as_dict()

    # ---- frames shared between ranks (CUDA IPC over NVLink) ---------------------------------------------------
    def frame_create(self, W, H):
        """Returns (device pointer, 64-byte IPC handle) of a zeroed W x H BGR8 frame on this scene's device."""
        p = C.c_void_p()
        h = C.create_string_buffer(64)
        self._check(self.L.snl_frame_create(self.h, W, H, C.byref(p), h))
        return p.value, h.raw

    def frame_open(self, handle: bytes):
        p = C.c_void_p()
        self._check(self.L.snl_frame_open(self.h, handle, C.byref(p)))
        return p.value

    def frame_close(self, ptr):
        self._check(self.L.snl_frame_close(self.h, C.c_void_p(ptr)))

    def frame_read(self, ptr, W, H, out=None):
        if out is None:
            out = np.zeros((H, W, 3), np.uint8)
        self._check(self.L.snl_frame_read(self.h, C.c_void_p(ptr), W, H, _ptr(out), out.strides[0]))
        return out

    def launch_count(self):
        return int(self.L.snl_scene_launch_count(self.h))

    def render_variant(self):
        return int(self.L.snl_scene_render_variant(self.h))

    def stream(self):
        return int(self.L.snl_scene_stream(self.h))


COMM_BLOB_BYTES = 96


def comm_plan(world: int, W: int, H: int, rank: int):
    """snl_comm_plan: (packets (n, 4) int32 {x, y, w, h} of `rank` in launch order, number of row bands, bandFirst[9])."""
    L = load_library()
    n, nb = C.c_int(), C.c_int()
    first = np.zeros(9, np.int32)
    r = L.snl_comm_plan(world, W, H, rank, None, C.byref(n), C.byref(nb), _ptr(first))
    if r:
        raise SnlError(r, L.snl_last_error().decode())
    packets = np.zeros((max(n.value, 1), 4), np.int32)
    L.snl_comm_plan(world, W, H, rank, _ptr(packets), C.byref(n), C.byref(nb), _ptr(first))
    return packets[: n.value], nb.value, first


class Comm:
    """Render() over the GPUs of one box (snl_comm_* / snl_render_multi, include/snail_b200.h): the frame lives on rank
    0, the packets are dealt to the ranks like the reference's taskMap and every rank's k_render stores its share into it
    over NVLink, one launch + one completion signal per row band.

      Comm.root(scene, world, W, H) -> (comm, blob)    rank 0; send `blob` (96 bytes) to the others by any transport
      Comm.attach(scene, rank, blob)                   ranks 1..world-1 (other processes)
      Comm.local([scene0, scene1, ...], W, H)          one process driving several devices
    render(cam, out) is collective: every rank calls it once per frame with the same camera."""

    def __init__(self, L, handle, scenes, rank):
        self.L, self.h, self.scenes, self.rank = L, handle, scenes, rank
        v = [C.c_int() for _ in range(4)]
        L.snl_comm_info(handle, *[C.byref(x) for x in v])
        self.world, self.W, self.H, self.local_ranks = [x.value for x in v]

    @staticmethod
    def _check(L, code):
        if code != 0:
            raise SnlError(code, L.snl_last_error().decode())

    @classmethod
    def root(cls, scene, world, W, H):
        blob = C.create_string_buffer(COMM_BLOB_BYTES)
        h = C.c_void_p()
        cls._check(scene.L, scene.L.snl_comm_create_root(scene.h, world, W, H, blob, C.byref(h)))
        return cls(scene.L, h, [scene], 0), blob.raw

    @classmethod
    def attach(cls, scene, rank, blob: bytes):
        h = C.c_void_p()
        cls._check(scene.L, scene.L.snl_comm_attach(scene.h, rank, blob, C.byref(h)))
        return cls(scene.L, h, [scene], rank)

    @classmethod
    def local(cls, scenes, W, H):
        arr = (C.c_void_p * len(scenes))(*[s.h for s in scenes])
        h = C.c_void_p()
        cls._check(scenes[0].L, scenes[0].L.snl_comm_create_local(arr, len(scenes), W, H, C.byref(h)))
        return cls(scenes[0].L, h, list(scenes), 0)

    def render(self, cam: Camera, out=None, host=True):
        """Returns (frame or None, stats). Rank 0: the (H, W, 3) frame in `out` (host=False: the frame stays on the
        device, see frame()) and TreeStats summed over the ranks; other ranks: (None, their own share)."""
        st = Stats()
        ptr, pitch = None, 0
        if self.rank == 0 and host:
            if out is None:
                out = np.zeros((self.H, self.W, 3), np.uint8)
            ptr, pitch = _ptr(out), out.strides[0]
        self._check(self.L, self.L.snl_render_multi(self.h, C.byref(cam), ptr, pitch, C.byref(st)))
        return (out if self.rank == 0 and host else None), st.as_dict()

    def frame(self):
        """(device pointer, pitch) of the frame the last render() produced (rank 0's address space)."""
        p, pitch = C.c_void_p(), C.c_size_t()
        self._check(self.L, self.L.snl_comm_frame(self.h, C.byref(p), C.byref(pitch)))
        return p.value, pitch.value

    def close(self):
        if getattr(self, "h", None):
            self.L.snl_comm_destroy(self.h)
            self.h = None


def bench_camera(bmin, bmax, inside=False):
    """SURVEY.md §8(d) bench camera: pos = C - 0.75*S*z (C when `inside`), identity axes, plane_dist 1; one point
    light at the camera, colour 1, radius S (what key 'j' creates, src/rtracer.cpp:270-281)."""
    bmin = np.asarray(bmin, np.float32)
    bmax = np.asarray(bmax, np.float32)
    c = (bmin + bmax) * np.float32(0.5)
    size = bmax - bmin
    s = np.float32(np.sqrt(np.float32(size[0] * size[0] + size[1] * size[1]) + np.float32(size[2] * size[2])))
    pos = c.copy()
    if not inside:
        pos[2] = c[2] - np.float32(0.75) * s
    return Camera(pos), Light(pos, (1, 1, 1), s)
