"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libsnailref_{sse,scalar}.so.

The libraries are the unmodified nadult/Snail reference compiled by oracle/build_ref.sh behind the
C API of oracle/ref_capi.cpp. Only tests/, tools/make_golden.py, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module; the product
(snail_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")

_libs: dict[str, C.CDLL] = {}


def available(variant: str = "scalar") -> bool:
    return os.path.exists(os.path.join(REF_DIR, f"libsnailref_{variant}.so"))


def lib(variant: str = "scalar") -> C.CDLL:
    if variant in _libs:
        return _libs[variant]
    path = os.path.join(REF_DIR, f"libsnailref_{variant}.so")
    # RTLD_LOCAL: the sse and scalar builds export the same C++ symbols (and the global gVals).
    L = C.CDLL(path, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    L.ref_scene_load.restype = C.c_void_p
    L.ref_scene_load.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int]
    L.ref_scene_free.argtypes = [C.c_void_p]
    L.ref_scene_counts.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 5
    for name in ("ref_scene_nodes", "ref_scene_tris", "ref_scene_shtris"):
        getattr(L, name).restype = C.c_void_p
        getattr(L, name).argtypes = [C.c_void_p]
    L.ref_scene_matids.argtypes = [C.c_void_p, _i32p]
    L.ref_scene_num_materials.argtypes = [C.c_void_p]
    L.ref_scene_material_info.argtypes = [C.c_void_p, C.c_int] + [C.POINTER(C.c_int)] * 4
    L.ref_scene_material_texture.restype = C.c_long
    L.ref_scene_material_texture.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.ref_scene_bbox.argtypes = [C.c_void_p, _f32p]
    L.ref_set_gvals.argtypes = [_i32p]
    L.ref_set_lights.argtypes = [C.c_void_p, _f32p, C.c_int]
    L.ref_render.restype = C.c_double
    L.ref_render.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, C.c_int, _u8p, _u32p]
    L.ref_render_tiles.restype = C.c_double
    L.ref_render_tiles.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _u8p, _i32p, _i32p, C.c_int, C.c_int, C.c_int, _u32p]
    L.ref_trace_primary.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _f32p, _f32p, _u32p]
    L.ref_trace_rays.argtypes = [C.c_void_p, _f32p, _f32p, C.c_void_p, C.c_int, C.c_int, _i32p, _f32p, _f32p]
    L.ref_trace_shadow.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, C.c_int]
    L.ref_raytrace.argtypes = [C.c_void_p, _f32p, _f32p, C.c_void_p, C.c_int, C.c_int, _f32p, _u32p]
    L.ref_primary_dirs.argtypes = [_f32p, C.c_int, C.c_int, _f32p]
    L.ref_scene_set_instances.argtypes = [C.c_void_p, _f32p, C.c_int]
    L.ref_dbvh_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.ref_dbvh_nodes.restype = C.c_void_p
    L.ref_dbvh_nodes.argtypes = [C.c_void_p]
    L.ref_dbvh_elements.argtypes = [C.c_void_p, _f32p]
    L.ref_trace_primary_inst.argtypes = [C.c_void_p, _f32p, C.c_int, C.c_int, _i32p, _i32p, _f32p, _f32p]
    L.ref_backend.restype = C.c_char_p
    _libs[variant] = L
    return L


# build flags of BVH::Construct (src/bvh/tree.h:21-25)
NO_SHADING_DATA, FAST_BUILD, USE_SAH = 1, 2, 4

NODE_DTYPE = np.dtype([("bmin", "<f4", 3), ("bmax", "<f4", 3), ("sub", "<i4"), ("axis", "<i2"), ("first_node", "<i2")])
TRI_DTYPE = np.dtype([("a", "<f4", 3), ("ba", "<f4", 3), ("ca", "<f4", 3), ("t0", "<f4"), ("it0", "<f4"),
                      ("pad", "<i4"), ("plane", "<f4", 4)])
SHTRI_DTYPE = np.dtype([("uv", "<f4", (3, 2)), ("nrm", "<f4", (3, 3)), ("mat", "<i4")])


def camera_array(pos, right=(1, 0, 0), up=(0, 1, 0), front=(0, 0, 1), plane_dist=1.0) -> np.ndarray:
    return np.array([*pos, *right, *up, *front, plane_dist], dtype=np.float32)


class RefScene:
    """Scene<BVH> built exactly as src/rtracer.cpp:505-586 does."""

    def __init__(self, obj_path: str, mtl_path: str | None = None, tex_dir: str = "", flip_normals: int = 0,
                 build_flags: int = USE_SAH, variant: str = "scalar"):
        self.L = lib(variant)
        self.variant = variant
        self.h = self.L.ref_scene_load(obj_path.encode(), (mtl_path or "").encode(), tex_dir.encode(),
                                       flip_normals, build_flags)
        n = [C.c_int() for _ in range(5)]
        self.L.ref_scene_counts(self.h, *[C.byref(x) for x in n])
        self.n_nodes, self.n_tris, self.n_shtris, self.n_matids, self.depth = [x.value for x in n]
        assert self.L.ref_sizeof_node() == 32 and self.L.ref_sizeof_triangle() == 64
        assert self.L.ref_sizeof_shtriangle() == 64

    def close(self):
        if self.h:
            self.L.ref_scene_free(self.h)
            self.h = None

    def _view(self, ptr, count, dtype):
        buf = (C.c_char * (count * dtype.itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=count).copy()

    def nodes(self):
        return self._view(self.L.ref_scene_nodes(self.h), self.n_nodes, NODE_DTYPE)

    def tris(self):
        return self._view(self.L.ref_scene_tris(self.h), self.n_tris, TRI_DTYPE)

    def shtris(self):
        return self._view(self.L.ref_scene_shtris(self.h), self.n_shtris, SHTRI_DTYPE)

    def matids(self):
        out = np.zeros(max(self.n_matids, 1), np.int32)
        self.L.ref_scene_matids(self.h, out)
        return out[: self.n_matids]

    def materials(self):
        out = []
        for i in range(self.L.ref_scene_num_materials(self.h)):
            v = [C.c_int() for _ in range(4)]
            self.L.ref_scene_material_info(self.h, i, *[C.byref(x) for x in v])
            flags, w, h, mips = [x.value for x in v]
            tex = None
            if w:
                nbytes = self.L.ref_scene_material_texture(self.h, i, None)
                tex = np.zeros(nbytes, np.uint8)
                self.L.ref_scene_material_texture(self.h, i, tex.ctypes.data)
            out.append(dict(flags=flags, tex_w=w, tex_h=h, tex_mips=mips, tex=tex))
        return out

    # ---- instanced form: Scene<DBVH> over ObjectInstances of this BVH (src/dbvh/tree.h) ------------------------
    def set_instances(self, instances):
        """instances: (n, 12) float32 {rotation rows, translation}. Set lights AFTER this or before, either works."""
        a = np.ascontiguousarray(np.asarray(instances, np.float32).reshape(-1, 12))
        self.L.ref_scene_set_instances(self.h, a, len(a))

    def dbvh(self):
        """(DBVH::nodes, elements (n, 12) in DBVH order, boxes (n, 6))."""
        nn, ni = C.c_int(), C.c_int()
        assert self.L.ref_dbvh_counts(self.h, C.byref(nn), C.byref(ni)) == 32
        nodes = self._view(self.L.ref_dbvh_nodes(self.h), nn.value, NODE_DTYPE)
        e = np.zeros((ni.value, 18), np.float32)
        self.L.ref_dbvh_elements(self.h, e)
        return nodes, e[:, :12].copy(), e[:, 12:].copy()

    def trace_primary_inst(self, cam, W, H):
        cam = np.ascontiguousarray(cam, np.float32)
        inst, tri = np.zeros((H, W), np.int32), np.zeros((H, W), np.int32)
        t, bary = np.zeros((H, W), np.float32), np.zeros((H, W, 2), np.float32)
        self.L.ref_trace_primary_inst(self.h, cam, W, H, inst, tri, t, bary)
        return inst, tri, t, bary

    def bbox(self):
        out = np.zeros(6, np.float32)
        self.L.ref_scene_bbox(self.h, out)
        return out[:3].copy(), out[3:].copy()

    def set_gvals(self, gvals):
        g = np.zeros(16, np.int32)
        for k, v in (gvals.items() if isinstance(gvals, dict) else enumerate(gvals)):
            g[k] = v
        self.L.ref_set_gvals(g)

    def set_lights(self, lights):
        arr = np.ascontiguousarray(np.array(lights, dtype=np.float32).reshape(-1, 7))
        self.L.ref_set_lights(self.h, arr.reshape(-1) if arr.size else np.zeros(7, np.float32), arr.shape[0])

    def render(self, cam, W, H, threads=1):
        img = np.zeros((H, W, 3), np.uint8)
        stats = np.zeros(4, np.uint32)
        ms = self.L.ref_render(self.h, cam, W, H, threads, img.reshape(-1), stats)
        return img, dict(intersects=int(stats[0]), iters=int(stats[1]), rays=int(stats[2]), skips=int(stats[3])), ms

    def render_tiles(self, cam, W, H, coords, offsets, nbytes, rank=0, threads=1):
        """Farm overload of Render(): planar delta-coded tiles at data + offsets[i] (src/render.cpp:140-166, 243-265)."""
        c = np.ascontiguousarray(coords, np.int32).reshape(-1)
        o = np.ascontiguousarray(offsets, np.int32)
        data = np.zeros(nbytes, np.uint8)
        stats = np.zeros(4, np.uint32)
        self.L.ref_render_tiles(self.h, cam, W, H, data, c, o, len(o), rank, threads, stats)
        return data, dict(intersects=int(stats[0]), iters=int(stats[1]), rays=int(stats[2]), skips=int(stats[3]))

    def trace_primary(self, cam, W, H):
        tri = np.zeros((H, W), np.int32)
        t = np.zeros((H, W), np.float32)
        bary = np.zeros((H, W, 2), np.float32)
        stats = np.zeros(4, np.uint32)
        self.L.ref_trace_primary(self.h, cam, W, H, tri.reshape(-1), t.reshape(-1), bary.reshape(-1), stats)
        return tri, t, bary

    def trace_rays(self, origins, dirs, mask=None, shared_origin=False):
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = dirs.shape[0]
        origins = np.ascontiguousarray(origins, np.float32)
        tri = np.zeros(n, np.int32)
        t = np.zeros(n, np.float32)
        bary = np.zeros((n, 2), np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self.L.ref_trace_rays(self.h, origins.reshape(-1), dirs.reshape(-1), None if m is None else m.ctypes.data,
                              n, int(shared_origin), tri, t, bary.reshape(-1))
        return tri, t, bary

    def trace_shadow(self, origin, dirs, tmax):
        dirs = np.ascontiguousarray(dirs, np.float32)
        out = np.ascontiguousarray(tmax, np.float32).copy()
        self.L.ref_trace_shadow(self.h, np.asarray(origin, np.float32), dirs.reshape(-1), out, dirs.shape[0])
        return out

    def raytrace(self, origins, dirs, mask=None, shared_origin=False):
        dirs = np.ascontiguousarray(dirs, np.float32)
        n = dirs.shape[0]
        origins = np.ascontiguousarray(origins, np.float32)
        rgb = np.zeros((n, 3), np.float32)
        stats = np.zeros(4, np.uint32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self.L.ref_raytrace(self.h, origins.reshape(-1), dirs.reshape(-1), None if m is None else m.ctypes.data,
                            n, int(shared_origin), rgb.reshape(-1), stats)
        return rgb, int(stats[2])

    def primary_dirs(self, cam, W, H):
        d = np.zeros((H, W, 3), np.float32)
        self.L.ref_primary_dirs(cam, W, H, d.reshape(-1))
        return d


def bench_camera(bmin, bmax, inside=False):
    """SURVEY.md §8(d) bench camera: pos = C - 0.75*S*z (or C when `inside`), identity axes, plane_dist 1;
    light = point light at the camera, colour 1, radius S (key 'j', src/rtracer.cpp:270-281)."""
    bmin = np.asarray(bmin, np.float32)
    bmax = np.asarray(bmax, np.float32)
    c = (bmin + bmax) * np.float32(0.5)
    size = bmax - bmin
    s = np.float32(np.sqrt(np.float32(size[0] * size[0] + size[1] * size[1]) + np.float32(size[2] * size[2])))
    pos = c.copy()
    if not inside:
        pos[2] = c[2] - np.float32(0.75) * s
    cam = camera_array(pos)
    light = np.array([*pos, 1.0, 1.0, 1.0, s], np.float32)
    return cam, light
