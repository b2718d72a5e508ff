"""TEST INFRASTRUCTURE — builds and binds oracle/snail_oracle.c (plain-C per-ray restatement of the reference's hot
path). Only tests/, __graft_entry__ and bench.py's cpu_baseline leg may import this; the product never does."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "snail_oracle.c")
LIB = os.path.join(HERE, "libsnail_oracle.so")

_lib = None


def build(force=False):
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c11", "-o", LIB + ".tmp", SRC, "-lm"],
                       check=True)
        os.replace(LIB + ".tmp", LIB)
    return LIB


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_create.restype = C.c_void_p
        vp, i = C.c_void_p, C.c_int
        _lib.oracle_create.argtypes = [vp, i, vp, vp, i, vp, i, vp, i, vp, i]
        _lib.oracle_destroy.argtypes = [vp]
        _lib.oracle_set_gvals.argtypes = [vp, vp]
        _lib.oracle_set_reflection_depth.argtypes = [vp, i]
        _lib.oracle_set_lights.argtypes = [vp, vp, i]
        _lib.oracle_counts.argtypes = [vp, vp]
        _lib.oracle_trace_primary.argtypes = [vp, vp, i, i, vp, vp, vp]
        _lib.oracle_trace_rays.argtypes = [vp, vp, vp, vp, i, i, vp, vp, vp]
        _lib.oracle_trace_shadow.argtypes = [vp, vp, vp, vp, i]
        _lib.oracle_render.argtypes = [vp, vp, i, i, vp]
        _lib.oracle_render_ex.argtypes = [vp, vp, i, i, vp, vp]
    return _lib


class _Mat(C.Structure):
    _fields_ = [("kind", C.c_int), ("flags", C.c_int), ("color", C.c_float * 3), ("specular", C.c_float * 3),
                ("opacity", C.c_float), ("texture", C.c_int), ("mask_texture", C.c_int)]


class _Tex(C.Structure):
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t), ("width", C.c_int), ("height", C.c_int), ("mips", C.c_int)]


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """nodes / tris / shtris / matids in the reference layouts (numpy structured arrays of 32 / 64 / 64 bytes),
    materials = list of dict(kind, flags, color, specular, opacity, texture), textures = list of dict(data, width,
    height, mips) — exactly what snail_b200.HostScene or snailref.RefScene hand out."""

    def __init__(self, nodes, tris, shtris, matids, materials=(), textures=()):
        self.L = lib()
        self.nodes, self.tris = np.ascontiguousarray(nodes), np.ascontiguousarray(tris)
        self.shtris = np.ascontiguousarray(shtris) if shtris is not None and len(shtris) else None
        self.matids = np.ascontiguousarray(matids, np.int32)
        self.mats = (_Mat * max(len(materials), 1))()
        for n, m in enumerate(materials):
            self.mats[n].kind, self.mats[n].flags = int(m["kind"]), int(m["flags"])
            self.mats[n].color = (C.c_float * 3)(*m["color"])
            self.mats[n].specular = (C.c_float * 3)(*m["specular"])
            self.mats[n].opacity, self.mats[n].texture = float(m["opacity"]), int(m["texture"])
            self.mats[n].mask_texture = int(m.get("mask_texture", -1))
        self.texs = (_Tex * max(len(textures), 1))()
        self._keep = []
        for n, t in enumerate(textures):
            d = np.ascontiguousarray(t["data"], np.uint8)
            self._keep.append(d)
            self.texs[n].data, self.texs[n].size = d.ctypes.data, d.size
            self.texs[n].width, self.texs[n].height, self.texs[n].mips = int(t["width"]), int(t["height"]), int(t["mips"])
        self.h = self.L.oracle_create(_p(self.nodes), len(self.nodes), _p(self.tris), _p(self.shtris), len(self.tris),
                                      _p(self.matids), len(self.matids), C.cast(self.mats, C.c_void_p), len(materials),
                                      C.cast(self.texs, C.c_void_p), len(textures))

    def close(self):
        if self.h:
            self.L.oracle_destroy(self.h)
            self.h = None

    def set_gvals(self, gvals):
        g = np.zeros(16, np.int32)
        for k, v in (gvals.items() if isinstance(gvals, dict) else enumerate(gvals)):
            g[k] = v
        self.L.oracle_set_gvals(self.h, _p(g))

    def set_reflection_depth(self, depth):
        """1 = the reference's cap; 2 = `cache.reflections < 2` (oracle/_ref/libsnailref_depth2.so is the reference patched so)."""
        self.L.oracle_set_reflection_depth(self.h, int(depth))

    def set_lights(self, lights):
        a = np.ascontiguousarray(np.array(lights, np.float32).reshape(-1, 7))
        self.L.oracle_set_lights(self.h, _p(a) if a.size else None, a.shape[0])

    def counts(self):
        c = np.zeros(7, np.uint64)
        self.L.oracle_counts(self.h, _p(c))
        return dict(zip(("node_visits", "tri_tests", "primary", "shadow", "reflection", "reflection_shadow", "transparency"),
                        [int(x) for x in c]))

    def trace_primary(self, cam, W, H):
        cam = np.ascontiguousarray(cam, np.float32)
        tri, t, bary = np.zeros((H, W), np.int32), np.zeros((H, W), np.float32), np.zeros((H, W, 2), np.float32)
        self.L.oracle_trace_primary(self.h, _p(cam), W, H, _p(tri), _p(t), _p(bary))
        return tri, t, bary

    def trace_rays(self, o, d, mask=None, shared_origin=False):
        o, d = np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)
        n = d.shape[0]
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        tri, t, bary = np.zeros(n, np.int32), np.zeros(n, np.float32), np.zeros((n, 2), np.float32)
        self.L.oracle_trace_rays(self.h, _p(o), _p(d), _p(m), n, int(shared_origin), _p(tri), _p(t), _p(bary))
        return tri, t, bary

    def trace_shadow(self, origin, d, tmax):
        d = np.ascontiguousarray(d, np.float32)
        out = np.ascontiguousarray(tmax, np.float32).copy()
        o = np.ascontiguousarray(origin, np.float32)
        self.L.oracle_trace_shadow(self.h, _p(o), _p(d), _p(out), out.size)
        return out

    def render(self, cam, W, H, undef=None):
        """undef: optional (H, W) uint8 array that receives 1 where the reference's pixel depends on an uninitialised
        Sample (snail_oracle.c shade_block)."""
        cam = np.ascontiguousarray(cam, np.float32)
        img = np.zeros((H, W, 3), np.uint8)
        self.L.oracle_render_ex(self.h, _p(cam), W, H, _p(img), _p(undef))
        c = self.counts()
        c["rays"] = c["primary"] + c["shadow"] + c["reflection"] + c["reflection_shadow"] + c["transparency"]
        return img, c
