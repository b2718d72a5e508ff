#!/usr/bin/env python
"""Authors data/panes.obj + data/panes.mtl: a small multi-material test scene for the transparency pass
(Scene::TraceTransparency, src/scene_trace.cpp:468-483,620-634) and TransparentMaterial. None of the reference's
scenes has more than one material or a transparent one, so parity for those paths needs an authored scene; the
golden images are still produced by the reference itself (tools/make_golden.py, case "panes").

Layout (camera on -z looking along +z, see bench_camera):
  room             floor, ceiling, three walls: default material (faces before the first usemtl); closed, so no ray misses
  back wall        TexMaterial (map_Kd 347.png)
  glass_red        Uber, d 0.5   pane at z = 1
  glass_green      Uber, d 0.3   pane at z = 2, overlaps glass_red -> chains of two transparent surfaces
  solid_blue       Uber, d 1.0   has fTransparency but opacity == 1: filtered by `opacity < 1`
  lattice          TransparentMaterial (map_Kd tex3.png + map_d 1669.png)
  mosaic           4-pixel-wide cells cycling glass_red / glass_green / lattice / wall / default at z = 3: one 16-lane
                   shading block sees several transparent materials (the `transSel[b4 + q] = mask[q]` overwrite quirk)
The textures are the reference's own data/*.png (copied to assets/data by __graft_entry__.build()).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

MTL = """# Authored for snail_b200 by tools/make_panes.py (transparency / TransparentMaterial parity scene)
newmtl wall
Kd 1.0 1.0 1.0
map_Kd 347.png

newmtl glass_red
Kd 0.8 0.2 0.2
Ks 0.3 0.3 0.3
d 0.5

newmtl glass_green
Kd 0.2 0.8 0.3
Ks 0.1 0.4 0.1
d 0.3

newmtl solid_blue
Kd 0.2 0.3 0.9
Ks 0.2 0.2 0.2
d 1.0

newmtl lattice
Kd 1.0 1.0 1.0
map_Kd tex3.png
map_d 1669.png
"""


class Obj:
    def __init__(self):
        self.v, self.vt, self.lines = [], [], []
        self.cur = None

    def usemtl(self, name):
        if name != self.cur:
            self.lines.append(f"usemtl {name}")
            self.cur = name

    def quad(self, p0, p1, p2, p3, uv=((0, 0), (1, 0), (1, 1), (0, 1)), nrm=1):
        """p0..p3 counter-clockwise; two triangles, per-vertex uv, shared normal index `nrm`."""
        base, tb = len(self.v), len(self.vt)
        self.v += [p0, p1, p2, p3]
        self.vt += list(uv)
        i = [base + k + 1 for k in range(4)]
        t = [tb + k + 1 for k in range(4)]
        self.lines.append(f"f {i[0]}/{t[0]}/{nrm} {i[1]}/{t[1]}/{nrm} {i[2]}/{t[2]}/{nrm}")
        self.lines.append(f"f {i[0]}/{t[0]}/{nrm} {i[2]}/{t[2]}/{nrm} {i[3]}/{t[3]}/{nrm}")

    def zquad(self, x0, x1, y0, y1, z, uvs=1.0):
        self.quad((x0, y0, z), (x1, y0, z), (x1, y1, z), (x0, y1, z),
                  uv=((0, 0), (uvs, 0), (uvs, uvs), (0, uvs)), nrm=1)

    def text(self):
        out = ["# Authored for snail_b200 by tools/make_panes.py", "mtllib panes.mtl", "vn 0 0 -1", "vn 0 1 0", "vn 0 -1 0", "vn 1 0 0",
               "vn -1 0 0", "vn 0 0 1"]
        out += ["v %.6f %.6f %.6f" % p for p in self.v]
        out += ["vt %.6f %.6f" % t for t in self.vt]
        return "\n".join(out + self.lines) + "\n"


def main():
    out_dir = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "data")
    o = Obj()
    # closed room, default material (no usemtl yet): floor, ceiling, side walls, wall behind the camera. No ray can
    # miss, which keeps the reference's uninitialised-normal quirk (tools/make_golden.py) out of this scene.
    o.quad((-4, -3, -4), (4, -3, -4), (4, -3, 4), (-4, -3, 4), nrm=2)
    o.quad((-4, 3, -4), (-4, 3, 4), (4, 3, 4), (4, 3, -4), nrm=3)
    o.quad((-4, -3, -4), (-4, -3, 4), (-4, 3, 4), (-4, 3, -4), nrm=4)
    o.quad((4, -3, -4), (4, 3, -4), (4, 3, 4), (4, -3, 4), nrm=5)
    o.quad((-4, -3, -4), (-4, 3, -4), (4, 3, -4), (4, -3, -4), nrm=6)
    o.usemtl("wall")
    o.zquad(-4, 4, -3, 3, 4.0, uvs=2.0)
    o.usemtl("glass_red")
    o.zquad(-3.5, -0.5, -2.0, 2.0, 1.0)
    o.usemtl("glass_green")
    o.zquad(-2.5, 1.5, -2.5, 1.5, 2.0)
    o.usemtl("solid_blue")
    o.zquad(0.5, 2.0, 0.0, 2.5, 1.5)
    o.usemtl("lattice")
    o.zquad(1.0, 3.8, -2.5, 0.5, 0.5)
    # mosaic strip: cells 0.2 wide, 5 materials in turn, two rows offset by one material
    names = ["glass_red", "glass_green", "lattice", "wall", None]
    for row, (y0, y1) in enumerate(((2.0, 2.45), (2.45, 2.9))):
        # the default material can only be used before the first usemtl, so `None` cells are left empty: the back wall
        # (TexMaterial) shows through there
        for c in range(36):
            name = names[(c + 2 * row) % len(names)]
            if name is None:
                continue
            o.usemtl(name)
            x0 = -3.6 + 0.2 * c
            o.zquad(x0, x0 + 0.2, y0, y1, 3.0, uvs=0.25)
    with open(os.path.join(out_dir, "panes.obj"), "w") as f:
        f.write(o.text())
    with open(os.path.join(out_dir, "panes.mtl"), "w") as f:
        f.write(MTL)
    print("panes:", len(o.lines), "lines ->", out_dir)


if __name__ == "__main__":
    main()
