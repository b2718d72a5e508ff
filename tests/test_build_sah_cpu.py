"""The binned-SAH builder of snail_b200/csrc/snl_build_sah.cu, run as plain loops on the HOST (snl_host_build_binned_sah:
the very functors the device build launches as kernels), pinned without a GPU:
  * structure: every triangle sits in exactly one leaf, child boxes are the exact bounds of their triangles, leaves hold
    at most 32 triangles, the depth fits the traversal stack;
  * the tree answers rays like the reference's: converted to BVH::Node records and walked by the oracle, hit distances
    and frames equal the committed goldens (which the reference's own SAH tree produced);
  * quality: the surface-area cost (sum over leaves SA x count, the reference's costs: traverse 0, intersect 1,
    src/bvh/tree.cpp:70-71) against the reference's full-sweep tree."""
import json
import os

import numpy as np
import pytest

from conftest import DATA, GOLDEN, ROOT

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))


def find(sub, name):
    if name.startswith("@"):
        return os.path.join(DATA, name[1:])
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        p = os.path.join(d, sub, name)
        if os.path.exists(p):
            return p
    pytest.skip(f"{sub}/{name}: scene assets not present")


def sa(lo, hi):
    d = hi - lo
    return 2.0 * (d[..., 0] * (d[..., 1] + d[..., 2]) + d[..., 1] * d[..., 2])


def tri_bounds(tris):
    a, b, c = tris["a"], tris["a"] + tris["ba"], tris["a"] + tris["ca"]
    return np.minimum(np.minimum(a, b), c), np.maximum(np.maximum(a, b), c)


def walk(nodes):
    """Yields (node index, child k, box lo, box hi, link, depth) for every child slot reachable from node 0."""
    stack = [(0, 1)]
    while stack:
        i, depth = stack.pop()
        n = nodes[i]
        for k in range(2):
            b = n["b0"] if k == 0 else n["b1"]
            lo = np.array([b[0], b[2], n["bz"][2 * k]])
            hi = np.array([b[1], b[3], n["bz"][2 * k + 1]])
            link = int(n["link"][k])
            yield i, k, lo, hi, link, depth
            if link >= 0:
                stack.append((link, depth + 1))


def to_ref_nodes(sb, nodes, root_lo, root_hi):
    """64-byte two-child nodes -> BVH::Node records (own box, children adjacent at `sub`), for the oracle's walker."""
    out = np.zeros(2 * len(nodes) + 1, sb.NODE_DTYPE)
    out[0]["bmin"], out[0]["bmax"] = root_lo, root_hi
    nxt = 1
    todo = [(0, 0)]  # (gpu node, ref index)
    while todo:
        g, r = todo.pop()
        out[r]["sub"] = nxt
        for k, (_, _, lo, hi, link, _) in enumerate(_children(nodes[g])):
            c = nxt + k
            out[c]["bmin"], out[c]["bmax"] = lo, hi
            if link >= 0:
                todo.append((link, c))
            else:
                first, count = (~link) & ((1 << 26) - 1), ((~link) >> 26) + 1
                out[c]["sub"] = first | -0x80000000
                out[c]["axis"] = count & 0x7fff  # BVH::Node::count shares the bytes of {axis, firstNode}
                out[c]["first_node"] = count >> 16
        nxt += 2
    return out[:nxt]


def _children(n):
    for k in range(2):
        b = n["b0"] if k == 0 else n["b1"]
        yield 0, k, np.array([b[0], b[2], n["bz"][2 * k]]), np.array([b[1], b[3], n["bz"][2 * k + 1]]), int(n["link"][k]), 0


@pytest.mark.parametrize("name", ["box", "lancia", "barracks_tex", "feline"])
def test_host_run_of_the_device_builder(sb, name):
    import snail_oracle as O
    c = MANIFEST["cases"][name]
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    mtl = os.path.join(DATA, c["mtl"]) if c["mtl"] else None
    texdir = (os.path.dirname(find("data", "347.png")) + "/") if c["mtl"] else None
    hs = sb.HostScene(find("scenes", c["obj"]), mtl, texdir)
    ref_nodes, tris, shtris, matids = hs.bvh()
    n = len(tris)
    rng = np.random.default_rng(3)
    soup = rng.permutation(n)  # the builder must not depend on the input order being the reference's leaf order
    nodes, order, height = sb.host_build_binned_sah(tris[soup])
    order = soup[order]
    assert sorted(order.tolist()) == list(range(n))
    lo_t, hi_t = tri_bounds(tris[order])
    seen = np.zeros(n, np.int32)
    leaf_cost, max_depth, max_leaf = 0.0, 0, 0
    for _, _, lo, hi, link, depth in walk(nodes):
        max_depth = max(max_depth, depth)
        if link < 0:
            first, count = (~link) & ((1 << 26) - 1), ((~link) >> 26) + 1
            seen[first:first + count] += 1
            max_leaf = max(max_leaf, count)
            assert (lo == lo_t[first:first + count].min(axis=0)).all() and (hi == hi_t[first:first + count].max(axis=0)).all()
            leaf_cost += float(sa(lo, hi)) * count
    if n > 4:
        assert (seen == 1).all()
    assert max_leaf <= 32 and max_depth == height <= 94
    # the reference's own tree, same cost function
    rl = ref_nodes[(ref_nodes["sub"] & -0x80000000) != 0]
    ref_counts = (rl["axis"].astype(np.int64) & 0xffff) | (rl["first_node"].astype(np.int64) << 16)
    ref_cost = float((sa(rl["bmin"].astype(np.float64), rl["bmax"].astype(np.float64)) * ref_counts).sum())
    print(f"{name}: {n} triangles, {len(nodes)} inner nodes, height {height}, largest leaf {max_leaf}; leaf SA cost {leaf_cost:.6g} vs the "
          f"reference's full-sweep tree {ref_cost:.6g} ({leaf_cost / ref_cost:.3f}x)")
    assert leaf_cost <= 1.25 * ref_cost
    # same answers as the reference tree: the oracle walks the converted tree against the goldens
    rn = to_ref_nodes(sb, nodes, lo_t.min(axis=0), hi_t.max(axis=0))
    orc = O.OracleScene(rn, tris[order], shtris[order] if len(shtris) else None, matids, hs.materials(), hs.textures())
    W, H = int(g["size"][0]), int(g["size"][1])
    tri_id, t, _ = orc.trace_primary(g["camera"], W, H)
    hit = g["tri"] >= 0
    assert ((tri_id >= 0) == hit).mean() >= 0.9995
    both = hit & (tri_id >= 0)
    assert float((t[both] == g["t"][both]).mean()) >= 0.999
    same_tri = order[np.maximum(tri_id, 0)] == _golden_ids(sb, hs, g)
    assert float(same_tri[both].mean()) >= 0.999
    orc.set_gvals({6: 1, 7: 1})
    orc.set_lights([g["light"]])
    img, _ = orc.render(g["camera"], W, H)
    err = np.abs(img.astype(np.int16) - g["img_whitted"].astype(np.int16)).max(axis=2)
    err[g["undef_whitted"]] = 0
    assert float((err <= 1).mean()) >= 0.999
    orc.close()


def _golden_ids(sb, hs, g):
    """The golden's triangle ids index the reference tree's triangle order = the order HostScene hands out."""
    return np.maximum(g["tri"], 0)


def test_degenerate_inputs(sb):
    """One triangle; 100 identical triangles (coincident centres cannot be separated by a plane: dealt out alternately)."""
    t = np.zeros(1, sb.TRI_DTYPE)
    t["a"], t["ba"], t["ca"] = (0, 0, 0), (1, 0, 0), (0, 1, 0)
    nodes, order, height = sb.host_build_binned_sah(t)
    assert len(nodes) == 1 and height == 1 and order.tolist() == [0]
    assert int(nodes[0]["link"][0]) == int(nodes[0]["link"][1]) == ~0
    many = np.repeat(t, 100)
    nodes, order, height = sb.host_build_binned_sah(many)
    counts = [((~link) >> 26) + 1 for _, _, _, _, link, _ in walk(nodes) if link < 0]
    assert sum(counts) == 100 and max(counts) <= 32 and sorted(order.tolist()) == list(range(100))


def test_bench_scene_quality(sb):
    """The bench scene (268 K triangles): the binned-SAH tree's leaf cost against the reference's full-sweep tree, and
    the build's level count (= stream synchronisations of the device build)."""
    import time
    from snail_b200 import workloads
    try:
        w = workloads.resolve("sponza_like_whitted")
    except FileNotFoundError as e:
        pytest.skip(str(e))
    hs = sb.HostScene(w["obj_path"], w["mtl_path"], w["tex_dir"])
    ref_nodes, tris, _, _ = hs.bvh()
    t0 = time.time()
    nodes, order, height = sb.host_build_binned_sah(tris)
    dt = time.time() - t0
    leaf_cost = 0.0
    for _, _, lo, hi, link, _ in walk(nodes):
        if link < 0:
            leaf_cost += float(sa(lo, hi)) * (((~link) >> 26) + 1)
    rl = ref_nodes[(ref_nodes["sub"] & -0x80000000) != 0]
    ref_counts = (rl["axis"].astype(np.int64) & 0xffff) | (rl["first_node"].astype(np.int64) << 16)
    ref_cost = float((sa(rl["bmin"].astype(np.float64), rl["bmax"].astype(np.float64)) * ref_counts).sum())
    print(f"sponza_like: {len(tris)} triangles -> {len(nodes)} inner nodes, height {height} (reference tree depth {hs.depth}), host run "
          f"{dt:.2f} s; leaf SA cost {leaf_cost / ref_cost:.3f}x the reference's full-sweep tree")
    assert sorted(order.tolist()) == list(range(len(tris)))
    assert leaf_cost <= 1.05 * ref_cost and height <= 40
