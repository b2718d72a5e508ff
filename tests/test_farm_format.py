"""The farm wire format: Render(scene, camera, resx, resy, data, coords, offsets, options, rank,
threads) writes every 16x64 tile as three byte planes x, y - x, z - x (src/render.cpp:140-166, 243-265) and the client
undoes it (DecompressTask, src/compression.cpp:112-142). Goldens tests/golden/farm_<scene>.npz are outputs of the
reference's own farm overload (tools/make_golden.py farm_cases): set A = every tile, shuffled back-to-back offsets,
gVals[8] rank tint (rank 5); set B = every other tile, gaps between the tiles, no tint.

CPU half: the decoder and the C oracle against the reference's tiles. GPU half: snl_render_farm against the tiles."""
import json
import os

import numpy as np
import pytest

from conftest import DATA, GOLDEN, ROOT

MANIFEST = json.load(open(os.path.join(GOLDEN, "manifest.json")))
FARM = list(MANIFEST.get("farm", {}))


def find(sub, name):
    if name.startswith("@"):
        return os.path.join(DATA, name[1:])
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        p = os.path.join(d, sub, name)
        if os.path.exists(p):
            return p
    pytest.skip(f"{sub}/{name}: scene assets not present")


def paths(name):
    c = MANIFEST["cases"][name]
    mtl = os.path.join(DATA, c["mtl"]) if c["mtl"] else None
    texdir = (os.path.dirname(find("data", "347.png")) + "/") if c["mtl"] else None
    return find("scenes", c["obj"]), mtl, texdir


def tile_mask(coords, W, H):
    m = np.zeros((H, W), bool)
    for x, y, w, h in coords:
        m[y:y + h, x:x + w] = True
    return m


def test_decoder_inverts_the_plane_layout():
    from snail_b200 import farm
    rng = np.random.default_rng(0)
    W, H = 64, 128
    img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    coords = farm.divide_image(W, H)[::-1]
    offsets = np.arange(len(coords)) * 3072
    data = np.zeros(len(coords) * 3072, np.uint8)
    for (x, y, w, h), o in zip(coords, offsets):  # the encoder of src/render.cpp:147-164, spelled out
        t = img[y:y + h, x:x + w]
        data[o:o + w * h] = t[:, :, 2].reshape(-1)
        data[o + w * h:o + 2 * w * h] = (t[:, :, 1] - t[:, :, 2]).reshape(-1)
        data[o + 2 * w * h:o + 3 * w * h] = (t[:, :, 0] - t[:, :, 2]).reshape(-1)
    assert (farm.decode_farm_tiles(data, coords, offsets, W, H) == img).all()


@pytest.mark.parametrize("name", FARM)
def test_c_oracle_vs_reference_farm_tiles(sb, name):
    """Set B (no tint) decoded = the frame the plain-C restatement renders, on the tiles the reference wrote."""
    import snail_oracle as O
    from snail_b200 import farm
    g = np.load(os.path.join(GOLDEN, "farm_" + name + ".npz"))
    W, H = int(g["size"][0]), int(g["size"][1])
    obj, mtl, texdir = paths(name)
    hs = sb.HostScene(obj, mtl, texdir)
    nodes, tris, shtris, matids = hs.bvh()
    orc = O.OracleScene(nodes, tris, shtris, matids, hs.materials(), hs.textures())
    orc.set_gvals({int(k): int(v) for k, v in g["gvals_B"]})
    orc.set_lights([g["light"]])
    img, _ = orc.render(g["camera"], W, H)
    dec = farm.decode_farm_tiles(g["data_B"], g["coords_B"], g["offsets_B"], W, H)
    m = tile_mask(g["coords_B"], W, H) & ~g["undef_B"]
    err = np.abs(img.astype(np.int16) - dec.astype(np.int16)).max(axis=2)
    assert float((err[m] <= 1).mean()) >= 0.999
    assert (dec[~tile_mask(g["coords_B"], W, H)] == 0).all()  # tiles that were not rendered stay empty
    orc.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", FARM)
@pytest.mark.parametrize("tag", ["A", "B"])
def test_render_farm_vs_reference_tiles(sb, name, tag):
    from snail_b200 import farm
    g = np.load(os.path.join(GOLDEN, "farm_" + name + ".npz"))
    W, H = int(g["size"][0]), int(g["size"][1])
    obj, mtl, texdir = paths(name)
    sc = sb.Scene(0)
    sc.load_obj(obj, mtl, texdir)
    lg = g["light"]
    sc.set_lights([sb.Light(lg[0:3], lg[3:6], lg[6])])
    sc.set_options({int(k): int(v) for k, v in g["gvals_" + tag]})
    cam = sb.Camera.from_array(g["camera"])
    coords, offsets, gold = g["coords_" + tag], g["offsets_" + tag], g["data_" + tag]
    data = np.full(len(gold), 0xAB, np.uint8)
    data, _, st = sc.render_farm(cam, W, H, coords, offsets, data=data, rank=int(g["rank_" + tag]))
    # bytes between the tiles are not touched
    used = np.zeros(len(gold), bool)
    for (x, y, w, h), o in zip(coords, offsets):
        used[o:o + 3 * w * h] = True
    assert (data[~used] == 0xAB).all()
    ours = farm.decode_farm_tiles(data, coords, offsets, W, H)
    ref = farm.decode_farm_tiles(gold, coords, offsets, W, H)
    m = tile_mask(coords, W, H) & ~g["undef_" + tag]
    err = np.abs(ours.astype(np.int16) - ref.astype(np.int16)).max(axis=2)
    frac = float((err[m] <= 1).mean())
    exact = float((data[used] == gold[used]).mean())
    print(f"farm {name}/{tag}: within 1/255 on {frac*100:.3f}% of the tile pixels, {exact*100:.2f}% of the wire bytes identical")
    assert frac >= 0.999
    assert abs(st["rays"] - int(g["rays_" + tag])) <= max(4 * int(g["undef_" + tag].sum()), 0.002 * int(g["rays_" + tag]))
    # the wire bytes are a pure re-arrangement of the frame snl_render produces for the same tiles
    img, _ = sc.render(cam, W, H, tiles=coords, rank=int(g["rank_" + tag]))
    for x, y, w, h in coords:
        assert (img[y:y + h, x:x + w] == ours[y:y + h, x:x + w]).all()
    sc.close()


@pytest.mark.gpu
def test_render_farm_ragged_and_unaligned(sb):
    """Tiles the reference leaves unwritten (not 16x64) use the same layout here; offsets that are not 4-byte aligned take
    the byte path of k_farm_pack."""
    from snail_b200 import farm
    obj, mtl, texdir = paths("box")
    sc = sb.Scene(0)
    sc.load_obj(obj, mtl, texdir)
    g = np.load(os.path.join(GOLDEN, "farm_box.npz"))
    lg = g["light"]
    sc.set_lights([sb.Light(lg[0:3], lg[3:6], lg[6])])
    cam = sb.Camera.from_array(g["camera"])
    W, H = 104, 90  # ragged right column (8 wide) and bottom row (26 high)
    coords = farm.divide_image(W, H)
    sizes = coords[:, 2] * coords[:, 3] * 3
    offsets = np.concatenate([[1], 1 + np.cumsum(sizes + 1)[:-1]]).astype(np.int32)  # odd offsets, 1-byte gaps
    data, _, _ = sc.render_farm(cam, W, H, coords, offsets)
    img, _ = sc.render(cam, W, H)
    assert (farm.decode_farm_tiles(data, coords, offsets, W, H) == img).all()
    with pytest.raises(sb.SnlError):
        sc.render_farm(cam, W, H, [[0, 0, 16, 200]], [0])
    sc.close()
