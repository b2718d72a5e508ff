"""Multi-GPU Render() behind the C ABI (snl_comm_* / snl_render_multi) and its Python bootstrap (farm.Farm).
  * world = 1 (runs on any GPU box): the per-band k_render launches with band signals and the banded
    device->host copy against the plain snl_render: frames byte-identical, TreeStats equal, on every kernel flavour
    (plain, transparency, instanced) and on frame sizes that are not multiples of the packet / the 16-byte store;
  * the plain-C caller tests/c/render_multi_test.c (no Python, no torch) on min(2, GPUs) devices in ONE process;
  * 2 ranks in 2 processes (needs >= 2 GPUs, skipped otherwise): "p2p" (snl_comm over CUDA IPC) and the NCCL "gather"
    compose, two frames with two cameras each (a rank may start frame k+1 while rank 0 still reads frame k)."""
import os
import socket

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, obj, compose, q):
    import torch
    import torch.distributed as dist
    import snail_b200 as sb
    from snail_b200 import farm
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    sc = sb.Scene(rank)
    sc.load_obj(obj)
    cam, light = sb.bench_camera(*sc.bbox())
    sc.set_options({6: 1, 7: 1})
    sc.set_lights([light])
    W, H = 640, 360
    f = farm.Farm(sc, dist, W, H, compose=compose)
    cam2, _ = sb.bench_camera(*sc.bbox())
    cam2.pos[0] += 0.1 * light.radius
    same, rays, rays1 = True, 0, 0
    # no barrier between the frames: the other rank is free to run ahead into the next one
    frames = [f.render(c) for c in (cam, cam2, cam, cam2)]
    if rank == 0:
        for (img, st), c in zip(frames, (cam, cam2, cam, cam2)):
            single, st1 = sc.render(c, W, H)
            same = same and bool((img == single).all())
            rays += st["rays"]
            rays1 += st1["rays"]
        q.put((same, rays, rays1))
    f.close()
    sc.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("compose", ["p2p", "gather"])
def test_two_rank_render_matches_single_gpu(sb, compose):
    L = sb.load_library()
    if L.snl_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    obj = None
    for d in (os.path.join(ROOT, "assets"), os.path.join(ROOT, "oracle", "_ref")):
        if os.path.exists(os.path.join(d, "scenes", "lancia.obj")):
            obj = os.path.join(d, "scenes", "lancia.obj")
    if obj is None:
        pytest.skip("scene assets missing")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, obj, compose, q)) for r in range(2)]
    for p in procs:
        p.start()
    same, rays, rays1 = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert same
    # the packet-wide selector.All() test is per packet, packets are never split across ranks: same ray count
    assert rays == rays1


def _scene(sb, name, mtl=None):
    import os
    from conftest import DATA, TEXDIR, scene_path
    sc = sb.Scene(0)
    if name == "panes":
        sc.load_obj(os.path.join(DATA, "panes.obj"), os.path.join(DATA, "panes.mtl"), TEXDIR + "/")
    else:
        sc.load_obj(scene_path(name), os.path.join(DATA, mtl) if mtl else None, TEXDIR + "/" if mtl else None)
    return sc


@pytest.mark.parametrize("name,mtl,gv,W,H", [
    ("feline", None, {6: 1, 7: 1}, 1920, 1080),          # plain kernel, 8 bands, ragged last packet row
    ("barracks", "barracks.mtl", {6: 1, 7: 1}, 1000, 700),  # Tex + Uber materials (packet vote), pitch 3000 -> 3008, cw < 16
    ("panes", None, {6: 1, 7: 1}, 648, 360),                # transparency kernel
    ("box", None, {}, 40, 24),                              # 6 packets: fewer than resident CTAs, fewer rows than bands
])
def test_comm_world1_equals_single_launch(sb, name, mtl, gv, W, H):
    sc = _scene(sb, name, mtl)
    inside = name == "box"
    cam, light = sb.bench_camera(*sc.bbox(), inside=inside)
    if name == "panes":
        cam, light = sb.Camera((0.0137, 0.0211, -2.0)), sb.Light((0.5, 1.0, -2.0), (1.0, 1.0, 1.0), 12.0)
    light2 = sb.Light((light.pos[0] + 0.2, light.pos[1] + 0.1, light.pos[2]), (0.4, 0.7, 1.0), light.radius * 0.4)
    sc.set_options(gv)
    sc.set_lights([light, light2])  # the short-radius light turns the per-packet light-cull accounting on
    comm = sb.Comm.local([sc], W, H)
    assert (comm.world, comm.W, comm.H, comm.local_ranks) == (1, W, H, 1)
    for k in range(3):
        cam.pos[0] += 0.01 * light.radius * k
        ref, st1 = sc.render(cam, W, H)
        out = np.full((H, W + 3, 3), 9, np.uint8)[:, :W]  # host pitch > 3 * W
        img, st = comm.render(cam, out=out)
        assert (img == ref).all(), (name, k)
        for key in ("rays", "primary", "shadow", "reflection", "reflection_shadow", "transparency"):
            assert st[key] == st1[key], (name, k, key)
    _, st = comm.render(cam, host=False)  # frame stays on the device
    assert st["rays"] == st1["rays"]
    ptr, pitch = comm.frame()
    assert ptr and pitch >= 3 * W and pitch % 16 == 0
    comm.close()
    sc.close()


def test_comm_world1_instanced(sb):
    from conftest import GOLDEN
    import os
    g = np.load(os.path.join(GOLDEN, "lancia_inst.npz"))
    sc = _scene(sb, "lancia")
    sc.set_instances(g["instances"])
    cam = sb.Camera.from_array(g["camera"])
    lg = g["light"]
    sc.set_options({7: 1})
    sc.set_lights([sb.Light(lg[0:3], lg[3:6], lg[6])])
    W, H = 480, 270
    ref, st1 = sc.render(cam, W, H)
    comm = sb.Comm.local([sc], W, H)
    img, st = comm.render(cam)
    assert (img == ref).all() and st["rays"] == st1["rays"]
    sc.set_options({7: 1, 9: 1})
    with pytest.raises(sb.SnlError):
        comm.render(cam)  # gVals[9] is not available in snl_render_multi
    comm.close()
    sc.close()


def test_c_caller_renders_on_the_gpus_of_the_box(sb, tmp_path):
    """tests/c/render_multi_test.c: plain C against include/snail_b200.h, one process driving min(2, GPUs) devices."""
    import subprocess
    from conftest import scene_path
    L = sb.load_library()
    n = min(2, L.snl_device_count())
    exe = os.path.join(str(tmp_path), "render_multi_test")
    lib_dir = os.path.join(ROOT, "snail_b200")
    subprocess.run(["gcc", "-std=c11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c", "render_multi_test.c"), "-L", lib_dir, "-lsnail_b200", "-lm",
                    f"-Wl,-rpath,{lib_dir}", "-o", exe], check=True)
    for ngpu in sorted({1, n}):  # one process, ngpu devices
        r = subprocess.run([exe, scene_path("lancia"), str(ngpu), "648", "360", "4"], capture_output=True, text=True, timeout=300)
        print(r.stdout)
        assert r.returncode == 0 and r.stdout.strip().endswith("PASS"), r.stdout + r.stderr
