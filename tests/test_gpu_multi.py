"""Multi-GPU parity (needs >= 2 GPUs, skipped otherwise): the tile-partitioned render over 2 ranks, composed on
rank 0 through CUDA-IPC peer memory ("p2p") and through an NCCL gather, is byte-identical to the 1-GPU frame."""
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
    img, st = f.render(cam)
    if rank == 0:
        single, st1 = sc.render(cam, W, H)
        q.put((bool((img == single).all()), st["rays"], st1["rays"]))
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
