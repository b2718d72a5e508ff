"""CPU tests of the multi-GPU host logic (snail_b200/farm.py): tile partition + compose on rank 0, exercised with
world_size 2 and 3 over the gloo backend. The "rendering" of a rank is emulated by copying its tiles out of a golden
frame, so what is tested is exactly what is not a kernel: partition, packing, gather, scatter, stats reduce."""
import os
import socket

import numpy as np
import pytest

from conftest import GOLDEN
from snail_b200 import farm


def test_partition_covers_image_once_and_is_balanced():
    for W, H in ((1920, 1080), (800, 600), (3840, 2160), (96, 80)):
        for world in (1, 2, 3, 4, 8):
            cover = np.zeros((H, W), np.int32)
            counts = []
            for r in range(world):
                t = farm.tiles_for_rank(W, H, r, world)
                counts.append(len(t))
                for x, y, w, h in t.tolist():
                    assert x % 16 == 0 and w > 0 and h > 0 and x + w <= W and y + h <= H
                    cover[y:y + h, x:x + w] += 1
            assert (cover == 1).all()
            assert max(counts) - min(counts) <= 1


def test_assignment_is_deterministic_and_shuffled():
    a = farm.assign_tiles(1000, 8)
    assert (a == farm.assign_tiles(1000, 8)).all()
    assert not (a == np.arange(1000) % 8).all()  # not plain round-robin
    for n in range(0, 1000, 8):
        assert sorted(a[n:n + 8].tolist()) == list(range(8))


def test_pack_unpack_roundtrip():
    rng = np.random.default_rng(0)
    img = rng.integers(0, 255, (80, 96, 3), dtype=np.uint8)
    tiles = farm.tiles_for_rank(96, 80, 1, 3)
    buf = farm.pack_tiles(img, tiles)
    assert buf.size == farm.tile_bytes(tiles)
    out = np.zeros_like(img)
    farm.unpack_tiles(buf, tiles, out)
    for x, y, w, h in tiles.tolist():
        assert (out[y:y + h, x:x + w] == img[y:y + h, x:x + w]).all()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, frame, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W = frame.shape[:2]
    full = torch.from_numpy(frame)
    img = torch.zeros_like(full)
    mine = farm.tiles_for_rank(W, H, rank, world)
    for x, y, w, h in mine.tolist():  # this rank's "render": only its own tiles
        img[y:y + h, x:x + w] = full[y:y + h, x:x + w]
    farm.gather_frame(dist, img, W, H, rank, world)
    stats = farm.reduce_stats(dist, {"rays": 100 + rank, "primary": 10 * (rank + 1), "kernel_ms": 1.0})
    if rank == 0:
        q.put((bool((img == full).all()), stats))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_compose_over_gloo(world):
    import torch.multiprocessing as mp
    frame = np.load(os.path.join(GOLDEN, "barracks_tex.npz"))["img_whitted"]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, frame, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok, stats = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert stats["rays"] == sum(100 + r for r in range(world))
    assert stats["primary"] == sum(10 * (r + 1) for r in range(world))
