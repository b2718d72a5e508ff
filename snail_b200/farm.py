"""Image-tile partitioning over the GPUs of one box and the compose of the tiles on rank 0 — the replacement of
the reference's MPI/TCP render farm (src/server.cpp:227-265 tile split, src/node.cpp + src/comm_mpi.cpp transport,
src/render.cpp:140-211 per-tile output). One process per GPU (torch.distributed); the scene is replicated, pixels
are partitioned, secondary rays never leave the GPU that spawned them.

Two compose paths:
  * "p2p"  — snl_comm_* / snl_render_multi (snail_b200/csrc/snl_comm.cu, C++ behind the C ABI): rank 0 owns the
             frame, every other rank maps it through CUDA IPC and its k_render stores its share of the packets (a static
             deal like the reference's taskMap) straight into rank 0's HBM over NVLink/NVSwitch: compute and gather
             are ONE kernel, no staging buffer, no collective on the data path. This module only bootstraps it.
  * "gather" — every rank packs its tiles into a contiguous buffer, torch.distributed gathers the buffers on
             rank 0 (NCCL on GPUs, gloo in the CPU tests), rank 0 scatters them into the frame.
"""
from __future__ import annotations

import numpy as np

BLOCK_W, BLOCK_H = 16, 64  # blockWidth x blockHeight, src/rtbase_math.h:31-34


def divide_image(W: int, H: int, bw: int = BLOCK_W, bh: int = BLOCK_H) -> np.ndarray:
    """DivideImage (src/server.cpp:233): row-major list of (x, y, w, h). Unlike the reference (which refuses sizes
    that are not multiples of the block, server.cpp:227-231) the last row / column may be ragged."""
    out = [(x, y, min(bw, W - x), min(bh, H - y)) for y in range(0, H, bh) for x in range(0, W, bw)]
    return np.asarray(out, np.int32).reshape(-1, 4)


def assign_tiles(n_tiles: int, world: int, seed: int = 1) -> np.ndarray:
    """taskMap of src/server.cpp:238-248: every run of `world` consecutive tiles goes to all `world` ranks in a
    freshly shuffled order (keeps every rank's tile count within 1 of the others and decorrelates a rank from
    image regions). The reference uses std::random_shuffle with the C library's generator; here a fixed LCG so
    that every rank computes the same map without communication."""
    order = np.arange(world)
    state = (seed * 2654435761 + 12345) & ((1 << 64) - 1)
    out = np.zeros(n_tiles, np.int32)
    for n in range(n_tiles):
        if n % world == 0:
            for i in range(world - 1, 0, -1):  # Fisher-Yates
                state = (state * 6364136223846793005 + 1442695040888963407) & ((1 << 64) - 1)
                j = (state >> 33) % (i + 1)
                order[i], order[j] = order[j], order[i]
        out[n] = order[n % world]
    return out


def tiles_for_rank(W: int, H: int, rank: int, world: int, seed: int = 1, bw: int = BLOCK_W, bh: int = BLOCK_H) -> np.ndarray:
    tiles = divide_image(W, H, bw, bh)
    if world == 1:
        return tiles
    return tiles[assign_tiles(len(tiles), world, seed) == rank]


def tile_bytes(tiles: np.ndarray) -> int:
    return int((tiles[:, 2].astype(np.int64) * tiles[:, 3] * 3).sum())


def pack_tiles(img, tiles: np.ndarray):
    """Concatenates the BGR8 pixels of `tiles` (each tile row-major) into one flat uint8 buffer. Works on numpy
    arrays and torch tensors (H, W, 3)."""
    parts = [img[y:y + h, x:x + w].reshape(-1) for x, y, w, h in tiles.tolist()]
    if isinstance(img, np.ndarray):
        return np.concatenate(parts) if parts else np.zeros(0, np.uint8)
    import torch
    return torch.cat(parts) if parts else torch.zeros(0, dtype=torch.uint8, device=img.device)


def unpack_tiles(buf, tiles: np.ndarray, img) -> None:
    off = 0
    for x, y, w, h in tiles.tolist():
        n = w * h * 3
        img[y:y + h, x:x + w] = buf[off:off + n].reshape(h, w, 3)
        off += n


_INDEX_CACHE = {}


def tile_byte_index(tiles: np.ndarray, W: int) -> np.ndarray:
    """Byte offsets (into a tightly packed H x W x 3 frame) of the pixels of `tiles`, in pack_tiles order."""
    parts = []
    for x, y, w, h in tiles.tolist():
        rows = (np.arange(y, y + h, dtype=np.int64)[:, None] * W + np.arange(x, x + w, dtype=np.int64)[None, :]) * 3
        parts.append((rows[:, :, None] + np.arange(3, dtype=np.int64)).reshape(-1))
    return np.concatenate(parts) if parts else np.zeros(0, np.int64)


def gather_frame(dist, img, W: int, H: int, rank: int, world: int, seed: int = 1):
    """"gather" compose: every rank contributes the tiles it rendered into its own copy of `img` (torch uint8
    (H, W, 3), CPU for gloo / CUDA for NCCL); on return rank 0's `img` holds the whole frame. Packing and scattering
    are one indexed copy each (index tensors cached per (W, H, world, device))."""
    import torch
    if world == 1:
        return img
    key = (W, H, world, seed, str(img.device))
    if key not in _INDEX_CACHE:
        idx = [torch.from_numpy(tile_byte_index(tiles_for_rank(W, H, r, world, seed), W)).to(img.device)
               for r in range(world)]
        _INDEX_CACHE[key] = (idx, max(int(i.numel()) for i in idx))
    idx, size = _INDEX_CACHE[key]
    flat = img.view(-1)
    mine = torch.zeros(size, dtype=torch.uint8, device=img.device)
    mine[: idx[rank].numel()] = flat[idx[rank]]
    bufs = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, bufs, dst=0)
    if rank == 0:
        for r in range(1, world):
            flat[idx[r]] = bufs[r][: idx[r].numel()]
    return img


def reduce_stats(dist, stats: dict, device=None) -> dict:
    """Sum of the per-rank TreeStats (replaces the per-node stats merge of src/server.cpp:403-418)."""
    import torch
    keys = [k for k in stats if k != "kernel_ms"]
    t = torch.tensor([float(stats[k]) for k in keys], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    out = {k: int(v) for k, v in zip(keys, t.tolist())}
    return out


class Farm:
    """Render() over the ranks of one box: the counterpart of the reference's server loop (src/server.cpp:266-420)
    with `dist` = an initialised torch.distributed process group (used ONLY to hand rank 0's 96-byte bootstrap blob to
    the other ranks, and by the "gather" compose) and `scene` = this rank's snail_b200.Scene holding a replica of the
    scene. compose = "p2p" (default): a thin caller of snl_comm_* / snl_render_multi (C++ behind the ABI: static
    deal, peer stores into rank 0's frame, in-stream completion, banded device->host copy; no per-frame host
    collective, two frame slots so a rank can start frame k+1 while rank 0 still reads frame k). compose = "gather":
    static tile split + torch.distributed gather of packed tiles (the NCCL collective north_star words; kept as the
    cross-check and for boxes without peer access)."""

    def __init__(self, scene, dist, W: int, H: int, compose: str = "p2p", seed: int = 1):
        from .api import Comm
        self.scene, self.dist, self.W, self.H, self.compose = scene, dist, W, H, compose
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.seed = seed
        self.image = self.comm = self.tiles = None
        if compose == "p2p":
            blob = [None]
            if self.rank == 0:
                self.comm, blob[0] = Comm.root(scene, self.world, W, H)
            dist.broadcast_object_list(blob, src=0)
            if self.rank != 0:
                self.comm = Comm.attach(scene, self.rank, blob[0])
        else:
            import torch
            self.tiles = tiles_for_rank(W, H, self.rank, self.world, seed) if self.world > 1 else None
            self.image = torch.zeros((H, W, 3), dtype=torch.uint8, device=torch.device("cuda", scene.device))

    def render(self, cam, out=None, host=True):
        """Whole frame, collective. Returns (image on rank 0 else None, TreeStats: summed over the ranks on rank 0)."""
        if self.comm is not None:
            return self.comm.render(cam, out=out, host=host)
        import torch
        self.scene.render_dev(cam, self.W, self.H, self.image.data_ptr(), self.W * 3, tiles=self.tiles, rank=self.rank)
        st = self.scene.sync()
        gather_frame(self.dist, self.image, self.W, self.H, self.rank, self.world, self.seed)
        total = reduce_stats(self.dist, st, device=torch.device("cuda", self.scene.device))
        img = None
        if self.rank == 0 and host:
            img = self.scene.frame_read(self.image.data_ptr(), self.W, self.H, out=out)
        return img, total

    def close(self):
        self.dist.barrier()
        if self.comm is not None:
            if self.rank != 0:
                self.comm.close()  # unmap before the owner frees the segment
            self.dist.barrier()
            if self.rank == 0:
                self.comm.close()


def decode_farm_tiles(data: np.ndarray, coords, offsets, W: int, H: int, out: np.ndarray | None = None) -> np.ndarray:
    """What the client does with the farm wire format (DecompressTask, src/compression.cpp:112-142): planes
    x, y - x, z - x of every tile back into an (H, W, 3) frame in ConvColor byte order (byte 2 = x)."""
    img = np.zeros((H, W, 3), np.uint8) if out is None else out
    c = np.asarray(coords, np.int64).reshape(-1, 4)
    for (x, y, w, h), off in zip(c, np.asarray(offsets, np.int64)):
        p = np.asarray(data[off:off + 3 * w * h], np.uint8).reshape(3, h, w)
        img[y:y + h, x:x + w, 2] = p[0]
        img[y:y + h, x:x + w, 1] = p[1] + p[0]
        img[y:y + h, x:x + w, 0] = p[2] + p[0]
    return img
