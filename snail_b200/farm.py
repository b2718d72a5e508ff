"""Image-tile partitioning over the GPUs of one box and the compose of the tiles on rank 0 — the replacement of
the reference's MPI/TCP render farm (src/server.cpp:227-265 tile split, src/node.cpp + src/comm_mpi.cpp transport,
src/render.cpp:140-211 per-tile output). One process per GPU (torch.distributed); the scene is replicated, pixels
are partitioned, secondary rays never leave the GPU that spawned them.

Two compose paths:
  * "p2p"  — rank 0 owns the frame (snl_frame_create), every other rank maps it through CUDA IPC
             (snl_frame_open) and k_render stores its pixels straight into rank 0's HBM over NVLink/NVSwitch:
             compute and gather are ONE kernel, no staging buffer, no collective on the data path.
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


def tiles_for_rank(W: int, H: int, rank: int, world: int, seed: int = 1) -> np.ndarray:
    tiles = divide_image(W, H)
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


def gather_frame(dist, img, W: int, H: int, rank: int, world: int, seed: int = 1):
    """"gather" compose: every rank contributes the tiles it rendered into its own copy of `img` (torch uint8
    (H, W, 3), CPU for gloo / CUDA for NCCL); on return rank 0's `img` holds the whole frame."""
    import torch
    if world == 1:
        return img
    all_tiles = [tiles_for_rank(W, H, r, world, seed) for r in range(world)]
    size = max(tile_bytes(t) for t in all_tiles)
    mine = torch.zeros(size, dtype=torch.uint8, device=img.device)
    packed = pack_tiles(img, all_tiles[rank])
    mine[: packed.numel()] = packed
    bufs = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, bufs, dst=0)
    if rank == 0:
        for r in range(1, world):
            unpack_tiles(bufs[r], all_tiles[r], img)
    return img


def reduce_stats(dist, stats: dict, device=None) -> dict:
    """Sum of the per-rank TreeStats (replaces the per-node stats merge of src/server.cpp:403-418)."""
    import torch
    keys = [k for k in stats if k != "kernel_ms"]
    t = torch.tensor([float(stats[k]) for k in keys], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    out = {k: int(v) for k, v in zip(keys, t.tolist())}
    return out
