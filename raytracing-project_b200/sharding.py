"""Frame sharding across GPUs (one process per GPU) and the final gather.

The reference is single-threaded; its pixels are independent once the position in the jitter
stream is known, so the frame shards with NO data-path collective:

* Standard mode: a tile is 156 consecutive pixels of the reference's loop order (y outer from the
  bottom row, x inner — tracer.cpp:288-299), i.e. exactly 8 MT19937 blocks; rank r renders tiles
  r, r+world, ... (interleaved, so cheap sky tiles and expensive model tiles mix).
* Paper mode: blocks of 16 image rows, interleaved; the 1-row halo of the edge detector
  (tracer.cpp:133-178) is recomputed locally.

The only exchange is the gather of each rank's finished RGBA8 pixels to rank 0
(`gather_frame`): NCCL over NVLink on GPUs, gloo on CPU for the tests.  This module is pure
index arithmetic + torch.distributed plumbing; it launches no kernels of its own.
"""
from __future__ import annotations

import numpy as np

TILE_PIXELS = 156      # RT_TILE_PIXELS
PAPER_ROWS = 16


def n_tiles(width: int, height: int) -> int:
    return (width * height + TILE_PIXELS - 1) // TILE_PIXELS


def owned_pixels(width: int, height: int, mode: int, rank: int, world: int) -> np.ndarray:
    """Output-layout pixel indices (row 0 = top, index = row*width + x) that `rank` renders, ascending in
    the order the rank produces them."""
    if not (0 <= rank < world):
        raise ValueError("need 0 <= rank < world")
    npx = width * height
    if mode == 1:   # paper: rows, no flip (tracer.cpp:279)
        rows = np.arange(height)
        mine = rows[(rows // PAPER_ROWS) % world == rank]
        return (mine[:, None] * width + np.arange(width)[None, :]).reshape(-1).astype(np.int64)
    p = np.arange(npx, dtype=np.int64)                       # loop-order pixel index
    p = p[((p // TILE_PIXELS) % world) == rank]
    y, x = p // width, p % width
    return (height - 1 - y) * width + x                      # row flip on store (tracer.cpp:297)


def gather_frame(local_rgba, width: int, height: int, mode: int, rank: int, world: int, group=None, index_cache=None):
    """Gather the ranks' slices of an [H, W, 4] uint8 torch tensor into the full frame on rank 0.

    Every rank passes its full-size buffer (only its own pixels are meaningful).  Only the owned
    pixels travel: 4 bytes per pixel in total across all ranks.  Returns the assembled tensor on
    rank 0 and None elsewhere."""
    import torch
    import torch.distributed as dist

    if world == 1:
        return local_rgba
    dev = local_rgba.device
    cache = index_cache if index_cache is not None else {}
    key = (width, height, mode, world, str(dev))
    if key not in cache:
        idx = [torch.from_numpy(owned_pixels(width, height, mode, r, world)).to(dev) for r in range(world)]
        cache[key] = (idx, max(int(i.numel()) for i in idx))
    idx, max_len = cache[key]
    flat = local_rgba.view(-1, 4).view(torch.int32).view(-1)            # one int32 per pixel
    mine = torch.zeros(max_len, dtype=torch.int32, device=dev)
    mine[: idx[rank].numel()] = flat[idx[rank]]
    if rank == 0:
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.gather(mine, parts, dst=0, group=group)
        out = torch.empty(width * height, dtype=torch.int32, device=dev)
        for r in range(world):
            out[idx[r]] = parts[r][: idx[r].numel()]
        return out.view(torch.uint8).view(height, width, 4)
    dist.gather(mine, None, dst=0, group=group)
    return None
