"""Frame sharding across GPUs (one process per GPU) and the final gather.

The reference is single-threaded; its pixels are independent once the position in the jitter
stream is known, so the frame shards with NO data-path collective:

* Standard mode: a tile is 156 consecutive pixels of the reference's loop order (y outer from the
  bottom row, x inner — tracer.cpp:288-299), i.e. exactly 8 MT19937 blocks; rank r renders tiles
  r, r+world, ... (interleaved, so cheap sky tiles and expensive model tiles mix).
* Paper mode: blocks of 16 image rows, interleaved; the 1-row halo of the edge detector
  (tracer.cpp:133-178) is recomputed locally.

The only exchange is getting each rank's finished RGBA8 pixels to rank 0.  Two ways:

* `PeerFrame` (GPUs): ONE frame allocated on rank 0 and mapped into every other process through a
  CUDA IPC handle (include/rt_b200.h, rt_peer_frame_*).  Every rank passes that pointer to
  rt_render_device, so the render kernel's own pixel store lands in rank 0's HBM over NVLink and
  nothing is moved afterwards; `complete()` is a stream-ordered 4-byte all-reduce that tells rank 0
  every peer's kernel has retired.
* `gather_frame`: pack owned pixels, NCCL/gloo gather, scatter on rank 0.  Used by the CPU (gloo)
  tests and as the path when peer mapping is not available on a box.

This module is index arithmetic + torch.distributed plumbing; it launches no kernels of its own.
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


class PeerFrame:
    """An [H, W, 4] uint8 frame in rank 0's HBM that every rank renders into directly (NVLink peer stores).

    `ptr` is what each rank passes to Tracer.render_device as the RGBA8 target.  Ranks write disjoint
    pixels (`owned_pixels`), so there is no ordering between them; `complete()` orders rank 0's readers
    after every rank's kernel.  Collective: construct, `complete()` and `close()` on all ranks."""

    def __init__(self, rt, width: int, height: int, rank: int, world: int, device, group=None):
        import torch
        import torch.distributed as dist

        self.rt, self.rank, self.world, self.group = rt, rank, world, group
        self.width, self.height, self.device = width, height, device
        self.ptr = 0
        self.owner = rank == 0
        box = torch.zeros(rt.IPC_HANDLE_BYTES, dtype=torch.uint8, device=device)
        ok = torch.ones(1, dtype=torch.int32, device=device)
        err = None
        if self.owner:
            try:
                self.ptr, handle = rt.peer_frame_create(width * height * 4)
                box.copy_(torch.frombuffer(bytearray(handle), dtype=torch.uint8))
            except Exception as e:          # noqa: BLE001 - reported below, collectively
                err = e; ok.zero_()
        if world > 1:
            dist.broadcast(box, src=0, group=group)
        if not self.owner:
            try:
                self.ptr = rt.peer_frame_open(bytes(box.cpu().numpy().tobytes()))
            except Exception as e:          # noqa: BLE001
                err = e; ok.zero_()
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        self._flag = torch.zeros(1, dtype=torch.int32, device=device)
        if int(ok.item()) == 0:
            self.close()
            raise RuntimeError("peer frame unavailable on this box: %s" % (err if err is not None else "another rank failed to map it"))

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.height, self.width, 4), "typestr": "|u1", "data": (self.ptr, False), "version": 2}

    def tensor(self):
        """The assembled frame as a torch tensor (rank 0 only; valid after `complete()`)."""
        import torch

        if not self.owner:
            return None
        return torch.as_tensor(self, device=self.device)

    def complete(self) -> None:
        """Stream-ordered: after this, work queued on rank 0's current stream sees every rank's pixels."""
        import torch.distributed as dist

        if self.world > 1:
            dist.all_reduce(self._flag, group=self.group)

    def close(self) -> None:
        import torch
        import torch.distributed as dist

        multi = self.world > 1 and dist.is_initialized()
        if multi:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)          # nobody unmaps while a peer kernel may still store
        if self.ptr and not self.owner:
            self.rt.peer_frame_close(self.ptr, False)
            self.ptr = 0
        if multi:
            dist.barrier(group=self.group)          # the owner frees only after every mapping is gone
        if self.ptr:
            self.rt.peer_frame_close(self.ptr, True)
            self.ptr = 0
