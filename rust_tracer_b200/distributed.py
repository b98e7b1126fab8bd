"""Multi-GPU plumbing: one process per GPU (torch.distributed), interleaved row tiles, one
gather of the quantised RGB8 tiles to rank 0 -- the only exchange step of the path
(SURVEY 8e).  Rendering itself has no collective: pixels are independent
(reference src/render.rs:31-38 writes one cell per iteration).

NCCL has no native gather to ragged sizes, so the gather is point-to-point: every rank
!= 0 sends its tile buffer, rank 0 receives each into place (batched isend/irecv, which
NCCL runs as one grouped send/recv over NVLink).  On CPU (gloo) the same code runs the
world_size-2 tests.
"""
import torch
import torch.distributed as dist

from .api import tile_rows


_row_index_cache = {}


def _row_index(height, rank, world, rows_per_tile, device):
    """Device tensor of the rows rank `rank` renders (cached: the gather runs every frame)."""
    key = (height, rank, world, rows_per_tile, str(device))
    idx = _row_index_cache.get(key)
    if idx is None:
        idx = torch.from_numpy(tile_rows(height, rank, world, rows_per_tile)).to(device, torch.long)
        _row_index_cache[key] = idx
    return idx


def assemble_rows(tiles, height, width, world, rows_per_tile):
    """tiles[r]: [rows_r, width, 3] uint8 tensor of rank r -> [height, width, 3] image."""
    full = torch.empty((height, width, 3), dtype=tiles[0].dtype, device=tiles[0].device)
    for r in range(world):
        full[_row_index(height, r, world, rows_per_tile, full.device)] = tiles[r]
    return full


def gather_rows(tile, height, width, rank, world, rows_per_tile, device=None):
    """Gather every rank's rendered rows to rank 0; returns the full image there, None elsewhere.

    tile: this rank's [rows, width, 3] uint8 tensor (device tensor for NCCL, CPU for gloo)."""
    if world == 1:
        return tile
    device = device if device is not None else tile.device
    if rank == 0:
        bufs = [tile]
        ops = []
        for r in range(1, world):
            n = len(tile_rows(height, r, world, rows_per_tile))
            b = torch.empty((n, width, 3), dtype=tile.dtype, device=device)
            bufs.append(b)
            if n:
                ops.append(dist.P2POp(dist.irecv, b, r))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return assemble_rows(bufs, height, width, world, rows_per_tile)
    if tile.numel():
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, tile.contiguous(), 0)]):
            req.wait()
    return None
