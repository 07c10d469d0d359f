"""Host-side sharding plans for one process per GPU (torch.distributed is only plumbing here).

The shader has no inter-pixel dependency (compute_shader.wgsl:333-364 touches only its own texel)
and pixel seeds use absolute coordinates (wgsl:309-310), so both plans need no data-path collective:
  rows   — one image, rank i owns the contiguous row block rt_shard_rows(height, world, i)
  bands  — one image, rank i owns every world-th band of band_rows rows (rt_create_bands): balanced whatever the
           image holds where (geometry rows cost more than sky rows; bounce rays cluster on the lower rows)
  frames — a camera batch, frame k goes to rank k % world
The only exchange is the final gather of finished images to rank 0.
"""
import numpy as np

from .api import shard_bands, shard_rows


def rows_for_rank(height: int, rank: int, world: int):
    return shard_rows(height, world, rank)


def frames_for_rank(n_frames: int, rank: int, world: int):
    return list(range(rank, n_frames, world))


def gather_row_blocks(dist, block, height: int, rank: int, world: int, dst: int = 0):
    """Gather per-rank row blocks (tensors of shape [rows_i, width, C], any device the backend
    supports) into the full image on `dst`.  Blocks may differ in height, so the exchange is
    point-to-point (send/recv), which is what the NCCL gather decomposes to anyway."""
    import torch
    if world == 1:
        return block
    if rank == dst:
        parts = []
        for src in range(world):
            b, c = shard_rows(height, world, src)
            if src == dst:
                parts.append(block)
            else:
                buf = torch.empty((c,) + tuple(block.shape[1:]), dtype=block.dtype, device=block.device)
                dist.recv(buf, src=src)
                parts.append(buf)
        return torch.cat(parts, dim=0)
    dist.send(block.contiguous(), dst=dst)
    return None


def band_rows_for_rank(height: int, band_rows: int, rank: int, world: int) -> np.ndarray:
    """Image rows of an interleaved shard, in the shard's window order (what Renderer(bands=...).read_accum returns)."""
    r = np.arange(shard_bands(height, band_rows, world, rank))
    return (r // band_rows * world + rank) * band_rows + r % band_rows


def gather_bands(dist, block, height: int, band_rows: int, rank: int, world: int, dst: int = 0):
    """Gather interleaved shards (tensors [rows_i, width, C] in window order) into the full image on `dst`."""
    import torch
    if rank != dst:
        dist.send(block.contiguous(), dst=dst)
        return None
    full = torch.empty((height,) + tuple(block.shape[1:]), dtype=block.dtype, device=block.device)
    for src in range(world):
        rows = torch.from_numpy(band_rows_for_rank(height, band_rows, src, world)).to(block.device)
        if src == dst:
            full[rows] = block
        else:
            buf = torch.empty((len(rows),) + tuple(block.shape[1:]), dtype=block.dtype, device=block.device)
            dist.recv(buf, src=src)
            full[rows] = buf
    return full
