"""Host-side sharding plans for one process per GPU (torch.distributed is only plumbing here).

The shader has no inter-pixel dependency (compute_shader.wgsl:333-364 touches only its own texel)
and pixel seeds use absolute coordinates (wgsl:309-310), so both plans need no data-path collective:
  rows   — one image, rank i owns the contiguous row block rt_shard_rows(height, world, i)
  frames — a camera batch, frame k goes to rank k % world
The only exchange is the final gather of finished images to rank 0.
"""
from .api import shard_rows


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
