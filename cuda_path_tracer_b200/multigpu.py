"""Multi-GPU frame partition: one process per GPU, interleaved row stripes, one gather per frame.

The reference is single-GPU (src/main.cpp:168).  Paths are independent and their RNG seeds depend
only on (global pixel index, iteration) (src/kernels/ray_gen.cu:41), so rank r renders the rows y
with (y // stripe_rows) % world == r against its own replica of the scene/LBVH, and the union of
the ranks' tiles is bit-identical to a 1-GPU render.  The only exchange is a gather of the resolved
tiles to rank 0 (torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests); no reduction,
so no float is ever summed across GPUs.
"""
import numpy as np


def owned_rows(rank, world, stripe_rows, height):
    """Rows of the frame rank `rank` renders, ascending (must match TileMap in csrc/cuda/common.cuh)."""
    y = np.arange(height)
    return y[(y // stripe_rows) % world == rank]


def owned_pixels(rank, world, stripe_rows, width, height):
    rows = owned_rows(rank, world, stripe_rows, height)
    return (rows[:, None] * width + np.arange(width)[None, :]).reshape(-1)


def gather_frame(local_tile, width, height, stripe_rows, dist, dst=0):
    """Gather every rank's resolved tile (torch tensor [owned_pixels, C]) to rank `dst` and place the
    rows at their global position.  Returns the [height*width, C] frame on `dst`, None elsewhere.
    Tiles differ in size by at most one stripe, so they are padded to a common length for all_gather
    -- one collective per frame."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    if world == 1:
        return local_tile
    counts = [owned_rows(r, world, stripe_rows, height).size * width for r in range(world)]
    pad = max(counts)
    buf = local_tile
    if local_tile.shape[0] != pad:
        buf = torch.zeros((pad,) + tuple(local_tile.shape[1:]), dtype=local_tile.dtype, device=local_tile.device)
        buf[:local_tile.shape[0]] = local_tile
    out = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, out, dst=dst)
    if rank != dst:
        return None
    frame = torch.empty((height * width,) + tuple(local_tile.shape[1:]), dtype=local_tile.dtype, device=local_tile.device)
    for r in range(world):
        idx = torch.from_numpy(owned_pixels(r, world, stripe_rows, width, height)).to(frame.device)
        frame[idx] = out[r][:counts[r]]
    return frame
