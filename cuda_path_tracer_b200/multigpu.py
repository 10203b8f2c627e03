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


class FrameGather:
    """Per-frame gather of the ranks' tiles into the full frame on rank `dst`, with every buffer and
    index allocated once.  `send` ([pad, C], pad = largest tile) is the buffer a rank resolves its tile
    into (rows beyond its own tile are padding); `gather()` issues ONE collective and, on `dst`, places
    the stripes at their global rows: a strided copy per rank when the stripes are whole, an index copy
    otherwise."""

    def __init__(self, width, height, stripe_rows, dist, channels=3, dtype=None, device="cpu", dst=0):
        import torch
        self.dist, self.dst = dist, dst
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.width, self.height, self.stripe = width, height, stripe_rows
        dtype = dtype or torch.float32
        self.counts = [owned_rows(r, self.world, stripe_rows, height).size * width for r in range(self.world)]
        self.owned = self.counts[self.rank]
        pad = max(self.counts)
        self.send = torch.zeros((pad, channels), dtype=dtype, device=device)
        self.recv = self.frame = self.index = None
        self.whole = height % stripe_rows == 0
        if self.rank == dst:
            self.recv = [torch.empty_like(self.send) for _ in range(self.world)]
            self.frame = torch.empty((height * width, channels), dtype=dtype, device=device)
            if not self.whole:
                self.index = [torch.from_numpy(owned_pixels(r, self.world, stripe_rows, width, height)).to(device) for r in range(self.world)]

    @property
    def tile(self):
        """View of `send` that holds exactly this rank's pixels."""
        return self.send[:self.owned]

    def gather(self):
        if self.world == 1:
            return self.tile
        self.dist.gather(self.send, self.recv, dst=self.dst)
        if self.rank != self.dst:
            return None
        if self.whole:
            per = self.stripe * self.width
            stripes = self.frame.view(self.height // self.stripe, per, -1)
            for r in range(self.world):
                stripes[r::self.world].copy_(self.recv[r][:self.counts[r]].view(-1, per, stripes.shape[-1]))
        else:
            for r in range(self.world):
                self.frame[self.index[r]] = self.recv[r][:self.counts[r]]
        return self.frame


def gather_frame(local_tile, width, height, stripe_rows, dist, dst=0):
    """One-shot form: gather every rank's resolved tile (torch tensor [owned_pixels, C]) to rank `dst`.
    Returns the [height*width, C] frame on `dst`, None elsewhere.  (Loops should keep a FrameGather.)"""
    if dist.get_world_size() == 1:
        return local_tile
    g = FrameGather(width, height, stripe_rows, dist, channels=local_tile.shape[1], dtype=local_tile.dtype,
                    device=local_tile.device, dst=dst)
    g.tile.copy_(local_tile)
    return g.gather()


class SampleReduce:
    """Sample-index partition (north_star's second mode): every rank renders the FULL frame,
    rank r taking iterations r+1, r+1+world, ...  -- the seeds depend only on (pixel, iteration), so the union
    of the ranks' samples is exactly the sample set of a 1-GPU render of the same iterations.  Per displayed
    frame the ranks' accumulated sums are added (ONE reduce to `dst`) and divided by the total sample count.
    Unlike the tile partition this changes the order in which a pixel's samples are summed, so it agrees with
    the 1-GPU render to float rounding (tests: <= 1e-6 relative), not bit for bit; in exchange every GPU keeps
    full-size queues, which is what scales when the frame is small."""

    def __init__(self, width, height, dist, channels=3, dtype=None, device="cpu", dst=0):
        import torch
        self.dist, self.dst = dist, dst
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.sum = torch.zeros((height * width, channels), dtype=dtype or torch.float32, device=device)

    def iteration(self, step):
        """1-based iteration this rank renders at 0-based step `step`."""
        return step * self.world + self.rank + 1

    def reduce(self, total_samples):
        """`self.sum` holds this rank's resolved SUM image; returns the mean frame on `dst`, None elsewhere."""
        if self.world > 1:
            self.dist.reduce(self.sum, dst=self.dst, op=self.dist.ReduceOp.SUM)
            if self.rank != self.dst:
                return None
        return self.sum.div_(float(total_samples))
