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


class PeerUnavailable(RuntimeError):
    """The ranks cannot map rank 0's buffer (no CUDA IPC / peer access between the GPUs): use FrameGather / SampleReduce."""


def hybrid_layout(rank, world, tile_groups):
    """(tile rank, sample group, sample groups) of `rank` when the world is cut into `tile_groups` stripe sets x world / tile_groups
    sample groups: ranks 0 .. G-1 are the tiles of sample group 0, and so on."""
    if tile_groups < 1 or world % tile_groups:
        raise ValueError("tile_groups must divide the world size")
    return rank % tile_groups, rank // tile_groups, world // tile_groups


class PeerFrame:
    """One process per GPU, NO collective on the data path: rank 0 owns the full-frame buffer (ptb_gather), exports it as a
    CUDA IPC handle, and every rank's resolve kernel stores the pixels it owns straight into it over NVLink
    (ptb_resolve_into).  The ranks then only need a barrier; a 4-byte all-reduce on the renderers' stream provides it in
    stream order.  `tile_groups` G cuts the frame into G interleaved stripe sets, the other world / G factor splits the
    sample indices (group s renders iterations s+1, s+1+S, ...); rank r is tile (r % G) of sample group (r // G).
    Two buffers alternate by step, so a rank that runs ahead writes step k + 1 while rank 0 may still fold step k; a buffer
    is written again only after the barrier of the step in between, which rank 0 joins after that fold (stream order).
    The NCCL gather / reduce classes above stay as the route for boxes without peer access (PeerUnavailable)."""

    def __init__(self, width, height, dist, device_index, tile_groups=0, stripe_rows=8):
        import torch
        from .capi import PtbError
        from .wavefront import Gather
        self.dist, self.torch = dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.tile_groups = tile_groups or self.world
        self.tile_rank, self.sample_group, self.sample_groups = hybrid_layout(self.rank, self.world, self.tile_groups)
        self.stripe_rows, self.width, self.height = stripe_rows, width, height
        self.token = torch.zeros(1, dtype=torch.float32, device=torch.device("cuda", device_index))
        self.gathers, ok = [], 1.0
        handles = [None]
        try:
            if self.rank == 0:
                self.gathers = [Gather(device_index, width, height, self.sample_groups) for _ in range(2)]
                handles[0] = [g.export() for g in self.gathers]
        except PtbError:
            ok = 0.0
        dist.broadcast_object_list(handles, src=0)
        if self.rank != 0:
            try:
                self.gathers = [Gather.open(device_index, h, width, height, self.sample_groups) for h in (handles[0] or [])]
                ok = 1.0 if len(self.gathers) == 2 else 0.0
            except PtbError:
                ok = 0.0
        flag = torch.tensor([ok], dtype=torch.float32, device=self.token.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if float(flag[0]) < 1.0:
            self.free()
            raise PeerUnavailable("rank 0's frame buffer cannot be mapped by every rank")

    def renderer_kwargs(self):
        """Keyword arguments of wavefront.Renderer for this rank's share of the frame."""
        return dict(rank=self.tile_rank, world=self.tile_groups, stripe_rows=self.stripe_rows)

    def iteration(self, step):
        """1-based iteration this rank renders at 0-based step `step`."""
        return step * self.sample_groups + self.sample_group + 1

    def submit(self, renderer, step):
        """Store this rank's pixels into rank 0's buffer of `step` and join the barrier (both on the current CUDA stream)."""
        self.gathers[step & 1].resolve_from(renderer, self.sample_group)
        self.dist.all_reduce(self.token)

    def combine_device(self, device_ptr, steps, fmt, stream, step):
        """Rank 0, after submit(step): fold the sample groups and write the frame (steps * sample_groups samples per pixel)."""
        self.gathers[step & 1].combine_device(device_ptr, steps * self.sample_groups, fmt, stream)

    def free(self):
        for g in self.gathers:
            g.free()
        self.gathers = []
