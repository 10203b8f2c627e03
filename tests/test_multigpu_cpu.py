"""CPU (gloo, world_size 2 and 3): tile partition + per-frame gather of the multi-GPU path."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cuda_path_tracer_b200 import multigpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, width, height, stripe, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pix = multigpu.owned_pixels(rank, world, stripe, width, height)
    # a tile whose value encodes the global pixel index, as the renderer's resolve would produce
    tile = torch.from_numpy(np.stack([pix, pix * 2, pix * 3], -1).astype(np.float32))
    frame = multigpu.gather_frame(tile, width, height, stripe, dist, dst=0)
    if rank == 0:
        q.put(frame.numpy())
    else:
        assert frame is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,width,height,stripe", [(2, 16, 20, 8), (3, 7, 33, 4), (2, 5, 3, 8), (2, 16, 32, 8), (3, 6, 28, 4)])
def test_gather_frame_gloo(world, width, height, stripe):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, width, height, stripe, q)) for r in range(world)]
    for p in procs:
        p.start()
    frame = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    idx = np.arange(width * height, dtype=np.float32)
    assert np.array_equal(frame, np.stack([idx, idx * 2, idx * 3], -1))


@pytest.mark.parametrize("world,stripe,height", [(1, 8, 10), (2, 8, 1080), (8, 8, 1080), (8, 32, 2160), (3, 5, 17)])
def test_partition_is_exact_cover(world, stripe, height):
    rows = np.concatenate([multigpu.owned_rows(r, world, stripe, height) for r in range(world)])
    assert np.array_equal(np.sort(rows), np.arange(height))
    sizes = [multigpu.owned_rows(r, world, stripe, height).size for r in range(world)]
    assert max(sizes) - min(sizes) <= stripe


def _reduce_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sr = multigpu.SampleReduce(4, 3, dist)
    its = [sr.iteration(k) for k in range(3)]
    # this rank's "resolved sum" after 3 steps: the sum of its iteration numbers in every channel
    sr.sum.fill_(float(sum(its)))
    frame = sr.reduce(3 * world)
    if rank == 0:
        q.put((its, frame.numpy()))
    else:
        assert frame is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sample_reduce_gloo(world):
    """Sample-index partition: the ranks' iterations tile 1..3*world exactly and the reduced frame is the mean."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_reduce_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    its, frame = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert its == [1, 1 + world, 1 + 2 * world]
    n = 3 * world
    assert np.allclose(frame, (n * (n + 1) / 2) / n) and frame.shape == (12, 3)


def test_hybrid_layout_covers_every_pixel_and_sample_once():
    """G tile groups x S sample groups (multigpu.hybrid_layout, the rank layout of PeerFrame and of ptb_multi_init): the ranks of one
    sample group tile the frame exactly once, and over `steps` steps the sample groups render iterations 1 .. steps * S exactly once."""
    width, height, stripe = 37, 50, 8
    for world, G in ((8, 2), (8, 8), (8, 1), (4, 2), (6, 3)):
        S = world // G
        seen = {}
        for rank in range(world):
            t, s, s_count = multigpu.hybrid_layout(rank, world, G)
            assert s_count == S and 0 <= t < G and 0 <= s < S
            pix = multigpu.owned_pixels(t, G, stripe, width, height)
            seen.setdefault(s, []).append(pix)
        for s, tiles in seen.items():
            allpix = np.sort(np.concatenate(tiles))
            assert np.array_equal(allpix, np.arange(width * height)), (world, G, s)
        steps = 5
        iters = sorted(step * S + s + 1 for step in range(steps) for s in range(S))      # PeerFrame.iteration
        assert iters == list(range(1, steps * S + 1))
    with pytest.raises(ValueError):
        multigpu.hybrid_layout(0, 8, 3)
