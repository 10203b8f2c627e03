"""Experiment (GPU box): does running K tile-partitioned renderers CONCURRENTLY on one GPU (each on its own stream, with
its own queues and graph) beat one renderer over the whole frame?  The partitions are the bit-identical interleaved
stripes of ptb_config.rank / world, so the union of the K accumulators equals the single renderer's; the question is
only whether one partition's RayCast fills the tails and Shade gaps of the others.  Prints ms per whole frame.
Measured: 12.32 ms for one renderer, 12.72-12.76 for two, 12.86 for three, 13.09 for four."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

scene = ptb.Scene(scenegen.fireplace("/tmp/dual_fire", res=(1920, 1080))[0])
N = int(os.environ.get("FRAMES", "30"))


def run(world, stripe):
    rs = [ptb.Renderer(scene, rank=k, world=world, stripe_rows=stripe) for k in range(world)]
    for it in range(1, 4):
        for r in rs:
            r.render(it, 1)
    for r in rs:
        r.sync()
    t0 = time.perf_counter()
    for it in range(4, 4 + N):
        for r in rs:
            r.render(it, 1)
    for r in rs:
        r.sync()
    ms = (time.perf_counter() - t0) / N * 1e3
    for r in rs:
        r.free()
    return ms


base_ms = run(1, 8)
print(f"world=1: {base_ms:.3f} ms/frame", flush=True)
for world, stripe in ((2, 8), (2, 64), (2, 270), (3, 8), (4, 8), (4, 135)):
    ms = run(world, stripe)
    print(f"world={world} stripe={stripe}: {ms:.3f} ms/frame ({base_ms / ms:.3f}x)", flush=True)
