#!/usr/bin/env python
"""One rank of an N-way tile partition on one GPU: ms per frame of its share of the bench scene (what bounds the tile split's
scaling), optionally under ncu for the per-kernel launch list.  usage: tile_rank.py WORLD [FRAMES]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 50
scene = ptb.Scene(scenegen.fireplace("/tmp/tile_rank_fire", res=(1920, 1080))[0])
r = ptb.Renderer(scene, rank=0, world=world, stripe_rows=8)
r.render(1, 5); r.sync()
t0 = time.perf_counter(); r.render(6, frames); r.sync()
ms = (time.perf_counter() - t0) / frames * 1e3
print(f"world {world}: {r.owned_pixels} paths per frame, {ms:.3f} ms per frame ({r.width * r.height / world / ms / 1e3:.1f} Mpaths/s per GPU)", flush=True)
r.free()
