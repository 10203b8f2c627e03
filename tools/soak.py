#!/usr/bin/env python
"""Soak run on the GPU box: (1) device memory after N ptb_init / ptb_free cycles (every front-end) must return to where it
started; (2) a few thousand frames in one renderer: frame time stays flat and the accumulators stay finite."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import scenegen  # noqa: E402

rt = C.CDLL("libcudart.so.12")


def free_mb():
    f, t = C.c_size_t(), C.c_size_t()
    assert rt.cudaMemGetInfo(C.byref(f), C.byref(t)) == 0
    return f.value / 2**20


scene = ptb.Scene(scenegen.fireplace("/tmp/soak_fire", res=(1920, 1080))[0])
r = ptb.Renderer(scene); r.render(1, 1); r.sync(); r.free()      # context, module load, graph pools
base = free_mb()
for k in range(24):
    kw = [dict(), dict(megakernel=True), dict(gbuffer=True, denoiser=True), dict(rank=1, world=4), dict(env_nee=True), dict(use_graph=False)][k % 6]
    r = ptb.Renderer(scene, **kw)
    r.render(1, 2); r.sync(); r.resolve(2)
    r.free()
leak = base - free_mb()
print(f"free memory before {base:.0f} MB, after 24 init/free cycles {free_mb():.0f} MB (difference {leak:.1f} MB)", flush=True)

r = ptb.Renderer(scene)
FR, CH = int(os.environ.get("FRAMES", "3000")), 500
it, times = 1, []
for c in range(FR // CH):
    t0 = time.perf_counter()
    r.render(it, CH); r.sync()
    times.append((time.perf_counter() - t0) / CH * 1e3)
    it += CH
d, i = r.accumulators()
print("ms/frame per 500-frame chunk:", " ".join(f"{t:.3f}" for t in times), flush=True)
print("finite:", bool(np.isfinite(d).all() and np.isfinite(i).all()), "mean radiance", float((d.sum() + i.sum()) / (it - 1) / (d.size / 3) / 3), flush=True)
r.free()
assert leak < 64.0, "device memory did not come back after init/free cycles"
assert max(times) < 1.15 * min(times), "frame time drifted"
