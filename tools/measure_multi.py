#!/usr/bin/env python
"""BASELINE configs 4 and 5 over the GPUs of one box through the C ABI's multi-GPU path (ptb_multi_*: ONE process, no torch,
no collective library; every GPU's resolve kernel stores its pixels into the root's frame over NVLink):
    config 4 (SCENE=stress): 10,000,000-triangle height field, 1920x1080, 4 spp
    config 5 (default):      262,279-triangle textured scene + HDR sky, 3840x2160, 64 spp
For every partition (tile groups G x sample groups N/G) the whole job is timed on the host around render + resolve into a host
buffer (after a warm-up job), and the frame is compared with ONE renderer's frame of the same samples.
usage: measure_multi.py N [G ...]      (G = tile groups; default: N, 2, 1 where they divide the sample count)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

STRESS = os.environ.get("SCENE", "sponza") == "stress"
SPP = int(os.environ.get("SPP", "4" if STRESS else "64"))
RES = (1920, 1080) if STRESS else (3840, 2160)
N = int(sys.argv[1]) if len(sys.argv) > 1 else capi.lib.ptb_device_count()
groups = [int(a) for a in sys.argv[2:]] or [N, 4, 2, 1]
t0 = time.perf_counter()
if STRESS:
    pos, nor, idx, mat = scenegen.heightfield_arrays(10_000_000)
    scene = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera(RES))
else:
    scene = ptb.Scene(scenegen.sponza("/tmp/measure_multi_sponza", res=RES)[0])
P = RES[0] * RES[1]
out = {"config": "4: 10M triangles, 1080p, 4 spp" if STRESS else "5: 262k textured + HDR sky, 3840x2160, 64 spp", "n_gpus": N, "spp": SPP,
       "scene_s": round(time.perf_counter() - t0, 1)}


def job(m, steps):
    m.render(1, steps)
    return m.resolve(capi.IMG_MEAN_RGB32F)


one = ptb.MultiRenderer(scene, [0])
job(one, SPP)
t0 = time.perf_counter(); ref = job(one, SPP); one_ms = (time.perf_counter() - t0) * 1e3
st = one.stats(0)
out["one_gpu_ms"] = round(one_ms, 2)
out["one_gpu_mpaths_per_s"] = round(P * SPP / one_ms / 1e3, 1)
out["lbvh_build_ms"] = round(st.lbvh_build_ms, 3); out["lbvh_max_depth"] = st.lbvh_max_depth
one.free()
for G in groups:
    if N % G or SPP % (N // G):
        continue
    S = N // G
    m = ptb.MultiRenderer(scene, list(range(N)), tile_groups=G)
    job(m, SPP // S)
    t0 = time.perf_counter(); img = job(m, SPP // S); ms = (time.perf_counter() - t0) * 1e3
    rel = float((np.abs(img - ref) / np.maximum(np.abs(ref), 1e-3)).max())
    out[f"G{G}xS{S}"] = {"ms": round(ms, 2), "mpaths_per_s": round(P * SPP / ms / 1e3, 1), "speedup": round(one_ms / ms, 2),
                         "bit_identical": bool(np.array_equal(img.view(np.uint32), ref.view(np.uint32))), "max_rel_err": rel,
                         "peer_stores": m.info.peer_stores}
    m.free()
print("MEASURE_MULTI " + json.dumps(out), flush=True)
