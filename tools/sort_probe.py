"""How much would re-ordering secondary rays buy?  Takes the bounce-1..3 extension rays of the bench scene and times
the batched closest-hit kernel (ptb_stats.last_batch_ms, device time) on: queue order, random order, and several sort keys."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen


def morton10(p, lo, hi):
    q = np.clip(((p - lo) / (hi - lo) * 1024).astype(np.int64), 0, 1023)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return (spread(q[:, 0]) << 2) | (spread(q[:, 1]) << 1) | spread(q[:, 2])


scene = ptb.Scene(scenegen.fireplace("/tmp/sp_fire", res=(1920, 1080))[0])
r = ptb.Renderer(scene)
r.stage(capi.STAGE_RAYGEN, 1)
for b in range(4):
    q = r.queue("extension")
    ori, dirs = r.path_field("ray_ori")[q].copy(), r.path_field("ray_dir_dist")[q].copy()
    if b >= 1:
        lo, hi = ori[:, :3].min(0), ori[:, :3].max(0) + 1e-3
        m = morton10(ori[:, :3], lo, hi)
        octant = ((dirs[:, 0] > 0).astype(np.int64) << 2) | ((dirs[:, 1] > 0).astype(np.int64) << 1) | (dirs[:, 2] > 0).astype(np.int64)
        # direction on an 8x8 octahedral-ish grid (6 bits)
        dq = np.clip(((dirs[:, :3] * 0.5 + 0.5) * 4).astype(np.int64), 0, 3)
        dkey = (dq[:, 0] << 4) | (dq[:, 1] << 2) | dq[:, 2]
        orders = {"queue": np.arange(len(q)), "random": np.random.RandomState(0).permutation(len(q)),
                  "origin_morton30": np.argsort(m, kind="stable"),
                  "octant|origin_morton30": np.argsort((octant << 30) | m, kind="stable"),
                  "origin_morton15|dir6|rest": np.argsort(((m >> 15) << 21) | (dkey << 15) | (m & 0x7fff), kind="stable"),
                  "dir6|origin_morton30": np.argsort((dkey << 30) | m, kind="stable"),
                  "origin_morton21|dir6": np.argsort(((m >> 9) << 6) | dkey, kind="stable")}
        for name, o in orders.items():
            best = 1e9
            for _ in range(3):
                r.trace_closest(ori[o], dirs[o])
                best = min(best, r.stats().last_batch_ms)
            print(f"bounce {b} {len(q)} rays  {name:28s} {best:.3f} ms  {len(q) / best / 1e3:.0f} Mrays/s", flush=True)
    r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
