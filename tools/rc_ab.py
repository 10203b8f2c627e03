"""A/B of the RayCast knobs on the bench scene (GPU box): per-bounce k_trace_extension time, frame time, and bit-equality
of the path state after every TRACE stage against the first configuration.  Knobs are the PTB_* environment overrides
ptb_init reads (PTB_RC_BLOCKS, PTB_WIDE, PTB_FUSE_LOGIC, PTB_OVERLAP_SHADOW, PTB_SH_BLOCKS); pass them as
`PTB_WIDE=3,PTB_RC_BLOCKS=8` arguments.  (The persistent-lane scheduler this tool was first written for has been removed.)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

configs = [dict(PTB_WIDE=1), dict(PTB_WIDE=0), dict(PTB_WIDE=3), dict(PTB_WIDE=1, PTB_RC_BLOCKS=8)]
if len(sys.argv) > 1:
    configs = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[1:]]
scene = ptb.Scene(scenegen.fireplace("/tmp/rc_ab_fire", res=(1920, 1080))[0])
base = None
for cfg in configs:
    for k in list(os.environ):
        if k.startswith("PTB_"):
            del os.environ[k]
    os.environ.update({k: str(v) for k, v in cfg.items()})
    r = ptb.Renderer(scene)
    r.stage(capi.STAGE_RAYGEN, 1)
    ms_b, sh_b, state = [], [], []
    for b in range(8):
        ms, rays = r.time_raycast(5)
        ms_b.append(ms)
        sh_b.append(r.time_kernel("shadow", 5)[0])
        r.stage(capi.STAGE_TRACE)
        if b < 3:
            state.append([r.path_field(n).copy() for n in ("hit_geom_id", "ray_dir_dist", "hit_uv", "material_id")])
        r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    d1, i1 = r.accumulators()
    r.render(1, 3); r.sync()
    t0 = time.perf_counter(); r.render(1, 10); r.sync(); fr = (time.perf_counter() - t0) / 10 * 1e3
    d, i = r.accumulators()
    if base is None:
        base = (state, d, i)
        same = "base"
    else:
        q_ok = all(np.array_equal(a.view(np.uint8), b.view(np.uint8)) for sa, sb in zip(state, base[0]) for a, b in zip(sa[:1] + sa[3:], sb[:1] + sb[3:]))
        same = f"hits_equal={q_ok} radiance_equal={np.array_equal(d.view(np.uint32), base[1].view(np.uint32)) and np.array_equal(i.view(np.uint32), base[2].view(np.uint32))}"
    print(cfg, "ext ms/bounce:", " ".join(f"{m:.3f}" for m in ms_b), f"sum={sum(ms_b):.3f} shadow_sum={sum(sh_b):.3f} frame={fr:.3f} ms", same, flush=True)
    r.free()
