#!/usr/bin/env python
"""One staged frame of the bench scene with k_trace_extension launched ALONE once per bounce (ptb_time_raycast, the launch
bench.py's roofline times) -- the ncu target for that kernel's DRAM traffic: `-k regex:"k_trace_extension<0"`."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import capi, scenegen  # noqa: E402

r = ptb.Renderer(ptb.Scene(scenegen.fireplace("/tmp/ptb_staged_rc", res=(1920, 1080))[0]))
r.stage(capi.STAGE_RAYGEN, 1)
SHADOW = len(sys.argv) > 1 and sys.argv[1] == "shadow"      # also launch k_trace_shadow alone (ptb_time_kernel) once per bounce
for b in range(8):
    ms, rays = r.time_raycast(1)
    print(f"bounce {b}: {rays} rays {ms:.4f} ms", flush=True)
    if SHADOW:
        ms, rays = r.time_kernel("shadow", 1)
        print(f"bounce {b}: {rays} shadow rays {ms:.4f} ms", flush=True)
    r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
r.free()
