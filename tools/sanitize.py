#!/usr/bin/env python
"""Small renders through every front-end of the library in one process (a crash / sticky-error check; written as a
compute-sanitizer target, but compute-sanitizer is closed on this GPU pool, so it runs plain: any illegal access surfaces
as a CUDA error code from the next ptb_* call and raises).
Covers the LBVH build, the wavefront graph and the staged API, both traversal trees, the megakernel, the G-buffer /
SVGF path, env NEE, a tile partition and the LBVH heat map on three small scenes."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import capi, scenegen  # noqa: E402

scenes = {"cornell": scenegen.cornell("/tmp/san_cornell", res=(96, 96)),
          "fire": scenegen.fireplace("/tmp/san_fire", res=(160, 90), scale=0.12)[0],
          "sponza": scenegen.sponza("/tmp/san_sponza", res=(128, 72), scale=0.12, tex_size=64)[0]}
for name, path in scenes.items():
    scene = ptb.Scene(path)
    for wide in ("0", "1", "3"):
        os.environ["PTB_WIDE"] = wide
        r = ptb.Renderer(scene)
        r.render(1, 2); r.sync()
        r.resolve(2)
        r.free()
    os.environ.pop("PTB_WIDE")
    r = ptb.Renderer(scene, use_graph=False)
    r.stage(capi.STAGE_RAYGEN, 1)
    for b in range(3):
        r.time_raycast(1)
        r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    r.free()
    r = ptb.Renderer(scene, megakernel=True)
    r.render(1, 2); r.sync(); r.resolve(2); r.free()
    r = ptb.Renderer(scene, gbuffer=True, denoiser=True)
    r.render(1, 2); r.sync(); r.resolve(2); r.visualize_lbvh(); r.free()
    r = ptb.Renderer(scene, env_nee=True, rank=1, world=3, stripe_rows=8)
    r.render(1, 1); r.sync(); r.free()
    scene.free()
    print("ok", name, flush=True)
