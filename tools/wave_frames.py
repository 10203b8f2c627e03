#!/usr/bin/env python
"""A few frames of the 143k / 1080p scene through the wavefront path (profiling target for ncu)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import scenegen  # noqa: E402

r = ptb.Renderer(ptb.Scene(scenegen.fireplace("/tmp/ptb_wave_frames", res=(1920, 1080))[0]))
r.render(1, int(sys.argv[1]) if len(sys.argv) > 1 else 2)
r.sync()
s = r.stats()
print("frames", s.frames, "rays", s.extension_rays + s.shadow_rays)
r.free()
