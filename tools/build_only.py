#!/usr/bin/env python
"""LBVH build of the 10 M-triangle height field, twice (profiling target for the build kernels)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import scenegen  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
pos, nor, idx, mat = scenegen.heightfield_arrays(n)
s = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera((64, 36)))
for _ in range(2):
    r = ptb.Renderer(s)
    st = r.stats()
    print("triangles", st.num_triangles, "build ms", st.lbvh_build_ms, "depth", st.lbvh_max_depth)
    r.free()
