#!/usr/bin/env python
"""Which paths make two knob sets differ in the INDIRECT accumulator after one whole frame, and how do those paths end
(environment miss / emissive hit, at which bounce)?  Textured atrium, full size."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

A = dict(kv.split("=") for kv in sys.argv[1].split(","))
B = dict(kv.split("=") for kv in sys.argv[2].split(","))
scene = ptb.Scene(scenegen.sponza("/tmp/wide_diff_sponza", res=(1920, 1080))[0])


def make(env):
    for k in list(os.environ):
        if k.startswith("PTB_"):
            del os.environ[k]
    os.environ.update(env)
    return ptb.Renderer(scene)


ra, rb = make(A), make(B)
ra.render(1); rb.render(1)
ia, ib = ra.accumulators()[1], rb.accumulators()[1]
bad = np.nonzero((ia.view(np.uint32) != ib.view(np.uint32)).any(axis=1))[0]
print("pixels whose indirect differs:", bad.size, "max rel", float((np.abs(ia - ib) / np.maximum(np.abs(ia), 1e-6)).max()))
mats = scene.array("materials")
ends = {}
ra.stage(capi.STAGE_RAYGEN, 1)
alive = np.ones(ia.shape[0], bool)
for b in range(8):
    ra.stage(capi.STAGE_TRACE)
    before = ra.accumulators()[1].copy()
    hit, mat = ra.path_field("hit_geom_id"), ra.path_field("material_id")
    ra.stage(capi.STAGE_LOGIC)
    after = ra.accumulators()[1]
    changed = (before.view(np.uint32) != after.view(np.uint32)).any(axis=1)
    for p in bad:
        if changed[p]:
            kind = "env miss" if hit[p] < 0 else f"emissive (mat {mat[p]}, emittance {mats[mat[p]]['emittance']:.1f})"
            ends.setdefault((b, kind), []).append(p)
    ra.stage(capi.STAGE_SHADE)
for k in sorted(ends):
    print(k, len(ends[k]))
