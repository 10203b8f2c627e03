#!/usr/bin/env python
"""Where does an extension-ray walk differ from the reference-order walk?  Staged frame of the textured atrium with two knob sets;
after every TRACE the hit records are compared and the first differing rays are printed."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

A = dict(kv.split("=") for kv in (sys.argv[1] if len(sys.argv) > 1 else "PTB_RC_MODE=1").split(","))
B = dict(kv.split("=") for kv in (sys.argv[2] if len(sys.argv) > 2 else "PTB_RC_MODE=3").split(","))
scene = ptb.Scene(scenegen.sponza("/tmp/wide_diff_sponza", res=(1920, 1080))[0])


def make(env):
    for k in list(os.environ):
        if k.startswith("PTB_"):
            del os.environ[k]
    os.environ.update(env)
    return ptb.Renderer(scene)


ra, rb = make(A), make(B)
mats = scene.array("materials")
for it in (1, 2):
    ra.stage(capi.STAGE_RAYGEN, it); rb.stage(capi.STAGE_RAYGEN, it)
    for b in range(8):
        ra.stage(capi.STAGE_TRACE); rb.stage(capi.STAGE_TRACE)
        q = ra.queue("extension")
        ha, hb = ra.path_field("hit_geom_id"), rb.path_field("hit_geom_id")
        da, db = ra.path_field("ray_dir_dist"), rb.path_field("ray_dir_dist")
        ua, ub = ra.path_field("hit_uv"), rb.path_field("hit_uv")
        bad = q[(ha[q] != hb[q]) | (da[q, 3].view(np.uint32) != db[q, 3].view(np.uint32))]
        print(f"iter {it} bounce {b}: {q.size} rays, {bad.size} differ", flush=True)
        for p in bad[:6]:
            ma, mb = ra.path_field("material_id")[p], rb.path_field("material_id")[p]
            o = ra.path_field("ray_ori")[p]
            print(f"   path {p}: A prim {ha[p]} t {da[p,3]!r} uv {ua[p,:2]} mat {ma} emit {mats[ma]['emittance'] if ma >= 0 else None} | "
                  f"B prim {hb[p]} t {db[p,3]!r} uv {ub[p,:2]} mat {mb} | ori {o[:3]} dir {da[p,:3]}")
        for r in (ra, rb):
            r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
