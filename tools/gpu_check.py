#!/usr/bin/env python
"""Exploratory parity + timing run on the GPU box (not a test; see tests/ for the gates).
Writes gpurun_out/gpu_check.json."""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _parity as par  # noqa: E402

import cuda_path_tracer_b200 as ptb  # noqa: E402
from cuda_path_tracer_b200 import scenegen  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
report = {}


def staged_case(name, scene_json, iters=1):
    d = f"/tmp/ref_{name}"
    t0 = time.time()
    log = par.run_refdump(scene_json, d, iters=iters, stages=True, selfcheck=True)
    ref = par.RefDump(d)
    scene = ptb.Scene(scene_json)
    r = ptb.Renderer(scene)
    rep = {"refdump_s": time.time() - t0, "selfcheck": [l for l in log.splitlines() if l.startswith("REFSELFCHECK")],
           "scene": par.compare_scene(ref, scene), "lbvh": par.compare_lbvh(ref, r), "meta": ref.meta}
    depth = ref.meta["trace_depth"]
    for it in range(1, iters):
        r.render(it)
    r.sync()
    st = par.compare_staged_frame(ref, r, iters, depth)
    rep["staged_ok"] = st["ok"]
    rep["raygen"] = st["raygen"]
    rep["steps"] = st["steps"]
    d_new, i_new = r.accumulators()
    # the staged frame skipped nothing: accumulators now hold `iters` samples
    rep["radiance_staged"] = par.radiance_report(ref.load("direct", "<f4", 3), ref.load("indirect", "<f4", 3), d_new, i_new, iters)
    # and the graph path
    r.render(1, iters)
    r.sync()
    d_new, i_new = r.accumulators()
    rep["radiance_graph"] = par.radiance_report(ref.load("direct", "<f4", 3), ref.load("indirect", "<f4", 3), d_new, i_new, iters)
    s = r.stats()
    rep["stats"] = dict(ext=s.extension_rays, shadow=s.shadow_rays, frames=s.frames, launches=s.kernel_launches,
                        lbvh_ms=s.lbvh_build_ms, depth=s.lbvh_max_depth)
    r.free()
    report[name] = rep
    print(name, "staged_ok", rep["staged_ok"], "lbvh", all(rep["lbvh"].values()), "scene", all(rep["scene"].values()),
          rep["radiance_graph"], flush=True)


def timing_case(name, scene_json, frames=20, warm=3):
    scene = ptb.Scene(scene_json)
    r = ptb.Renderer(scene)
    r.render(1, warm)
    r.sync()
    r.stats_reset()
    t0 = time.time()
    r.render(warm + 1, frames)
    r.sync()
    dt = time.time() - t0
    s = r.stats()
    P = r.width * r.height
    rep = {"ms_per_frame": dt * 1e3 / frames, "mpaths_per_s": frames * P / dt / 1e6,
           "mrays_per_s": (s.extension_rays + s.shadow_rays) / dt / 1e6, "lbvh_ms": s.lbvh_build_ms, "depth": s.lbvh_max_depth}
    r.stage(0, 1)
    ms, rays = r.time_raycast(20)
    rep["raycast_bounce0_ms"] = ms
    rep["raycast_bounce0_mrays"] = rays / ms / 1e3
    r.free()
    p = subprocess.run([par.REFDUMP, scene_json, f"/tmp/refb_{name}", "--no-dump", "--bench", str(frames), str(warm)],
                       capture_output=True, text=True, timeout=1200)
    rep["ref_wave"] = [l for l in p.stdout.splitlines() if l.startswith("REFBENCH")]
    p = subprocess.run([par.REFDUMP, scene_json, f"/tmp/refb_{name}", "--no-dump", "--mega", "--bench", str(frames), str(warm)],
                       capture_output=True, text=True, timeout=1200)
    rep["ref_mega"] = [l for l in p.stdout.splitlines() if l.startswith("REFBENCH")]
    report["timing_" + name] = rep
    print("timing", name, rep, flush=True)


def debug_shade(scene_json, name, step_limit=1):
    """Break step-0 shading mismatches down by material queue and print examples."""
    from cuda_path_tracer_b200 import capi
    d = f"/tmp/ref_{name}"
    if not os.path.exists(os.path.join(d, "meta.json")):
        par.run_refdump(scene_json, d, iters=1, stages=True)
    ref = par.RefDump(d)
    scene = ptb.Scene(scene_json)
    r = ptb.Renderer(scene)
    r.stage(capi.STAGE_RAYGEN, 1)
    for step in range(step_limit):
        r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    T = f"s{step_limit - 1}"
    mats = scene.array("materials")
    matid = ref.load(T + "_shaded_material_id", "<i4")
    for qn in ("pbr", "diffuse", "reflection", "refraction"):
        q = ref.load(f"{T}_{qn}_queue", "<i4")
        alive = ref.load(T + "_shaded_remaining_bounces", "<i4")[q] >= 0
        qa = q[alive]
        line = {"queue": qn, "n": int(q.size), "alive": int(qa.size)}
        for fld, rname in (("ray_dir_dist", "ray_dir_dist"), ("throughput_pdf", "throughput_pdf"), ("ray_ori", "ray_ori")):
            a = ref.load(f"{T}_shaded_{rname}", "<f4", 4)[qa]
            b = r.path_field(fld)[qa]
            neq = (a.view(np.uint32) != b.view(np.uint32))
            line[fld] = [int(x) for x in neq.sum(axis=0)]
            rows = np.nonzero(neq.any(axis=1))[0][:3]
            for k in rows:
                ulp = (a[k].view(np.int32).astype(np.int64) - b[k].view(np.int32).astype(np.int64)).tolist()
                print("   ", qn, fld, "path", int(qa[k]), "mat", int(matid[qa[k]]), "ref", a[k].tolist(), "new", b[k].tolist(), "ulp", ulp)
        print(line, flush=True)
    r.free()


if __name__ == "__main__":
    which = sys.argv[1:] or ["cornell", "fire_small", "fire", "timing"]
    try:
        if "cornell" in which:
            staged_case("cornell256", scenegen.cornell("/tmp/sc_cornell256", res=(256, 256)))
        if "fire_small" in which:
            staged_case("fire_small", scenegen.fireplace("/tmp/sc_fire_small", res=(480, 270), scale=0.25)[0])
        if "fire" in which:
            staged_case("fire", scenegen.fireplace("/tmp/sc_fire", res=(1920, 1080))[0])
        if "debug" in which:
            debug_shade(scenegen.fireplace("/tmp/sc_fire_small", res=(480, 270), scale=0.25)[0], "fire_small")
        if "timing" in which:
            timing_case("cornell800", scenegen.cornell("/tmp/sc_cornell800", res=(800, 800)))
            timing_case("fire", scenegen.fireplace("/tmp/sc_fire", res=(1920, 1080))[0])
    finally:
        def default(o):
            if isinstance(o, (np.integer,)):
                return int(o)
            if isinstance(o, (np.floating,)):
                return float(o)
            return str(o)
        json.dump(report, open(os.path.join(OUT, "gpu_check.json"), "w"), indent=1, default=default)

