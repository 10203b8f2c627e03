"""Per-config measurement on the GPU box (not the bench line): Mpaths/s, Mrays/s, LBVH build, per-bounce RayCast,
and -- where the scene is a scene.json and oracle/_ref/refdump travelled -- the reference's wavefront and megakernel
builds on the same files.  Usage: python tools/measure_configs.py cornell fireplace sponza stress [--tris N]"""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

REFDUMP = os.path.join(ROOT, "oracle", "_ref", "refdump")


def ref_bench(scene_json, mega=False, steps=10, warm=3):
    if not os.path.exists(REFDUMP):
        return None
    p = subprocess.run([REFDUMP, scene_json, "/tmp/ptb_mc_ref", "--no-dump", "--bench", str(steps), str(warm)] + (["--mega"] if mega else []),
                       capture_output=True, text=True, timeout=3000)
    line = [l for l in p.stdout.splitlines() if l.startswith("REFBENCH")]
    return json.loads(line[0][9:]) if line else {"error": (p.stderr or p.stdout)[-200:]}


def measure(name, scene, scene_json=None, steps=10, spp_note=""):
    t0 = time.perf_counter()
    r = ptb.Renderer(scene)
    init_s = time.perf_counter() - t0
    P = r.width * r.height
    r.render(1, 3); r.sync()
    r.stats_reset()
    t0 = time.perf_counter(); r.render(1, steps); r.sync(); dt = time.perf_counter() - t0
    st = r.stats()
    rays = st.extension_rays + st.shadow_rays
    out = {"config": name, "triangles": st.num_triangles, "resolution": [r.width, r.height], "lbvh_build_ms": round(st.lbvh_build_ms, 3),
           "lbvh_mtris_per_s": round(st.num_triangles / st.lbvh_build_ms / 1e3, 1), "lbvh_max_depth": st.lbvh_max_depth, "init_s": round(init_s, 2),
           "ms_per_frame": round(dt / steps * 1e3, 3), "mpaths_per_s": round(P * steps / dt / 1e6, 2), "mrays_per_s": round(rays / dt / 1e6, 1),
           "rays_per_path": round(rays / (P * steps), 2)}
    # per-bounce extension RayCast alone
    depth = scene.info.trace_depth
    r.stage(capi.STAGE_RAYGEN, 1)
    per = []
    for b in range(depth):
        ms, n = r.time_raycast(3)
        r.stage(capi.STAGE_TRACE)
        q = r.queue("extension")
        hits = int((r.path_field("hit_geom_id")[q] >= 0).sum())
        per.append({"rays": n, "hits": hits, "ms": round(ms, 4), "alg_GBps": round((36.0 * n + 28.0 * hits + 8.0 * (n - hits)) / (ms * 1e-3) / 1e9, 1) if ms > 0 else None})
        r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    out["raycast_ext_per_bounce"] = per
    tot_ms = sum(p["ms"] for p in per); tot_rays = sum(p["rays"] for p in per)
    out["raycast_ext_mrays_per_s"] = round(tot_rays / (tot_ms * 1e-3) / 1e6, 1)
    r.free()
    if scene_json:
        for mega in (False, True):
            rb = ref_bench(scene_json, mega, steps)
            if rb:
                out["reference_megakernel" if mega else "reference_wavefront"] = rb
        if "reference_wavefront" in out and "mpaths_per_s" in out["reference_wavefront"]:
            out["speedup_vs_reference_wavefront"] = round(out["mpaths_per_s"] / out["reference_wavefront"]["mpaths_per_s"], 2)
    print("MEASURE " + json.dumps(out), flush=True)
    return out


if __name__ == "__main__":
    which = [a for a in sys.argv[1:] if not a.startswith("--")] or ["cornell", "fireplace", "sponza"]
    tris = int(sys.argv[sys.argv.index("--tris") + 1]) if "--tris" in sys.argv else 10_000_000
    for w in which:
        if w == "cornell":
            p = scenegen.cornell("/tmp/mc_cornell", res=(800, 800)); measure("1: cornell 800x800", ptb.Scene(p), p)
        elif w == "fireplace":
            p = scenegen.fireplace("/tmp/mc_fire")[0]; measure("2: fireplace-like 143k 1080p", ptb.Scene(p), p)
        elif w == "sponza":
            p = scenegen.sponza("/tmp/mc_sponza")[0]; measure("3: sponza-like 262k textured + HDR env 1080p", ptb.Scene(p), p)
        elif w == "stress":
            t0 = time.perf_counter()
            pos, nor, idx, mat = scenegen.heightfield_arrays(tris)
            s = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera())
            print(f"generated {len(idx)} triangles in {time.perf_counter() - t0:.1f} s", flush=True)
            measure(f"4: stress {len(idx)} tris 1080p", s, None, steps=4)
