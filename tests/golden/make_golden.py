#!/usr/bin/env python
"""Generate the golden fixtures in this directory by running the UNMODIFIED reference build
(oracle/_ref/refdump, compiled from /root/reference by oracle/Makefile) on a GPU box:

    gpurun -- python tests/golden/make_golden.py          # writes gpurun_out/golden/*.npz
    cp gpurun_out/golden/*.npz tests/golden/

Each .npz holds, for one small procedural scene (cuda_path_tracer_b200/scenegen.py, deterministic):
the reference's uploaded mesh arrays, its LBVH arrays, the RayGen path state, bounce-0/1 RayCast
results, the bounce-1 shadow queue, per-bounce counters and the final accumulators at 1 spp.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _parity as par  # noqa: E402
from cuda_path_tracer_b200 import scenegen  # noqa: E402

CASES = {
    "cornell64": lambda d: scenegen.cornell(d, res=(64, 64)),
    "fire_tiny": lambda d: scenegen.fireplace(d, res=(96, 54), scale=0.1)[0],
}
KEEP = [("mesh_pos", "<f4", 4), ("mesh_nor", "<f4", 4), ("mesh_uv", "<f4", 2), ("mesh_indices_matid", "<i4", 4), ("mesh_nor_geom", "<f4", 4),
        ("light_tri_idx", "<i4", None), ("light_cdf", "<f4", None),
        ("bvh_aabb_min", "<f4", 4), ("bvh_aabb_max", "<f4", 4), ("bvh_centroid", "<f4", 4), ("bvh_primitive_indices", "<i4", None),
        ("bvh_morton_codes", "<u8", None), ("bvh_child_nodes", "<i4", 2), ("bvh_parent", "<i4", None), ("bvh_escape_indices", "<i4", None),
        ("bvh_world_aabb", "<f4", 4),
        ("raygen_ray_ori", "<f4", 4), ("raygen_ray_dir_dist", "<f4", 4), ("raygen_rng_state", "<u4", None),
        ("s0_traced_ray_dir_dist", "<f4", 4), ("s0_traced_hit_normal", "<f4", 4), ("s0_traced_hit_geom_id", "<i4", None),
        ("s0_traced_material_id", "<i4", None),
        ("s1_ext_queue", "<i4", None), ("s0_shaded_ray_ori", "<f4", 4), ("s0_shaded_ray_dir_dist", "<f4", 4),
        ("s1_traced_ray_dir_dist", "<f4", 4), ("s1_traced_hit_geom_id", "<i4", None),
        ("s1_shadow_ori_tmax", "<f4", 4), ("s1_shadow_dir", "<f4", 4), ("s1_shadow_radiance", "<f4", 4), ("s1_shadow_pixel", "<i4", None),
        ("s0_direct", "<f4", 3), ("s1_direct", "<f4", 3), ("counters", "<i4", 6), ("direct", "<f4", 3), ("indirect", "<f4", 3)]

if __name__ == "__main__":
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    for name, gen in CASES.items():
        scene_json = gen(f"/tmp/golden_{name}")
        d = f"/tmp/golden_ref_{name}"
        log = par.run_refdump(scene_json, d, iters=1, stages=True, selfcheck=True)
        assert '"mismatching_floats": 0' in log, log
        ref = par.RefDump(d)
        arrays = {k: ref.load(k, dt, w) for k, dt, w in KEEP}
        arrays["materials"] = np.fromfile(os.path.join(d, "materials.bin"), dtype=np.uint8)
        arrays["camera"] = np.fromfile(os.path.join(d, "camera.bin"), dtype="<f4")
        arrays["meta"] = np.array([ref.meta[k] for k in ("num_triangles", "num_vertices", "width", "height", "trace_depth", "num_materials", "num_lights")], np.int64)
        np.savez_compressed(os.path.join(out, name + ".npz"), **arrays)
        print(name, "ok", os.path.getsize(os.path.join(out, name + ".npz")), "bytes")
