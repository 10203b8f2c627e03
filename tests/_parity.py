"""Parity harness: drive the UNMODIFIED reference (oracle/_ref/refdump, built from /root/reference by
oracle/Makefile and shipped prebuilt to the GPU box) and this library on the same scene files, and
compare LBVH arrays, per-bounce queues / path state and the accumulators.

Test infrastructure only: nothing in the product imports this.
"""
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDUMP = os.path.join(ROOT, "oracle", "_ref", "refdump")


def have_refdump():
    return os.path.exists(REFDUMP) and os.access(REFDUMP, os.X_OK)


SHIMDUMP = os.path.join(ROOT, "oracle", "_ref", "shimdump")


def word_sum(a):
    """refdump --sums / --load-only checksum of an array: position-weighted sum of its 32-bit words (wrapping uint64)."""
    w = np.ascontiguousarray(a).view(np.uint32).reshape(-1).astype(np.uint64)
    n = np.arange(w.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = ((w + np.uint64(1)) * (n * np.uint64(0x9E3779B97F4A7C15) + np.uint64(0x632BE59BD9B4E019))).sum(dtype=np.uint64)
    return "%016x" % int(h)


def run_shimdump(scene_json, out_dir, iters, mode="file", pbo=True, timeout=600):
    """The reference's headless loop over the SHIM (shim/pathtrace_wavefront_ptb200.cpp compiled against the reference's
    unchanged headers, oracle/Makefile): returns (direct, indirect, pbo or None)."""
    os.makedirs(out_dir, exist_ok=True)
    env = dict(os.environ, PTB_SHIM_SCENE=mode)
    p = subprocess.run([SHIMDUMP, scene_json, out_dir, str(iters)] + (["--pbo"] if pbo else []), capture_output=True, text=True, timeout=timeout, env=env)
    if p.returncode != 0 or "SHIMDUMP" not in p.stdout:
        raise RuntimeError(f"shimdump failed ({p.returncode}):\n{p.stdout[-2000:]}\n{p.stderr[-2000:]}")
    d = np.fromfile(os.path.join(out_dir, "direct.bin"), dtype="<f4").reshape(-1, 3)
    i = np.fromfile(os.path.join(out_dir, "indirect.bin"), dtype="<f4").reshape(-1, 3)
    px = np.fromfile(os.path.join(out_dir, "pbo.bin"), dtype="u1").reshape(-1, 4) if pbo else None
    return d, i, px


def run_refdump(scene_json, out_dir, iters=1, stages=True, selfcheck=False, timeout=900, views=False, denoise=False, mega=False, extra=()):
    os.makedirs(out_dir, exist_ok=True)
    cmd = [REFDUMP, scene_json, out_dir, "--iters", str(iters)] + list(extra)
    if stages:
        cmd.append("--stages")
    if selfcheck:
        cmd.append("--selfcheck")
    if views:
        cmd.append("--views")
    if denoise:
        cmd.append("--denoise")
    if mega:   # the reference's megakernel front-end: dumps mega_image / mega_samples only
        cmd = [REFDUMP, scene_json, out_dir, "--iters", str(iters), "--mega"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError(f"refdump failed ({p.returncode}):\n{p.stdout[-2000:]}\n{p.stderr[-2000:]}")
    return p.stdout


class RefDump:
    def __init__(self, d):
        self.d = d
        self.meta = json.load(open(os.path.join(d, "meta.json")))

    def has(self, name):
        return os.path.exists(os.path.join(self.d, name + ".bin"))

    def load(self, name, dtype, width=None):
        a = np.fromfile(os.path.join(self.d, name + ".bin"), dtype=dtype)
        return a.reshape(-1, width) if width else a


def bits_equal(a, b):
    """Bit-level equality of two float arrays (NaN == NaN with the same payload, -0 != +0)."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    if a.shape != b.shape:
        return False
    return bool(np.array_equal(a.view(np.uint8), b.view(np.uint8)))


def count_bit_mismatch(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
    return int((a.view(np.uint32) != b.view(np.uint32)).sum())


def compare_scene(ref, scene):
    """Loader parity: device-upload arrays of the reference vs ptb_scene_get_array."""
    out = {}
    for name, rname, dt, w in [("pos", "mesh_pos", "<f4", 4), ("nor", "mesh_nor", "<f4", 4), ("tangent", "mesh_tangent", "<f4", 4),
                               ("uv", "mesh_uv", "<f4", 2), ("indices_matid", "mesh_indices_matid", "<i4", 4),
                               ("nor_geom", "mesh_nor_geom", "<f4", 4), ("light_tri_idx", "light_tri_idx", "<i4", None),
                               ("light_cdf", "light_cdf", "<f4", None)]:
        a = ref.load(rname, dt, w)
        b = scene.array(name)
        out[name] = bits_equal(a, b)
    m_ref = np.fromfile(os.path.join(ref.d, "materials.bin"), dtype=np.uint8).reshape(-1, 48)[:, :44]   # _pad is uninitialised in the reference
    m_new = scene.array("materials").view(np.uint8).reshape(-1, 48)[:, :44]
    out["materials"] = bool(np.array_equal(m_ref, m_new))
    if ref.has("env_probs"):
        out["env_probs"] = bits_equal(ref.load("env_probs", "<f4"), scene.array("env_probs"))
        out["env_aliases"] = bool(np.array_equal(ref.load("env_aliases", "<i4"), scene.array("env_aliases")))
    return out


LBVH_ARRAYS = [("aabb_min", "<f4", 4), ("aabb_max", "<f4", 4), ("centroid", "<f4", 4), ("primitive_indices", "<i4", None),
               ("morton_codes", "<u8", None), ("child_nodes", "<i4", 2), ("parent", "<i4", None), ("escape_indices", "<i4", None)]


def compare_lbvh(ref, renderer):
    """Bit-equality of every LBVH array.  Index N-1 of the 2N-sized arrays is the unused gap
    (src/scene_structs.h:152): aabb/centroid keep per-primitive values there, compared too."""
    out = {}
    for name, dt, w in LBVH_ARRAYS:
        a = ref.load("bvh_" + name, dt, w)
        b = renderer.lbvh(name)
        out[name] = bits_equal(a, b)
    w_ref = ref.load("bvh_world_aabb", "<f4", 4)
    out["world_aabb"] = bits_equal(w_ref, renderer.lbvh("world_aabb"))
    return out


PATH_STATE = [("ray_ori", "ray_ori", "<f4", 4), ("ray_dir_dist", "ray_dir_dist", "<f4", 4), ("hit_uv", "hit_normal", "<f4", 4),
              ("throughput_pdf", "throughput_pdf", "<f4", 4), ("hit_geom_id", "hit_geom_id", "<i4", None),
              ("material_id", "material_id", "<i4", None), ("remaining_bounces", "remaining_bounces", "<i4", None),
              ("rng_state", "rng_state", "<u4", None)]


def _state_diff(ref, tag, renderer, live_mask=None, fields=None):
    """Number of path slots whose field differs bitwise, per field (restricted to live_mask)."""
    out = {}
    for name, rname, dt, w in PATH_STATE:
        if fields and name not in fields:
            continue
        a = ref.load(f"{tag}_{rname}", dt, w)
        b = renderer.path_field(name)
        if w:
            neq = (a.view(np.uint32) != b.view(np.uint32)).any(axis=1)
        else:
            neq = a.view(np.uint32) != b.view(np.uint32)
        if live_mask is not None:
            neq = neq & live_mask
        out[name] = int(neq.sum())
    return out


def compare_staged_frame(ref, renderer, iteration, depth):
    """Re-drive one frame stage by stage and compare against the reference's dumps after each stage.
    Returns a report dict; report['ok'] is True when every integer item is equal and every float
    item is bit-equal."""
    from cuda_path_tracer_b200 import capi
    rep = {"steps": [], "ok": True}
    P = renderer.width * renderer.height
    counters = ref.load("counters", "<i4", 6)
    renderer.stage(capi.STAGE_RAYGEN, iteration)
    # RayGen does not write hit_normal (stale memory on both sides, src/kernels/ray_gen.cu:65-75)
    d = _state_diff(ref, "raygen", renderer, None, ("ray_ori", "ray_dir_dist", "throughput_pdf", "hit_geom_id", "material_id", "remaining_bounces", "rng_state"))
    rep["raygen"] = d
    if any(d.values()):
        rep["ok"] = False
    for step in range(depth):
        s = {"step": step}
        T = f"s{step}"
        n_ext, n_sh, n_pbr, n_dif, n_rfl, n_rfr = (int(x) for x in counters[step])
        # queues before tracing
        q = np.sort(renderer.queue("extension"))
        s["ext_count"] = (renderer.queue_count("extension"), n_ext)
        s["ext_set_equal"] = bool(np.array_equal(q, np.sort(ref.load(T + "_ext_queue", "<i4"))))
        s["shadow_count"] = (renderer.queue_count("shadow"), n_sh)
        if n_sh > 0 and s["shadow_count"][0] == n_sh:
            o, dd, r, p = renderer.shadow_queue()
            ro, rd, rr, rp = (ref.load(T + "_shadow_ori_tmax", "<f4", 4), ref.load(T + "_shadow_dir", "<f4", 4),
                              ref.load(T + "_shadow_radiance", "<f4", 4), ref.load(T + "_shadow_pixel", "<i4"))
            a, b = np.argsort(p, kind="stable"), np.argsort(rp, kind="stable")
            s["shadow_pixels_equal"] = bool(np.array_equal(p[a], rp[b]))
            if s["shadow_pixels_equal"]:
                s["shadow_ori_mismatch"] = int((o[a].view(np.uint32) != ro[b].view(np.uint32)).any(axis=1).sum())
                s["shadow_dir_mismatch"] = int((dd[a].view(np.uint32) != rd[b].view(np.uint32)).any(axis=1).sum())
                s["shadow_rad_mismatch"] = int((r[a].view(np.uint32) != rr[b].view(np.uint32)).any(axis=1).sum())
        live = np.zeros(P, bool)
        live[ref.load(T + "_ext_queue", "<i4")] = True
        renderer.stage(capi.STAGE_TRACE)
        s["traced"] = _state_diff(ref, T + "_traced", renderer, live, ("ray_dir_dist", "hit_uv", "hit_geom_id", "material_id"))
        # hit_uv / material_id are only written on a hit
        hit = ref.load(T + "_traced_hit_geom_id", "<i4") >= 0
        s["traced_hit"] = _state_diff(ref, T + "_traced", renderer, live & hit, ("hit_uv", "material_id"))
        s["traced"]["hit_uv"] = s["traced_hit"]["hit_uv"]
        s["traced"]["material_id"] = s["traced_hit"]["material_id"]
        del s["traced_hit"]
        direct_ref = ref.load(T + "_direct", "<f4", 3)
        dnew, _ = renderer.accumulators()
        s["direct_mismatch"] = count_bit_mismatch(direct_ref, dnew)
        renderer.stage(capi.STAGE_LOGIC)
        for qn, rq, n in (("pbr", "_pbr_queue", n_pbr), ("diffuse", "_diffuse_queue", n_dif), ("reflection", "_reflection_queue", n_rfl),
                          ("refraction", "_refraction_queue", n_rfr)):
            mine = np.sort(renderer.queue(qn))
            theirs = np.sort(ref.load(T + rq, "<i4"))
            s[qn + "_count"] = (int(mine.size), n)
            s[qn + "_set_equal"] = bool(np.array_equal(mine, theirs))
        _, inew = renderer.accumulators()
        s["indirect_mismatch"] = count_bit_mismatch(ref.load(T + "_indirect", "<f4", 3), inew)
        rb = renderer.path_field("remaining_bounces")
        s["logic_remaining_mismatch"] = int((rb != ref.load(T + "_logic_remaining_bounces", "<i4")).sum())
        renderer.stage(capi.STAGE_SHADE)
        shaded = np.zeros(P, bool)
        for rq in ("_pbr_queue", "_diffuse_queue", "_reflection_queue", "_refraction_queue"):
            shaded[ref.load(T + rq, "<i4")] = True
        alive = ref.load(T + "_shaded_remaining_bounces", "<i4") >= 0
        s["shaded"] = _state_diff(ref, T + "_shaded", renderer, shaded, ("remaining_bounces", "rng_state"))
        s["shaded"].update(_state_diff(ref, T + "_shaded", renderer, shaded & alive, ("ray_ori", "ray_dir_dist", "throughput_pdf")))
        rep["steps"].append(s)
        bad = (s["ext_count"][0] != s["ext_count"][1] or not s["ext_set_equal"] or s["shadow_count"][0] != s["shadow_count"][1]
               or any(s["traced"].values()) or s["direct_mismatch"] or s["indirect_mismatch"] or s["logic_remaining_mismatch"]
               or any(s["shaded"].values()) or not all(s[k + "_set_equal"] for k in ("pbr", "diffuse", "reflection", "refraction"))
               or s.get("shadow_ori_mismatch", 0) or s.get("shadow_dir_mismatch", 0) or s.get("shadow_rad_mismatch", 0)
               or not s.get("shadow_pixels_equal", True))
        if bad:
            rep["ok"] = False
    return rep


def radiance_report(ref_direct, ref_indirect, new_direct, new_indirect, iters):
    """Per-pixel relative error of radiance = (direct + indirect) / iter (src/pathtrace_wavefront.cu:74-75)."""
    a = (ref_direct.astype(np.float64) + ref_indirect) / iters
    b = (new_direct.astype(np.float64) + new_indirect) / iters
    finite = np.isfinite(a).all(axis=1) & np.isfinite(b).all(axis=1)
    nan_same = (np.isnan(a) == np.isnan(b)).all()
    err = np.abs(a - b)[finite]
    den = np.maximum(np.abs(a)[finite], 1e-3)
    rel = (err / den).max(axis=1) if err.size else np.zeros(0)
    rmse = float(np.sqrt(np.mean(err ** 2))) if err.size else 0.0
    return {"pixels": int(a.shape[0]), "non_finite_ref": int((~np.isfinite(a).all(axis=1)).sum()), "nan_pattern_equal": bool(nan_same),
            "bit_mismatch_direct": count_bit_mismatch(ref_direct, new_direct), "bit_mismatch_indirect": count_bit_mismatch(ref_indirect, new_indirect),
            "max_rel_err": float(rel.max()) if rel.size else 0.0, "outliers_gt_1e-4": int((rel > 1e-4).sum()),
            "outlier_fraction": float((rel > 1e-4).mean()) if rel.size else 0.0, "rmse": rmse}


def failures(rep):
    """The non-matching items of a compare_staged_frame report (for assertion messages)."""
    out = {"raygen": {k: v for k, v in rep["raygen"].items() if v}}
    for s in rep["steps"]:
        bad = {}
        for k, v in s.items():
            if isinstance(v, dict):
                if any(v.values()):
                    bad[k] = {a: b for a, b in v.items() if b}
            elif k.endswith("mismatch") and v:
                bad[k] = v
            elif k.endswith("equal") and not v:
                bad[k] = v
            elif k.endswith("count") and v[0] != v[1]:
                bad[k] = v
        if bad:
            out[s["step"]] = bad
    return out
