"""Scene ingestion at scale: this library's loader (all host threads, then one thread)
beside the reference's own Scene constructor (oracle/_ref/refdump --load-only), with a checksum over vertex
positions / uvs in id order, index triples and material ids that both sides must agree on; then the binary cache."""
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

REFDUMP = os.path.join(ROOT, "oracle", "_ref", "refdump")


def checksum(scene):
    pos = scene.array("pos")[:, :3].view(np.uint32)
    uv = scene.array("uv").view(np.uint32)
    im = scene.array("indices_matid")
    words = np.concatenate([np.concatenate([pos, uv], axis=1).reshape(-1), im[:, :3].reshape(-1).view(np.uint32), im[:, 3].view(np.uint32)]).astype(np.uint64)
    n = np.arange(words.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        h = ((words + np.uint64(1)) * (n * np.uint64(0x9E3779B97F4A7C15) + np.uint64(0x632BE59BD9B4E019))).sum(dtype=np.uint64)
    return "%016x" % int(h)


def load_with_stage_line(path, device=None):
    """(scene, seconds, the loader's own PTB_LOADER_VERBOSE line about its de-duplication + pack stage)"""
    import tempfile
    os.environ["PTB_LOADER_VERBOSE"] = "1"
    sys.stderr.flush()
    keep = os.dup(2)
    with tempfile.TemporaryFile() as f:
        os.dup2(f.fileno(), 2)
        try:
            t0 = time.perf_counter(); s = ptb.Scene(path, device=device); dt = time.perf_counter() - t0
        finally:
            os.dup2(keep, 2); os.close(keep)
            os.environ.pop("PTB_LOADER_VERBOSE", None)
        f.seek(0)
        lines = [l for l in f.read().decode(errors="replace").splitlines() if "de-duplication" in l]
    return s, dt, (lines[-1].replace("ptb200: ", "") if lines else "")


def main():
    scale = float(sys.argv[1]) if len(sys.argv) > 1 else 3.8
    out_dir = sys.argv[2] if len(sys.argv) > 2 else "/tmp/ingest_scene"
    path, ntri = scenegen.fireplace(out_dir, scale=scale)
    size = sum(os.path.getsize(os.path.join(out_dir, f)) for f in os.listdir(out_dir))
    res = {"triangles": int(ntri), "obj_bytes": size, "host_threads": os.cpu_count()}
    for label, env in (("load_s", None), ("load_1thread_s", "1")):
        if env: os.environ["PTB_LOADER_THREADS"] = env
        else: os.environ.pop("PTB_LOADER_THREADS", None)
        t0 = time.perf_counter(); s = ptb.Scene(path); res[label] = round(time.perf_counter() - t0, 3)
        res.setdefault("checksums", []).append(checksum(s))
        if env is None and hasattr(s, "save_cache"):
            cache = os.path.join(out_dir, "scene.ptbscene")
            t0 = time.perf_counter(); s.save_cache(cache); res["cache_save_s"] = round(time.perf_counter() - t0, 3)
            res["cache_bytes"] = os.path.getsize(cache)
            t0 = time.perf_counter(); c = ptb.Scene.load_cache(cache); res["cache_load_s"] = round(time.perf_counter() - t0, 3)
            res["checksums"].append(checksum(c)); c.free()
        s.free()
    os.environ.pop("PTB_LOADER_THREADS", None)
    s, dt, line = load_with_stage_line(path)
    res["host_stage"] = line; s.free()
    if ptb.capi.lib.ptb_device_count() > 0:
        # de-duplication + repack on GPU 0 (ptb_scene_load_device): the first call pays for the CUDA context, the second does not
        for label in ("load_gpu_first_s", "load_gpu_s"):
            s, dt, line = load_with_stage_line(path, device=0)
            res[label] = round(dt, 3); res["gpu_stage"] = line
            res["checksums"].append(checksum(s)); s.free()
    if os.path.exists(REFDUMP):
        q = subprocess.run([REFDUMP, path, "/tmp/ingest_ref", "--load-only"], capture_output=True, text=True)
        for l in q.stdout.splitlines():
            if l.startswith("REFLOAD"):
                r = json.loads(l[8:]); res["reference_load_s"] = r["load_s"]; res["reference_checksum"] = r["checksum"]
                res["reference_vertices"] = r["vertices"]
    res["checksums_agree"] = len(set(res["checksums"] + ([res["reference_checksum"]] if "reference_checksum" in res else []))) == 1
    print("MEASURE_INGEST " + json.dumps(res))


if __name__ == "__main__":
    main()
