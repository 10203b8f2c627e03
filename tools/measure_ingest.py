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
