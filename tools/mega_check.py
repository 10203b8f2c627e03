#!/usr/bin/env python
"""Exploratory: ptb_config.megakernel against the reference's megakernel image (refdump --mega), and timing of both.
usage: mega_check.py [cornell|fire_small|sponza_small|fire_full|sponza_full ...] [--iters N] [--bench]"""
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

GEN = {
    "cornell": lambda d: scenegen.cornell(d, res=(200, 160)),
    "cornell800": lambda d: scenegen.cornell(d, res=(800, 800)),
    "fire_small": lambda d: scenegen.fireplace(d, res=(480, 270), scale=0.25)[0],
    "sponza_small": lambda d: scenegen.sponza(d, res=(320, 180), scale=0.2, tex_size=128)[0],
    "fire_full": lambda d: scenegen.fireplace(d, res=(1920, 1080))[0],
    "sponza_full": lambda d: scenegen.sponza(d, res=(1920, 1080))[0],
}


def main():
    iters = int(sys.argv[sys.argv.index("--iters") + 1]) if "--iters" in sys.argv else 2
    args = [a for a in sys.argv[1:] if a in GEN]
    bench = "--bench" in sys.argv
    report = {}
    for name in args or ["cornell", "fire_small", "sponza_small"]:
        d = f"/tmp/mega_{name}"
        scene_json = GEN[name](d)
        par.run_refdump(scene_json, d + "/ref", iters=iters, mega=True)
        ref = np.fromfile(d + "/ref/mega_image.bin", dtype="<f4").reshape(-1, 3)
        smp = np.fromfile(d + "/ref/mega_samples.bin", dtype="<i4")
        scene = ptb.Scene(scene_json)
        r = ptb.Renderer(scene, megakernel=True)
        r.render(1, iters)
        r.sync()
        direct, indirect = r.accumulators()
        bad = (direct.view(np.uint32) != ref.view(np.uint32)).any(axis=1)
        rel = np.abs(direct - ref) / np.maximum(np.abs(ref), 1e-6)
        rep = {"pixels": int(ref.shape[0]), "iters": iters, "samples_ok": bool((smp == iters).all()), "mismatching_pixels": int(bad.sum()),
               "max_rel": float(rel.max()), "indirect_zero": bool((indirect == 0).all()), "mean_ref": float(ref.mean()), "mean_new": float(direct.mean())}
        if bad.any():
            idx = np.nonzero(bad)[0][:8]
            rep["examples"] = [{"pixel": int(i), "ref": ref[i].tolist(), "new": direct[i].tolist()} for i in idx]
        if bench:
            s = r.stats()
            for mode in ("mega", "wave"):
                rr = ptb.Renderer(scene, megakernel=(mode == "mega"))
                rr.render(1, 3)
                rr.sync()
                t0 = time.time()
                rr.render(4, 20)
                rr.sync()
                dt = (time.time() - t0) / 20
                st = rr.stats()
                rep[f"new_{mode}_ms"] = dt * 1e3
                rep[f"new_{mode}_mpaths"] = ref.shape[0] / dt / 1e6
                rep[f"new_{mode}_rays_per_frame"] = (st.extension_rays + st.shadow_rays) / st.frames
                rr.free()
            for mode in ("mega", "wave"):
                cmd = [par.REFDUMP, scene_json, d + "/bench", "--bench", "20", "3", "--no-dump"] + (["--mega"] if mode == "mega" else [])
                out = subprocess.run(cmd, capture_output=True, text=True).stdout
                for line in out.splitlines():
                    if line.startswith("REFBENCH"):
                        rep[f"ref_{mode}"] = json.loads(line[len("REFBENCH "):])
        r.free()
        report[name] = rep
        print(name, json.dumps(rep), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(report, open(os.path.join(ROOT, "gpurun_out", "mega_check.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
