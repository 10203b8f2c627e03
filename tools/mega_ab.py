#!/usr/bin/env python
"""A/B of the megakernel's scheduling knobs (env overrides read by ptb_init): ms/frame per scene and setting.
usage: mega_ab.py  (runs itself once per setting in a subprocess; writes gpurun_out/mega_ab.json)"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SETTINGS = [{"PTB_MEGA_REGEN": r, "PTB_MEGA_REFILL": f, "PTB_MEGA_BLOCKS": b} for b in ("8", "10") for r, f in (("0", "1"), ("1", "1"))]


def child():
    import cuda_path_tracer_b200 as ptb
    from cuda_path_tracer_b200 import scenegen
    out = {}
    for name, gen in (("fire_full", lambda d: scenegen.fireplace(d, res=(1920, 1080))[0]), ("sponza_full", lambda d: scenegen.sponza(d, res=(1920, 1080))[0]),
                      ("cornell800", lambda d: scenegen.cornell(d, res=(800, 800)))):
        d = f"/tmp/megaab_{name}"
        path = os.path.join(d, "scene.json")
        scene_json = path if os.path.exists(path) else gen(d)
        r = ptb.Renderer(ptb.Scene(scene_json), megakernel=True)
        r.render(1, 3); r.sync()
        t0 = time.time()
        r.render(4, 20); r.sync()
        out[name] = (time.time() - t0) / 20 * 1e3
        r.free()
    print("AB " + json.dumps(out))


if __name__ == "__main__":
    if "--child" in sys.argv:
        child()
    else:
        res = []
        for s in SETTINGS:
            env = dict(os.environ, **s)
            o = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True)
            line = [l for l in o.stdout.splitlines() if l.startswith("AB ")]
            res.append({"setting": s, "ms": json.loads(line[0][3:]) if line else o.stderr[-500:]})
            print(res[-1], flush=True)
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        json.dump(res, open(os.path.join(ROOT, "gpurun_out", "mega_ab.json"), "w"), indent=1)
