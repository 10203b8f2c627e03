#!/usr/bin/env python
"""A/B of the 4-wide quantised traversal (PTB_WIDE bits: 1 shadow, 2 extension, 4 megakernel): ms/frame and whether the
accumulators stay bit-identical to the binary walk.  usage: wide_ab.py [setting ...]   e.g. wide_ab.py 0 1 2 3"""
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SCENES = ("fire_full", "sponza_full", "cornell800")


def child(mega):
    import numpy as np
    import cuda_path_tracer_b200 as ptb
    from cuda_path_tracer_b200 import scenegen
    gens = {"fire_full": lambda d: scenegen.fireplace(d, res=(1920, 1080))[0], "sponza_full": lambda d: scenegen.sponza(d, res=(1920, 1080))[0],
            "cornell800": lambda d: scenegen.cornell(d, res=(800, 800))}
    out = {}
    for name in SCENES:
        d = f"/tmp/wideab_{name}"
        path = os.path.join(d, "scene.json")
        scene_json = path if os.path.exists(path) else gens[name](d)
        r = ptb.Renderer(ptb.Scene(scene_json), megakernel=mega)
        r.render(1, 3); r.sync()
        a, b = r.accumulators()
        digest = hashlib.md5(a.tobytes() + b.tobytes()).hexdigest()
        np.save(f"/tmp/wideab_{name}_{os.environ.get('AB_TAG', '0')}_{int(mega)}.npy", np.concatenate([a, b], axis=1))
        t0 = time.time()
        r.render(4, 20); r.sync()
        out[name] = {"ms": round((time.time() - t0) / 20 * 1e3, 3), "md5": digest}
        r.free()
    print("AB " + json.dumps(out))


if __name__ == "__main__":
    if "--child" in sys.argv:
        child("--mega" in sys.argv)
    else:
        import numpy as np
        mega = "--mega" in sys.argv
        settings = [a for a in sys.argv[1:] if a.isdigit()] or ["0", "1", "2", "3"]
        extra = [a for a in sys.argv[1:] if "=" in a]
        res = {}
        for s in settings:
            env = dict(os.environ, PTB_WIDE=s, **dict(e.split("=") for e in extra))
            env["AB_TAG"] = s
            o = subprocess.run([sys.executable, __file__, "--child"] + (["--mega"] if mega else []), env=env, capture_output=True, text=True)
            line = [l for l in o.stdout.splitlines() if l.startswith("AB ")]
            res[s] = json.loads(line[0][3:]) if line else o.stderr[-800:]
            print(s, res[s], flush=True)
        base = settings[0]
        for s in settings[1:]:
            for name in SCENES:
                try:
                    a = np.load(f"/tmp/wideab_{name}_{base}_{int(mega)}.npy"); b = np.load(f"/tmp/wideab_{name}_{s}_{int(mega)}.npy")
                    bad = int((a.view(np.uint32) != b.view(np.uint32)).any(axis=1).sum())
                    print(f"PTB_WIDE={s} vs {base} {name}: {bad} mismatching pixels of {a.shape[0]}", flush=True)
                except Exception as e:  # noqa: BLE001
                    print("compare failed", s, name, e)
