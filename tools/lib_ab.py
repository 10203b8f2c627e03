#!/usr/bin/env python
"""A/B of whole library builds on the GPU box: for every .so given, a subprocess renders the bench scene (and the textured
atrium) and prints ms per frame and an md5 of the accumulators after 3 iterations.  Usage: python tools/lib_ab.py a.so b.so"""
import hashlib
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, ROOT)
    import cuda_path_tracer_b200 as ptb
    from cuda_path_tracer_b200 import scenegen
    for name, gen in (("fireplace", scenegen.fireplace), ("sponza", scenegen.sponza)):
        scene = ptb.Scene(gen(f"/tmp/lib_ab_{name}", res=(1920, 1080))[0])
        r = ptb.Renderer(scene)
        r.render(1, 3); r.sync()
        d, i = r.accumulators()
        md5 = hashlib.md5(d.tobytes() + i.tobytes()).hexdigest()
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter(); r.render(4, 30); r.sync()
            best = min(best, (time.perf_counter() - t0) / 30 * 1e3)
        print(f"  {name}: {best:.3f} ms/frame  md5 {md5}", flush=True)
        r.free(); scene.free()
else:
    for lib in sys.argv[1:]:
        print(lib, flush=True)
        subprocess.run([sys.executable, __file__, "--child"], env=dict(os.environ, PTB200_LIB=os.path.abspath(lib)), check=True)
