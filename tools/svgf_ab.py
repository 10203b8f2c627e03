"""A/B of the SVGF kernel variants (env PTB_SVGF_VARIANT): frame time with the denoiser on, and a hash of the
denoised image so that every variant is checked to produce the same bits."""
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) > 1 and sys.argv[1] == "--child":
    import cuda_path_tracer_b200 as ptb
    from cuda_path_tracer_b200 import scenegen
    p = scenegen.sponza("/tmp/msv_sponza")[0]
    scene = ptb.Scene(p)
    r = ptb.Renderer(scene, denoiser=True)
    r.render(1, 8); r.sync()
    best = 1e9
    for rep in range(3):
        t0 = time.perf_counter(); r.render(9 + 30 * rep, 30); r.sync()
        best = min(best, (time.perf_counter() - t0) / 30 * 1e3)
    from cuda_path_tracer_b200 import capi
    img = r.resolve(98, capi.IMG_DENOISED_RGB32F)
    print("AB " + json.dumps({"variant": os.environ.get("PTB_SVGF_VARIANT", "0"), "ms_per_frame": round(best, 4),
                              "sha": hashlib.sha1(img.tobytes()).hexdigest()[:12]}), flush=True)
else:
    for v in sys.argv[1:]:
        env = dict(os.environ, PTB_SVGF_VARIANT=v)
        q = subprocess.run([sys.executable, __file__, "--child"], env=env, capture_output=True, text=True)
        lines = [l for l in q.stdout.splitlines() if l.startswith("AB ")]
        print(lines[0] if lines else ("FAILED " + v + " " + q.stderr[-400:]), flush=True)
