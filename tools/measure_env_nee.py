"""ptb_config.env_nee on the 262k-triangle textured scene with the sun-and-sky environment (1920x1080): frame time
with the flag off / on, the mean radiance of both estimators, and the noise of an 8-spp image of each against a
512-spp image of the flag-on estimator (relative RMS over pixels of the luminance)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

p = scenegen.sponza("/tmp/msv_sponza")[0]
scene = ptb.Scene(p)
out = {}
img = {}
for flag in (False, True):
    r = ptb.Renderer(scene, env_nee=flag)
    r.render(1, 8); r.sync()
    d, i = r.accumulators()
    img[flag] = (d + i) / 8.0
    t0 = time.perf_counter(); r.render(9, 40); r.sync()
    out["ms_per_frame_" + ("on" if flag else "off")] = round((time.perf_counter() - t0) / 40 * 1e3, 4)
    if flag:
        r.render(49, 512 - 48); r.sync()
        d, i = r.accumulators()
        ref = (d + i) / 512.0
    st = r.stats()
    out["shadow_rays_per_frame_" + ("on" if flag else "off")] = int(st.shadow_rays // (512 if flag else 48))
    r.free()
lum = lambda a: a.reshape(-1, 3) @ np.array([0.2126, 0.7152, 0.0722])
L = lum(ref)
for flag in (False, True):
    e = lum(img[flag]) - L
    out["mean_" + ("on" if flag else "off")] = round(float(lum(img[flag]).mean()), 5)
    out["rel_rms_8spp_" + ("on" if flag else "off")] = round(float(np.sqrt((e ** 2).mean()) / L.mean()), 4)
out["mean_512spp_on"] = round(float(L.mean()), 5)
print("MEASURE_ENV_NEE " + json.dumps(out))
