"""SVGF cost at 1920x1080 on the 262k-triangle textured scene: this library (frame with / without the denoiser, CUDA
events) beside the reference's own DenoiserOn mode (oracle/_ref/refdump --denoise-bench)."""
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

REFDUMP = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "refdump")
p = scenegen.sponza("/tmp/msv_sponza")[0]
scene = ptb.Scene(p)
out = {}
for den in (False, True):
    r = ptb.Renderer(scene, denoiser=den)
    r.render(1, 5); r.sync()
    t0 = time.perf_counter(); r.render(6, 30); r.sync()
    out["ms_per_frame_denoised" if den else "ms_per_frame_plain"] = round((time.perf_counter() - t0) / 30 * 1e3, 4)
    r.free()
out["svgf_ms"] = round(out["ms_per_frame_denoised"] - out["ms_per_frame_plain"], 4)
out["svgf_note"] = "includes the per-frame accumulator clear and the G-buffer kernel that the plain frame skips"
print("MEASURE_SVGF " + json.dumps(out), flush=True)
if os.path.exists(REFDUMP):
    q = subprocess.run([REFDUMP, p, "/tmp/msv_ref", "--iters", "3", "--denoise-bench", "30"], capture_output=True, text=True, timeout=1200)
    for l in q.stdout.splitlines():
        if l.startswith("REFDENOISE"):
            print(l, flush=True)
