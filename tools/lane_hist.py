"""Histogram of live lanes per warp-iteration of trace_closest() (needs a -DPTB_WIDE_DEBUG build: make EXTRA=-DPTB_WIDE_DEBUG
with the __activemask() counter of raycast.cuh enabled); DESIGN.md section 5b."""
import ctypes as C, sys
sys.path.insert(0, "/root/repo")
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen, capi
r = ptb.Renderer(ptb.Scene(scenegen.fireplace("/tmp/dbg_fire", res=(1920, 1080))[0]))
r.render(1, 1); r.sync()
out = (C.c_uint64 * 8)()
capi.lib.ptb_debug_wide(out)
v = list(out); tot = sum(v)
print("warp-iterations by active lanes (1-4,5-8,...,29-32):", v, [round(x / tot, 3) for x in v])
