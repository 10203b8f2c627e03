"""How many popped stack entries of the 4-wide closest-hit walk could have been culled by their entry distance (needs a\n-DPTB_WIDE_DEBUG -DPTB_EXP_POPCULL build of raycast.cu with the counting variant of WalkWide::step; profiles/r02_raycast_notes.md section 5)."""
import ctypes as C, sys
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen, capi
r = ptb.Renderer(ptb.Scene(scenegen.fireplace("/tmp/dbg_fire", res=(1920, 1080))[0]))
r.render(1, 1); r.sync()
out = (C.c_uint64 * 8)()
capi.lib.ptb_debug_wide(out)
v = list(out)
print("node steps", v[0], "popped", v[1], "cullable", v[2], "| leaf steps", v[3], "popped", v[4], "cullable", v[5])
print("cullable share of all steps: %.3f (nodes %.3f of node steps, leaves %.3f of leaf steps)" % ((v[2]+v[5])/(v[0]+v[3]), v[2]/v[0], v[5]/v[3]))
