"""How often WalkWide::finish() / WalkAnyWide::confirm() take their re-walk paths (needs a -DPTB_WIDE_DEBUG build with PTB_WDBG counters\nin finish(): 0 rays, 1 near ties / overflow, 2 refused winners, 3 shadow hits confirmed, 4 refused; profiles/r02_raycast_notes.md section 5)."""
import ctypes as C, sys
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen, capi
for name, path in (("fireplace", scenegen.fireplace("/tmp/dbg_fire", res=(1920, 1080))[0]), ("sponza", scenegen.sponza("/tmp/dbg_sp", res=(1920, 1080), tex_size=64)[0]), ("cornell", scenegen.cornell("/tmp/dbg_c", res=(800, 800)))):
    r = ptb.Renderer(ptb.Scene(path))
    before = (C.c_uint64 * 8)(); capi.lib.ptb_debug_wide(before)
    r.render(1, 4); r.sync()
    out = (C.c_uint64 * 8)(); capi.lib.ptb_debug_wide(out)
    v = [int(a) - int(b) for a, b in zip(out, before)]
    print(name, "ext rays", v[0], "re-walked for near ties / overflow", v[1], "for the leaf-box test", v[2], "| shadow hits confirmed", v[3], "refused", v[4])
    r.free()
