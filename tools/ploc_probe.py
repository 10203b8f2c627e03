"""SAH cost of the reference-exact LBVH against a PLOC tree over the same Morton order (CPU, numpy): would a better
traversal tree pay?  usage: ploc_probe.py [fire|sponza]   (no)"""
import sys, time
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import _oracle as orc
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen

which = sys.argv[1] if len(sys.argv) > 1 else "fire"
if which == "fire":
    path = scenegen.fireplace("/tmp/ploc_fire", res=(64, 36))[0]
else:
    path = scenegen.sponza("/tmp/ploc_sponza", res=(64, 36), tex_size=16)[0]
s = ptb.Scene(path)
pos, im = s.array("pos"), s.array("indices_matid")
t0 = time.time()
bvh = orc.build_lbvh(pos, im)
N = im.shape[0]
print("N", N, "lbvh build (cpu oracle)", time.time() - t0)
amin, amax = bvh["aabb_min"][:, :3].astype(np.float64), bvh["aabb_max"][:, :3].astype(np.float64)

def area(mn, mx):
    d = mx - mn
    return 2 * (d[..., 0] * d[..., 1] + d[..., 1] * d[..., 2] + d[..., 2] * d[..., 0])

root_area = area(amin[0], amax[0])
sah_lbvh_inner = area(amin[:N - 1], amax[:N - 1]).sum() / root_area
sah_leaves = area(amin[N:], amax[N:]).sum() / root_area
print("LBVH   SAH inner %.2f  leaves %.2f" % (sah_lbvh_inner, sah_leaves))

# PLOC over the Morton-sorted leaves
def ploc(cmin, cmax, r):
    n = cmin.shape[0]
    inner_area = 0.0
    it = 0
    while n > 1:
        best = np.full(n, np.inf); nn = np.full(n, -1)
        for off in range(1, r + 1):
            if off >= n: break
            a = area(np.minimum(cmin[:-off], cmin[off:]), np.maximum(cmax[:-off], cmax[off:]))
            # candidate for i: i+off ; for i+off: i
            idx = np.arange(n - off)
            m = a < best[idx]; best[idx[m]] = a[m]; nn[idx[m]] = idx[m] + off
            j = idx + off
            m = a < best[j]; best[j[m]] = a[m]; nn[j[m]] = idx[m]
        i = np.arange(n)
        mutual = (nn[nn] == i)
        lead = mutual & (i < nn)
        gone = mutual & (i > nn)
        newmin = cmin.copy(); newmax = cmax.copy()
        li = i[lead]; lj = nn[lead]
        newmin[li] = np.minimum(cmin[li], cmin[lj]); newmax[li] = np.maximum(cmax[li], cmax[lj])
        inner_area += area(newmin[li], newmax[li]).sum()
        keep = ~gone
        cmin, cmax = newmin[keep], newmax[keep]
        n = cmin.shape[0]; it += 1
    return inner_area, it

for r in (4, 10, 16):
    t0 = time.time()
    ia, it = ploc(amin[N:].copy(), amax[N:].copy(), r)
    print("PLOC r=%d SAH inner %.2f (%.2fx of LBVH) iterations %d  %.1fs" % (r, ia / root_area, ia / root_area / sah_lbvh_inner, it, time.time() - t0))
