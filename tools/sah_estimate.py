"""Summed internal surface area of the LBVH against a PLOC tree (radius argv[2]) over the same Morton-sorted leaves, numpy, CPU only:\nwould a better tree under the 4-wide records pay?  (profiles/r02_raycast_notes.md section 5: 0.99x / 0.88x -- no.)  Uses the CPU oracle: a tool, not product."""
import sys, time
ROOT = __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tests")
import numpy as np
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import scenegen
import _oracle

which = sys.argv[1] if len(sys.argv) > 1 else "fireplace"
if which == "fireplace":
    s = ptb.Scene(scenegen.fireplace("/tmp/sah_fire", res=(64, 36))[0])
else:
    s = ptb.Scene(scenegen.sponza("/tmp/sah_sp", res=(64, 36), tex_size=32)[0])
pos, im = s.array("pos"), s.array("indices_matid")
n = im.shape[0]
b = _oracle.build_lbvh(pos, im)
mn, mx = b["aabb_min"][:, :3], b["aabb_max"][:, :3]
def sa(lo, hi):
    d = np.maximum(hi - lo, 0); return 2 * (d[:, 0] * d[:, 1] + d[:, 1] * d[:, 2] + d[:, 2] * d[:, 0])
ch = b["child_nodes"]
root_sa = sa(mn[:1], mx[:1])[0]
print("n", n, "child range", ch.min(), ch.max())
lbvh_int = sa(mn[:n - 1], mx[:n - 1]).sum() / root_sa
leaf_sa = sa(mn[n - 1:2 * n - 1], mx[n - 1:2 * n - 1]).sum() / root_sa
print("LBVH: sum SA internal / root", lbvh_int, "leaves", leaf_sa)
# wide estimate: internal nodes at even depth
depth = np.zeros(2 * n - 1, np.int32)
order = [0]
# BFS
import collections
q = collections.deque([0])
while q:
    i = q.popleft()
    for c in ch[i]:
        if c < 0: continue
        depth[c] = depth[i] + 1
        q.append(c)
ev = (depth[:n - 1] % 2 == 0)
print("LBVH: even-depth internal SA", sa(mn[:n - 1][ev], mx[:n - 1][ev]).sum() / root_sa, "max depth", depth.max())

# PLOC over the Morton-sorted leaves
lo = mn[n - 1:2 * n - 1].astype(np.float64).copy(); hi = mx[n - 1:2 * n - 1].astype(np.float64).copy()
R = int(sys.argv[2]) if len(sys.argv) > 2 else 16
tot = 0.0; it = 0; t0 = time.time()
dep = np.zeros(n, np.int32)
while lo.shape[0] > 1:
    m = lo.shape[0]
    best = np.full(m, np.inf); nn = np.full(m, -1)
    for off in range(1, R + 1):
        if off >= m: break
        u_lo = np.minimum(lo[:-off], lo[off:]); u_hi = np.maximum(hi[:-off], hi[off:])
        a = sa(u_lo, u_hi)
        # pair (i, i+off)
        idx = np.arange(m - off)
        better = a < best[idx]
        best[idx[better]] = a[better]; nn[idx[better]] = idx[better] + off
        j = idx + off
        better = a < best[j]
        best[j[better]] = a[better]; nn[j[better]] = idx[better]
    i = np.arange(m)
    mutual = (nn[nn] == i) & (i < nn)
    keep = np.ones(m, bool)
    a_i = i[mutual]; b_i = nn[mutual]
    lo[a_i] = np.minimum(lo[a_i], lo[b_i]); hi[a_i] = np.maximum(hi[a_i], hi[b_i])
    dep[a_i] = np.maximum(dep[a_i], dep[b_i]) + 1
    tot += sa(lo[a_i], hi[a_i]).sum()
    keep[b_i] = False
    lo, hi, dep = lo[keep], hi[keep], dep[keep]
    it += 1
print("PLOC r=%d: sum SA internal / root %.3f, iterations %d, depth %d, %.1fs" % (R, tot / root_sa, it, dep[0], time.time() - t0))
print("ratio PLOC / LBVH internal SA: %.3f" % (tot / root_sa / lbvh_int))
