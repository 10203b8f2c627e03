"""ctypes wrapper of the CPU oracle (oracle/libptoracle.so, built from oracle/cpu/*.c).

Test infrastructure: imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "libptoracle.so")
if not os.path.exists(LIB):
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "cpu"], check=True)
_lib = C.CDLL(LIB)
_P = C.c_void_p


def _p(a):
    return a.ctypes.data_as(_P)


def build_lbvh(pos4, indices_matid4):
    pos4 = np.ascontiguousarray(pos4, np.float32).reshape(-1, 4)
    im = np.ascontiguousarray(indices_matid4, np.int32).reshape(-1, 4)
    n = im.shape[0]
    out = dict(aabb_min=np.zeros((2 * n, 4), np.float32), aabb_max=np.zeros((2 * n, 4), np.float32),
               centroid=np.zeros((2 * n, 4), np.float32), primitive_indices=np.zeros(2 * n, np.int32),
               morton_codes=np.zeros(n, np.uint64), child_nodes=np.zeros((max(n - 1, 1), 2), np.int32),
               parent=np.zeros(2 * n, np.int32), escape_indices=np.zeros(2 * n, np.int32), world_aabb=np.zeros((2, 4), np.float32))
    rc = _lib.ptb_oracle_build_lbvh(_p(pos4), pos4.shape[0], _p(im), n, _p(out["aabb_min"]), _p(out["aabb_max"]), _p(out["centroid"]),
                                    _p(out["primitive_indices"]), _p(out["morton_codes"]), _p(out["child_nodes"]), _p(out["parent"]),
                                    _p(out["escape_indices"]), _p(out["world_aabb"]))
    if rc != 0:
        raise ValueError("oracle LBVH needs at least 2 triangles")
    out["child_nodes"] = out["child_nodes"][:n - 1]
    return out


def check_hierarchy(child_nodes, parent):
    child = np.ascontiguousarray(child_nodes, np.int32)
    par = np.ascontiguousarray(parent, np.int32)
    return _lib.ptb_oracle_check_hierarchy(_p(child), _p(par), par.size // 2)


def trace_closest(bvh, pos4, im4, ori4, dir4, threads=1):
    pos4 = np.ascontiguousarray(pos4, np.float32); im4 = np.ascontiguousarray(im4, np.int32)
    ori4 = np.ascontiguousarray(ori4, np.float32).reshape(-1, 4); dir4 = np.ascontiguousarray(dir4, np.float32).reshape(-1, 4)
    n, count = im4.shape[0], ori4.shape[0]
    hit = np.zeros(count, np.int32)
    t, u, v = (np.zeros(count, np.float32) for _ in range(3))
    child = np.ascontiguousarray(bvh["child_nodes"], np.int32)
    trace_closest.max_stack = _lib.ptb_oracle_trace_closest(_p(bvh["aabb_min"]), _p(bvh["aabb_max"]), _p(bvh["primitive_indices"]), _p(child),
                                                            _p(pos4), _p(im4), n, _p(ori4), _p(dir4), count, _p(hit), _p(t), _p(u), _p(v), threads)
    return hit, t, u, v


def trace_occluded(bvh, pos4, im4, ori_tmax4, dir4, threads=1):
    pos4 = np.ascontiguousarray(pos4, np.float32); im4 = np.ascontiguousarray(im4, np.int32)
    ori = np.ascontiguousarray(ori_tmax4, np.float32).reshape(-1, 4); dir4 = np.ascontiguousarray(dir4, np.float32).reshape(-1, 4)
    n, count = im4.shape[0], ori.shape[0]
    occ = np.zeros(count, np.int32)
    child = np.ascontiguousarray(bvh["child_nodes"], np.int32)
    _lib.ptb_oracle_trace_occluded(_p(bvh["aabb_min"]), _p(bvh["aabb_max"]), _p(bvh["primitive_indices"]), _p(child), _p(bvh["escape_indices"]),
                                   _p(pos4), _p(im4), n, _p(ori), _p(dir4), count, _p(occ), threads)
    return occ


def raygen_seeds(first, count, iteration):
    rng = np.zeros(count, np.uint32)
    jx, jy = np.zeros(count, np.float32), np.zeros(count, np.float32)
    _lib.ptb_oracle_raygen_seeds(first, count, iteration, _p(rng), _p(jx), _p(jy))
    return rng, jx, jy


def sample_environment(probs, aliases, pdf_map, width, height, rnd):
    """sampleEnvironmentMap (src/interactions.cu:447-488) on the host-built alias table."""
    probs = np.ascontiguousarray(probs, np.float32); aliases = np.ascontiguousarray(aliases, np.int32)
    pdf_map = np.ascontiguousarray(pdf_map, np.float32); rnd = np.ascontiguousarray(rnd, np.float32).reshape(-1, 2)
    n = rnd.shape[0]
    d, p, t = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
    _lib.ptb_oracle_sample_environment(_p(probs), _p(aliases), _p(pdf_map), int(width), int(height), _p(rnd), n, _p(d), _p(p), _p(t))
    return d, p, t
