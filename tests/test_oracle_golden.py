"""CPU: the oracle (oracle/cpu/*.c) against golden vectors dumped from the UNMODIFIED reference build
(tests/golden/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

import _oracle as orc

CASES = ["cornell64", "fire_tiny"]


def _bits(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint8), np.ascontiguousarray(b).view(np.uint8))


@pytest.mark.parametrize("case", CASES)
def test_lbvh_bit_exact(golden, case):
    g = golden(case)
    b = orc.build_lbvh(g["mesh_pos"], g["mesh_indices_matid"])
    for name in ("aabb_min", "aabb_max", "centroid", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices", "world_aabb"):
        assert _bits(b[name], g["bvh_" + name]), name
    # the reference's own structural self-test (src/bvh.cu:291-414)
    assert orc.check_hierarchy(b["child_nodes"], b["parent"]) == 0
    assert orc.check_hierarchy(g["bvh_child_nodes"], g["bvh_parent"]) == 0


@pytest.mark.parametrize("case", CASES)
def test_morton_sorted_and_unique(golden, case):
    b = orc.build_lbvh(golden(case)["mesh_pos"], golden(case)["mesh_indices_matid"])
    m = b["morton_codes"]
    assert (np.diff(m.astype(np.uint64)) > 0).all()
    n = m.size
    assert np.array_equal(np.sort(b["primitive_indices"][:n]), np.arange(n))
    assert np.array_equal((m & np.uint64(0xffffffff)).astype(np.int32), b["primitive_indices"][:n])


@pytest.mark.parametrize("case", CASES)
def test_closest_hit_bit_exact(golden, case):
    g = golden(case)
    b = orc.build_lbvh(g["mesh_pos"], g["mesh_indices_matid"])
    hit, t, u, v = orc.trace_closest(b, g["mesh_pos"], g["mesh_indices_matid"], g["raygen_ray_ori"], g["raygen_ray_dir_dist"])
    assert np.array_equal(hit, g["s0_traced_hit_geom_id"])
    assert _bits(t, g["s0_traced_ray_dir_dist"][:, 3])          # t on a hit, -1 on a miss
    m = hit >= 0
    assert _bits(u[m], g["s0_traced_hit_normal"][m, 0]) and _bits(v[m], g["s0_traced_hit_normal"][m, 1])
    mat = g["mesh_indices_matid"][hit[m], 3]
    assert np.array_equal(mat, g["s0_traced_material_id"][m])
    assert orc.trace_closest.max_stack < 32                      # the reference's stack abort never triggers here


@pytest.mark.parametrize("case", CASES)
def test_closest_hit_bounce1(golden, case):
    g = golden(case)
    b = orc.build_lbvh(g["mesh_pos"], g["mesh_indices_matid"])
    q = g["s1_ext_queue"]
    hit, t, u, v = orc.trace_closest(b, g["mesh_pos"], g["mesh_indices_matid"], g["s0_shaded_ray_ori"][q], g["s0_shaded_ray_dir_dist"][q], threads=4)
    assert np.array_equal(hit, g["s1_traced_hit_geom_id"][q])
    assert _bits(t, g["s1_traced_ray_dir_dist"][q, 3])


@pytest.mark.parametrize("case", CASES)
def test_shadow_rays(golden, case):
    g = golden(case)
    b = orc.build_lbvh(g["mesh_pos"], g["mesh_indices_matid"])
    occ = orc.trace_occluded(b, g["mesh_pos"], g["mesh_indices_matid"], g["s1_shadow_ori_tmax"], g["s1_shadow_dir"])
    # a visible shadow ray adds its (non-zero) radiance to the direct image of its pixel
    pix = g["s1_shadow_pixel"]
    changed = (g["s1_direct"][pix] != g["s0_direct"][pix]).any(axis=1)
    assert np.array_equal(occ == 0, changed)
    expect = g["s0_direct"].copy()
    np.add.at(expect, pix[occ == 0], g["s1_shadow_radiance"][occ == 0, :3])
    assert _bits(expect, g["s1_direct"])


def test_raygen_seeds(golden):
    g = golden("cornell64")
    rng, jx, jy = orc.raygen_seeds(0, 64 * 64, 1)
    assert np.array_equal(rng, g["raygen_rng_state"])
    assert (np.abs(jx) <= 0.5).all() and (np.abs(jy) <= 0.5).all()


def test_lbvh_rejects_single_triangle():
    pos = np.array([[0, 0, 0, 1], [1, 0, 0, 1], [0, 1, 0, 1]], np.float32)
    with pytest.raises(ValueError):
        orc.build_lbvh(pos, np.array([[0, 1, 2, 0]], np.int32))


def test_lbvh_duplicate_centroids():
    """All-equal Morton codes: order falls back to the triangle index (src/bvh.cu:95)."""
    pos = np.array([[0, 0, 0, 1], [1, 0, 0, 1], [0, 1, 0, 1]], np.float32)
    im = np.tile(np.array([[0, 1, 2, 0]], np.int32), (9, 1))
    b = orc.build_lbvh(pos, im)
    assert np.array_equal(b["primitive_indices"][:9], np.arange(9))
    assert orc.check_hierarchy(b["child_nodes"], b["parent"]) == 0
