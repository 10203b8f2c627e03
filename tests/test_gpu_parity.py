"""GPU: the CUDA path (through the C ABI) against the oracle, the golden fixtures and -- where the
prebuilt reference build oracle/_ref/refdump travelled to the box -- the unmodified reference itself.
Bars: integer/index work bit-exact; radiance bit-exact at 1 spp (north_star asks for <= 1e-4 relative
per pixel; the tolerance actually met is 0)."""
import os

import numpy as np
import pytest

import _oracle as orc
import _parity as par
import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, multigpu, scenegen

pytestmark = pytest.mark.gpu

LBVH_NAMES = ("aabb_min", "aabb_max", "centroid", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices", "world_aabb")


def _bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


@pytest.fixture(scope="module")
def fire_small(tmp_path_factory):
    return scenegen.fireplace(str(tmp_path_factory.mktemp("fire_small")), res=(480, 270), scale=0.25)[0]


@pytest.fixture(scope="module")
def sponza_small(tmp_path_factory):
    return scenegen.sponza(str(tmp_path_factory.mktemp("sponza_small")), res=(320, 180), scale=0.2, tex_size=128)[0]


@pytest.fixture(scope="module")
def sponza_full(tmp_path_factory):
    return scenegen.sponza(str(tmp_path_factory.mktemp("sponza_full")), res=(1920, 1080))[0]


@pytest.fixture(scope="module")
def fire_full(tmp_path_factory):
    return scenegen.fireplace(str(tmp_path_factory.mktemp("fire_full")), res=(1920, 1080))[0]


# ------------------------------------------------------------------------------------------ LBVH
@pytest.mark.parametrize("case,gen", [("cornell64", lambda d: scenegen.cornell(d, res=(64, 64))),
                                      ("fire_tiny", lambda d: scenegen.fireplace(d, res=(96, 54), scale=0.1)[0])])
def test_lbvh_and_radiance_vs_golden(golden, tmp_path, case, gen):
    g = golden(case)
    r = ptb.Renderer(ptb.Scene(gen(str(tmp_path / case))))
    for name in LBVH_NAMES:
        assert _bits(r.lbvh(name), g["bvh_" + name]), name
    r.render(1)
    d, i = r.accumulators()
    assert _bits(d, g["direct"]) and _bits(i, g["indirect"])
    r.free()


def test_lbvh_vs_oracle_143k(fire_full):
    s = ptb.Scene(fire_full)
    pos, im = s.array("pos"), s.array("indices_matid")
    ref = orc.build_lbvh(pos, im)
    out = ptb.build_lbvh(pos, im)            # BuildLBVH through the standalone C-ABI call, host buffers
    for name in ("aabb_min", "aabb_max", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices", "world_aabb"):
        assert _bits(out[name], ref[name]), name
    assert orc.check_hierarchy(out["child_nodes"], out["parent"]) == 0
    assert out["build_ms"] > 0


@pytest.mark.parametrize("n", [2, 3, 33, 4097, 300000])
def test_lbvh_vs_oracle_synthetic(n):
    """Ragged sizes around the sort tile (4096 keys), duplicates and degenerate extents."""
    pos, nor, idx, mat = scenegen.heightfield_arrays(max(n, 64), seed=n)
    idx = idx[:n]
    pos4 = np.concatenate([pos, np.ones((len(pos), 1), np.float32)], 1)
    im = np.concatenate([idx, np.zeros((n, 1), np.int32)], 1).astype(np.int32)
    if n == 33:
        im[:] = im[0]                         # every centroid identical: order decided by the index bits
    ref = orc.build_lbvh(pos4, im)
    out = ptb.build_lbvh(pos4, im)
    for name in ("aabb_min", "aabb_max", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices", "world_aabb"):
        assert _bits(out[name], ref[name]), (n, name)


def test_lbvh_single_triangle_is_rejected():
    pos4 = np.array([[0, 0, 0, 1], [1, 0, 0, 1], [0, 1, 0, 1]], np.float32)
    with pytest.raises(capi.PtbError) as e:
        ptb.build_lbvh(pos4, np.array([[0, 1, 2, 0]], np.int32))
    assert e.value.code == capi.ERR_UNSUPPORTED


# ------------------------------------------------------------------------------------------ RayCast
def test_raycast_vs_oracle(fire_small):
    s = ptb.Scene(fire_small)
    r = ptb.Renderer(s)
    pos, im = s.array("pos"), s.array("indices_matid")
    bvh = {k: r.lbvh(k) for k in LBVH_NAMES}
    rng = np.random.RandomState(3)
    n = 100000
    ori = np.zeros((n, 4), np.float32)
    ori[:, :3] = rng.uniform([-4, 0.2, -4], [4, 4.5, 4], (n, 3))
    d = rng.normal(size=(n, 4)).astype(np.float32)
    d[:, :3] /= np.linalg.norm(d[:, :3], axis=1, keepdims=True)
    d[:100, 0] = 0.0                            # axis-parallel rays: inf / NaN slabs
    hit, t, u, v = r.trace_closest(ori, d)
    h2, t2, u2, v2 = orc.trace_closest(bvh, pos, im, ori, d, threads=8)
    assert np.array_equal(hit, h2)
    m = hit >= 0
    assert _bits(t, t2) and _bits(u[m], u2[m]) and _bits(v[m], v2[m])
    ori[:, 3] = rng.uniform(0.05, 6.0, n)       # tmax
    occ = r.trace_occluded(ori, d)
    assert np.array_equal(occ, orc.trace_occluded(bvh, pos, im, ori, d, threads=8))
    hit0, _, _, _ = r.trace_closest(ori[:0], d[:0])   # empty batch
    assert hit0.size == 0
    r.free()


# ------------------------------------------------------------------------------------------ frames
def _need_ref():
    if not par.have_refdump():
        pytest.skip("oracle/_ref/refdump not present (built only where /root/reference exists)")


@pytest.mark.parametrize("which", ["cornell", "fire_small", "sponza_small"])
def test_staged_frame_vs_reference(which, fire_small, sponza_small, tmp_path):
    """Every stage of every bounce against the unmodified reference: queue counters equal, queue
    contents equal as sets, path state / shadow queue / accumulators bit-equal."""
    _need_ref()
    scene_json = {"cornell": lambda: scenegen.cornell(str(tmp_path / "c"), res=(200, 160)), "fire_small": lambda: fire_small,
                  "sponza_small": lambda: sponza_small}[which]()        # sponza: MTL materials, textures, normal maps, HDR env + pdf-map MIS
    d = str(tmp_path / "ref")
    log = par.run_refdump(scene_json, d, iters=2, stages=True, selfcheck=True)
    assert '"mismatching_floats": 0' in log
    ref = par.RefDump(d)
    s = ptb.Scene(scene_json)
    assert all(par.compare_scene(ref, s).values())
    r = ptb.Renderer(s)
    assert all(par.compare_lbvh(ref, r).values())
    r.render(1)
    rep = par.compare_staged_frame(ref, r, 2, ref.meta["trace_depth"])
    assert rep["ok"], par.failures(rep)
    dn, inn = r.accumulators()
    rr = par.radiance_report(ref.load("direct", "<f4", 3), ref.load("indirect", "<f4", 3), dn, inn, 2)
    assert rr["bit_mismatch_direct"] == 0 and rr["bit_mismatch_indirect"] == 0 and rr["max_rel_err"] <= 1e-4
    r.free()


def test_radiance_vs_reference_full_size(fire_full, tmp_path):
    """BASELINE config[1]: 143,173 triangles, 1920x1080, 8 bounces, MIS + NEE; 3 spp."""
    _need_ref()
    d = str(tmp_path / "ref")
    par.run_refdump(fire_full, d, iters=3, stages=False)
    ref = par.RefDump(d)
    r = ptb.Renderer(ptb.Scene(fire_full))
    r.render(1, 3)
    dn, inn = r.accumulators()
    rr = par.radiance_report(ref.load("direct", "<f4", 3), ref.load("indirect", "<f4", 3), dn, inn, 3)
    assert rr["max_rel_err"] <= 1e-4 and rr["rmse"] <= 1e-3, rr       # north_star tolerance
    assert rr["bit_mismatch_direct"] == 0 and rr["bit_mismatch_indirect"] == 0, rr   # what is actually met
    r.free()


def test_radiance_vs_reference_262k_textured(sponza_full, tmp_path):
    """BASELINE config[2]: 262,279 triangles, 15 bindless textures, HDR sky with alias/pdf map, 1920x1080, 8 bounces; 2 spp."""
    _need_ref()
    d = str(tmp_path / "ref")
    par.run_refdump(sponza_full, d, iters=2, stages=False)
    ref = par.RefDump(d)
    s = ptb.Scene(sponza_full)
    assert all(par.compare_scene(ref, s).values())
    r = ptb.Renderer(s)
    assert all(par.compare_lbvh(ref, r).values())
    r.render(1, 2)
    dn, inn = r.accumulators()
    rr = par.radiance_report(ref.load("direct", "<f4", 3), ref.load("indirect", "<f4", 3), dn, inn, 2)
    assert rr["max_rel_err"] <= 1e-4 and rr["rmse"] <= 1e-3, rr       # north_star tolerance
    assert rr["bit_mismatch_direct"] == 0 and rr["bit_mismatch_indirect"] == 0, rr
    r.free()


def test_full_size_properties(fire_full):
    """Size-independent properties at the benchmark size: determinism, graph == plain launches,
    staged == fused frame, tiles of a 3-way partition == the full frame, ray statistics."""
    s = ptb.Scene(fire_full)
    a = ptb.Renderer(s)
    a.render(1, 2)
    d0, i0 = a.accumulators()
    st = a.stats()
    assert st.frames == 2 and st.extension_rays > 2 * 1920 * 1080 and st.shadow_rays > 0 and st.kernel_launches > 0
    a.render(1, 2)
    d1, i1 = a.accumulators()
    assert _bits(d0, d1) and _bits(i0, i1)
    b = ptb.Renderer(s, use_graph=False)
    b.render(1, 2)
    d2, i2 = b.accumulators()
    assert _bits(d0, d2) and _bits(i0, i2)
    b.free()
    img = a.resolve(2)
    assert _bits(img, (d0 + i0) / np.float32(2))
    rgba = a.resolve(2, capi.IMG_RGBA8)
    assert rgba[:, 3].min() == 255
    full = np.zeros_like(img)
    for rank in range(3):
        t = ptb.Renderer(s, rank=rank, world=3, stripe_rows=8)
        t.render(1, 2)
        pix = t.owned_pixel_indices()
        assert np.array_equal(pix, multigpu.owned_pixels(rank, 3, 8, 1920, 1080))
        full[pix] = t.resolve(2)
        t.free()
    assert _bits(full, img)
    a.free()


def test_call_order_and_arguments(tmp_path):
    s = ptb.Scene(scenegen.cornell(str(tmp_path / "c"), res=(32, 32)))
    r = ptb.Renderer(s)
    with pytest.raises(capi.PtbError) as e:
        r.render(0)
    assert e.value.code == capi.ERR_INVALID
    r.stage(capi.STAGE_RAYGEN, 1)
    assert r.queue_count("extension") == 32 * 32 and r.queue_count("shadow") == 0
    for _ in range(8):
        r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    with pytest.raises(capi.PtbError) as e:
        r.stage(capi.STAGE_TRACE)
    assert e.value.code == capi.ERR_STATE
    with pytest.raises(capi.PtbError):
        ptb.Renderer(s, rank=2, world=2)
    r.free()


def test_sample_index_partition_matches_single_renderer(fire_small):
    """Two renderers taking alternate iterations (what two ranks of multigpu.SampleReduce do) sum to the
    1-renderer accumulation of the same four iterations up to float summation order."""
    s = ptb.Scene(fire_small)
    one = ptb.Renderer(s)
    one.render(1, 4)
    ref = one.resolve(4, capi.IMG_SUM_RGB32F)
    parts = []
    for rank in range(2):
        r = ptb.Renderer(s)
        r.clear()
        for it in (rank + 1, rank + 3):
            r.render(it)
        parts.append(r.resolve(1, capi.IMG_SUM_RGB32F))
        r.free()
    both = parts[0] + parts[1]
    err = np.abs(both - ref) / np.maximum(np.abs(ref), 1e-3)
    assert err.max() <= 1e-5, err.max()
    # iteration 1 alone must still be bit-identical, whichever renderer produced it
    a = ptb.Renderer(s)
    a.clear(); a.render(3)
    b = ptb.Renderer(s)
    b.render(1); b.clear(); b.render(3)
    assert _bits(a.resolve(1, capi.IMG_SUM_RGB32F), b.resolve(1, capi.IMG_SUM_RGB32F))
    one.free(); a.free(); b.free()


def test_stress_10m_lbvh_and_raycast_vs_oracle():
    """BASELINE config[3]: 10,000,000 procedural triangles injected as arrays.  LBVH bit-exact against the C oracle,
    closest / any hit bit-exact on a ray sample, hierarchy invariants of the reference's TestHierarchyLogic."""
    pos, nor, idx, mat = scenegen.heightfield_arrays(10_000_000)
    s = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera((640, 360)))
    r = ptb.Renderer(s)
    pos4, im = s.array("pos"), s.array("indices_matid")
    ref = orc.build_lbvh(pos4, im)
    for name in ("aabb_min", "aabb_max", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices", "world_aabb"):
        assert _bits(r.lbvh(name), ref[name]), name
    assert orc.check_hierarchy(ref["child_nodes"], ref["parent"]) == 0
    st = r.stats()
    assert st.num_triangles == 10_000_000 and 0 < st.lbvh_build_ms < 100 and st.lbvh_max_depth <= 64
    # the rays of a real first bounce
    r.stage(capi.STAGE_RAYGEN, 1)
    r.stage(capi.STAGE_TRACE); r.stage(capi.STAGE_LOGIC); r.stage(capi.STAGE_SHADE)
    q = r.queue("extension")[:60000]
    ori, d = r.path_field("ray_ori")[q].copy(), r.path_field("ray_dir_dist")[q].copy()
    hit, t, u, v = r.trace_closest(ori, d)
    h2, t2, u2, v2 = orc.trace_closest(ref, pos4, im, ori, d, threads=8)
    assert np.array_equal(hit, h2)
    m = hit >= 0
    assert _bits(t, t2) and _bits(u[m], u2[m]) and _bits(v[m], v2[m])
    ori[:, 3] = 25.0
    assert np.array_equal(r.trace_occluded(ori, d), orc.trace_occluded(ref, pos4, im, ori, d, threads=8))
    r.render(1, 2)
    a = r.resolve(2)
    assert np.isfinite(a).all() and a.max() > 0
    r.free()


def test_deep_tree_stack_abort_vs_reference(tmp_path):
    """The reference's closest-hit walk gives up with the best hit so far once 32 entries are stacked
    (src/kernels/ray_cast.cu:106).  scenegen.deepchain builds an LBVH 38 levels deep in which most camera rays run into
    that limit (58 % of the pixels show a hit that is NOT the closest one), so the reference's own abort is what pins
    raycast.cuh's: every stage of every bounce, both accumulators and the megakernel image, bit for bit."""
    _need_ref()
    scene_json = scenegen.deepchain(str(tmp_path / "chain"), res=(160, 120), depth=4)
    s = ptb.Scene(scene_json)
    pos, im = s.array("pos"), s.array("indices_matid")
    d = str(tmp_path / "ref")
    log = par.run_refdump(scene_json, d, iters=2, stages=True, selfcheck=True)
    assert '"mismatching_floats": 0' in log
    ref = par.RefDump(d)
    assert all(par.compare_scene(ref, s).values())
    r = ptb.Renderer(s)
    assert r.stats().lbvh_max_depth > 32
    assert all(par.compare_lbvh(ref, r).values())
    # the limit is live: the CPU restatement, which counts stack use, reaches 32 on the camera rays of this frame
    r.stage(capi.STAGE_RAYGEN, 1)
    ori, dirs = r.path_field("ray_ori"), r.path_field("ray_dir_dist")
    bvh = {k: r.lbvh(k) for k in LBVH_NAMES}
    hit_cpu, _, _, _ = orc.trace_closest(bvh, pos, im, ori, dirs, threads=8)
    assert orc.trace_closest.max_stack == 32
    hit_gpu, _, _, _ = r.trace_closest(ori, dirs)
    assert np.array_equal(hit_gpu, hit_cpu)
    r.render(1)
    rep = par.compare_staged_frame(ref, r, 2, ref.meta["trace_depth"])
    assert rep["ok"], par.failures(rep)
    dn, inn = r.accumulators()
    assert par.count_bit_mismatch(dn, ref.load("direct", "<f4", 3)) == 0 and par.count_bit_mismatch(inn, ref.load("indirect", "<f4", 3)) == 0
    assert dn.max() > 0
    r.free()
    # every scheduler / node format of the extension rays must fall back to the reference-order walk on such a tree
    base = (dn, inn)
    for env in ({"PTB_RC_MODE": "2", "PTB_RC_MINB": "8", "PTB_RC_REFILL": "16"}, {"PTB_RC_MODE": "1", "PTB_RC_REFILL": "24", "PTB_RC_STEPS": "4"}, {"PTB_WIDE": "3"}):
        got = _render_with(env, lambda: ptb.Renderer(s))
        assert _bits(got[0], base[0]) and _bits(got[1], base[1]), env
    mref, cnt = _mega_ref(scene_json, str(tmp_path / "mref"), 2)
    m = ptb.Renderer(s, megakernel=True)
    m.render(1, 2)
    assert par.count_bit_mismatch(m.accumulators()[0], mref) == 0
    m.free()


def test_stress_10m_vs_reference(tmp_path):
    """BASELINE config[3] against the reference ITSELF: the 10,000,000-triangle scene written as OBJ files, loaded by the
    reference's own loader and by this one, LBVH built by both (depth 35 > the 32-entry stack), 2 spp at 1920x1080.
    LBVH arrays are compared through position-weighted checksums (refdump --sums; dumping them would be 1.4 GB), the
    accumulators bit for bit."""
    _need_ref()
    import json as _json
    import shutil
    d_scene, d_ref = str(tmp_path / "stress"), str(tmp_path / "ref")
    scene_json, ntri = scenegen.stress_scene(d_scene, 10_000_000, res=(1920, 1080))
    try:
        log = par.run_refdump(scene_json, d_ref, iters=2, stages=False, extra=("--no-dump", "--sums"), timeout=1500)
        sums = _json.loads([l for l in log.splitlines() if l.startswith("REFSUMS")][0][len("REFSUMS "):])
        s = ptb.Scene(scene_json)
        assert s.info.num_triangles == sums["num_triangles"] == ntri and s.info.num_vertices == sums["num_vertices"]
        assert par.word_sum(s.array("pos")) == sums["pos"] and par.word_sum(s.array("indices_matid")) == sums["indices_matid"]
        r = ptb.Renderer(s)
        assert r.stats().lbvh_max_depth > 32
        for name in ("aabb_min", "aabb_max", "centroid", "primitive_indices", "morton_codes", "child_nodes", "parent", "escape_indices"):
            assert par.word_sum(r.lbvh(name)) == sums[name], name
        r.render(1, 2)
        dn, inn = r.accumulators()
        rd = np.fromfile(os.path.join(d_ref, "direct.bin"), dtype="<f4").reshape(-1, 3)
        ri = np.fromfile(os.path.join(d_ref, "indirect.bin"), dtype="<f4").reshape(-1, 3)
        rr = par.radiance_report(rd, ri, dn, inn, 2)
        assert rr["max_rel_err"] <= 1e-4 and rr["rmse"] <= 1e-3, rr       # north_star tolerance
        assert rr["bit_mismatch_direct"] == 0 and rr["bit_mismatch_indirect"] == 0, rr
        assert dn.max() > 0 and inn.max() > 0
        r.free(); s.free()
    finally:
        shutil.rmtree(d_scene, ignore_errors=True)


# ------------------------------------------------------------------------------------------ the reference-side shim
@pytest.mark.parametrize("which,mode", [("cornell", "file"), ("cornell", "live"), ("fire_small", "live"), ("sponza_small", "file"), ("sponza_small", "live")])
def test_shim_vs_reference(which, mode, fire_small, sponza_small, tmp_path):
    """The source-level drop-in: shim/pathtrace_wavefront_ptb200.cpp compiled against the reference's UNCHANGED headers and driven
    by a copy of its headless loop (shim/shim_driver.cpp: Scene(file) -> PathtraceInit -> Pathtrace(pbo, 0, 1..n) ->
    PathtraceFree) must leave the accumulators, and the RGBA8 image in the caller's device buffer, that the reference's own
    pathtrace_wavefront.cu leaves.  "live" hands the scene over from the reference's Scene object (arrays, textures read back
    from its CUDA arrays, alias table); "file" re-parses the scene file."""
    _need_ref()
    if not os.path.exists(par.SHIMDUMP):
        pytest.skip("oracle/_ref/shimdump not present")
    scene_json = {"cornell": lambda: scenegen.cornell(str(tmp_path / "c"), res=(200, 160)), "fire_small": lambda: fire_small,
                  "sponza_small": lambda: sponza_small}[which]()
    d = str(tmp_path / "ref")
    par.run_refdump(scene_json, d, iters=3, stages=False, views=True)
    ref = par.RefDump(d)
    sd, si, px = par.run_shimdump(scene_json, str(tmp_path / "shim"), 3, mode=mode)
    assert par.count_bit_mismatch(sd, ref.load("direct", "<f4", 3)) == 0
    assert par.count_bit_mismatch(si, ref.load("indirect", "<f4", 3)) == 0
    assert np.array_equal(px, ref.load("view0", "u1", 4))


# ------------------------------------------------------------------------------------------ headless views
@pytest.mark.parametrize("which,iters", [("cornell", 1), ("sponza_small", 1), ("sponza_small", 2)])
def test_gbuffer_views_and_heatmap_vs_reference(which, iters, sponza_small, tmp_path):
    """What the reference only shows through its GL preview: the first-bounce G-buffer its Logic stage writes, the
    display modes of SendImageToPBOKernel, and the VisualizeLBVH heat map -- against the reference's own kernels."""
    _need_ref()
    scene_json = scenegen.cornell(str(tmp_path / "c"), res=(200, 160)) if which == "cornell" else sponza_small
    d = str(tmp_path / "ref")
    par.run_refdump(scene_json, d, iters=iters, stages=False, views=True)
    ref = par.RefDump(d)
    r = ptb.Renderer(ptb.Scene(scene_json), gbuffer=True)
    r.render(1, iters)
    # iteration 1 sees prev_view_proj = identity (large motion); later ones prev == current, where the reference's
    # fma(u_cur, res, -(u_prev * res)) leaves the rounding error of the product, not an exact zero -- so do we
    assert _bits(r.gbuffer("depth"), ref.load("gb_depth", "<f4"))
    assert _bits(r.gbuffer("normal_matid"), ref.load("gb_normal_matid", "<f4", 4))
    assert _bits(r.gbuffer("albedo"), ref.load("gb_albedo", "<f4", 4))
    mine, theirs = r.gbuffer("motion"), ref.load("gb_motion", "<f4", 2)
    assert _bits(mine, theirs), float(np.abs(mine - theirs).max())
    if iters > 1:
        assert np.abs(mine).max() < 1e-3
    assert np.array_equal(r.resolve(iters, capi.IMG_RGBA8), ref.load("view0", "u1", 4))
    for k, fmt in enumerate((capi.IMG_VIEW_NORMALS, capi.IMG_VIEW_DEPTH, capi.IMG_VIEW_ALBEDO, capi.IMG_VIEW_MOTION), 1):
        assert np.array_equal(r.resolve(iters, fmt), ref.load(f"view{k}", "u1", 4)), k
    heat = ref.load("bvh_heatmap", "<f4", 3)
    assert _bits(r.visualize_lbvh(), heat)
    steps = r.visualize_lbvh(steps=True)
    assert steps.min() >= 0 and steps.max() > 1
    r.free()
    plain = ptb.Renderer(ptb.Scene(scene_json))
    with pytest.raises(capi.PtbError) as e:
        plain.gbuffer("depth")
    assert e.value.code == capi.ERR_STATE
    plain.free()


# ------------------------------------------------------------------------------------------ SVGF denoiser
@pytest.mark.parametrize("which,iters", [("cornell", 1), ("cornell", 6), ("sponza_small", 3), ("sponza_small", 8)])
def test_svgf_vs_reference(which, iters, sponza_small, tmp_path):
    """The reference's DenoiserOn mode through its untouched Pathtrace(): 1-spp frames, SVGF (temporal accumulation with
    history, spatial variance, 5 a-trous levels, modulation) into d_final_image.  Bit-exact, including frames where the
    temporal variance takes over from the spatial estimate (history >= 4)."""
    _need_ref()
    scene_json = scenegen.cornell(str(tmp_path / "c"), res=(200, 160)) if which == "cornell" else sponza_small
    d = str(tmp_path / "ref")
    par.run_refdump(scene_json, d, iters=iters, stages=False, denoise=True)
    ref = par.RefDump(d)
    r = ptb.Renderer(ptb.Scene(scene_json), denoiser=True)
    r.render(1, iters)
    dn, inn = r.accumulators()            # the last frame's 1-spp estimate: the accumulators are cleared every frame
    assert _bits(dn, ref.load("direct", "<f4", 3)) and _bits(inn, ref.load("indirect", "<f4", 3))
    assert _bits(r.resolve(iters, capi.IMG_DENOISED_RGB32F), ref.load("denoised", "<f4", 3))
    assert np.array_equal(r.resolve(iters, capi.IMG_DENOISED_RGBA8), ref.load("denoised_view0", "u1", 4))
    # iteration 1 restarts the history: rendering the same frames again gives the same image
    first = r.resolve(iters, capi.IMG_DENOISED_RGB32F)
    r.render(1, iters)
    assert _bits(r.resolve(iters, capi.IMG_DENOISED_RGB32F), first)
    r.free()
    with pytest.raises(capi.PtbError) as e:
        ptb.Renderer(ptb.Scene(scene_json), denoiser=True, rank=0, world=2)
    assert e.value.code == capi.ERR_UNSUPPORTED
    plain = ptb.Renderer(ptb.Scene(scene_json))
    with pytest.raises(capi.PtbError) as e:
        plain.resolve(1, capi.IMG_DENOISED_RGB32F)
    assert e.value.code == capi.ERR_STATE
    plain.free()


# ---- environment-map next-event estimation (behind a flag, it changes the image) --------------
def _furnace_scene(tmp, res=(96, 64), env_rgb=(1.0, 1.0, 1.0), albedo=0.6, depth=3):
    """One large diffuse quad seen from above under a constant environment: nothing occludes anything, so the exact
    radiance leaving the quad is albedo * L_env."""
    d = str(tmp)
    os.makedirs(d, exist_ok=True)
    scenegen.write_obj(os.path.join(d, "quad.obj"), *scenegen.grid([-50, 0, -50], [100, 0, 0], [0, 0, 100], 2, 2, [0, 1, 0]))
    hdr = np.empty((16, 32, 3), np.float32)
    hdr[:] = np.asarray(env_rgb, np.float32)
    scenegen.write_hdr(os.path.join(d, "env.hdr"), hdr)
    cam = {"RES": list(res), "FOVY": 30.0, "ITERATIONS": 1, "DEPTH": depth, "FILE": "furnace", "EYE": [0.0, 10.0, 0.01], "LOOKAT": [0.0, 0.0, 0.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(d, "scene.json")
    scenegen.write_scene(path, cam, {"m": scenegen._mat([albedo] * 3)}, [scenegen._obj_entry(os.path.join(d, "quad.obj"), "m")], env=os.path.join(d, "env.hdr"))
    return path


@pytest.mark.gpu
def test_sample_environment_vs_oracle(tmp_path):
    """ptb_sample_environment (the device's sampleEnvironmentMap) against the CPU restatement on the sky of the
    textured scene: chosen texel and pdf exactly, direction to a few ulp (sinf / cosf differ between libm and CUDA)."""
    path = scenegen.sponza(str(tmp_path / "sp"), res=(64, 36), scale=0.05, tex_size=32)[0]
    scene = ptb.Scene(path)
    r = ptb.Renderer(scene)
    rng = np.random.RandomState(5)
    rnd = rng.random_sample((200000, 2)).astype(np.float32)
    rnd[:4, 0] = [0.0, np.float32(1.0) - np.float32(2.0 ** -24), 0.5, 1.0]   # the clamp to N-1 included
    d, p, t = r.sample_environment(rnd)
    i = scene.info
    od, op, ot = orc.sample_environment(scene.array("env_probs"), scene.array("env_aliases"), scene.array("env_pdf"), i.env_width, i.env_height, rnd)
    assert np.array_equal(t, ot)
    assert np.array_equal(p, op)
    assert np.abs(d - od).max() < 4e-7
    # the sampler's own consistency: texel frequencies follow luminance * sin(theta), the table's target
    hist = np.bincount(t, minlength=i.env_width * i.env_height).astype(np.float64)
    row = hist.reshape(i.env_height, i.env_width).sum(axis=1)
    assert row.argmax() < i.env_height // 2      # the sun is in the upper hemisphere
    r.free(); scene.free()


@pytest.mark.gpu
def test_sample_environment_vs_reference(tmp_path):
    """ptb_sample_environment against the reference's OWN sampleEnvironmentMap (src/interactions.cu:447-488; defined there,
    called by nothing) run on the device by refdump --sample-env over the same random pairs: direction and pdf bit for bit."""
    _need_ref()
    path = scenegen.sponza(str(tmp_path / "sp"), res=(64, 36), scale=0.05, tex_size=32)[0]
    scene = ptb.Scene(path)
    r = ptb.Renderer(scene)
    rng = np.random.RandomState(9)
    rnd = rng.random_sample((300000, 2)).astype(np.float32)
    rnd[:4, 0] = [0.0, np.float32(1.0) - np.float32(2.0 ** -24), 0.5, 1.0]
    rnd.tofile(str(tmp_path / "rnd.bin"))
    par.run_refdump(path, str(tmp_path / "ref"), stages=False, extra=("--sample-env", str(tmp_path / "rnd.bin")))
    rd = np.fromfile(str(tmp_path / "ref" / "env_dir.bin"), dtype="<f4").reshape(-1, 3)
    rp = np.fromfile(str(tmp_path / "ref" / "env_pdf.bin"), dtype="<f4")
    d, p, t = r.sample_environment(rnd)
    assert _bits(p, rp)
    assert _bits(d, rd)
    r.free(); scene.free()


@pytest.mark.gpu
def test_env_nee_flag(tmp_path):
    """env_nee = 0 is the reference's estimator (checked bit for bit elsewhere).  env_nee = 1 must leave the indirect
    image bit-identical, only add to the direct image, be deterministic, and restore the energy the reference loses:
    on the furnace scene the mean converges to albedo * L, while the reference's estimator stays visibly below."""
    albedo, L = 0.6, 1.0
    scene = ptb.Scene(_furnace_scene(tmp_path / "furnace", albedo=albedo))
    spp = 64
    off = ptb.Renderer(scene)
    on = ptb.Renderer(scene, env_nee=True)
    on2 = ptb.Renderer(scene, env_nee=True, use_graph=False)
    for r in (off, on, on2):
        r.render(1, spp); r.sync()
    d0, i0 = off.accumulators()
    d1, i1 = on.accumulators()
    d2, i2 = on2.accumulators()
    assert np.array_equal(i0, i1) and np.array_equal(i1, i2)
    assert np.array_equal(d1, d2)
    assert (d0 == 0).all() and (d1 >= 0).all() and d1.mean() > 0     # no mesh lights: direct is the environment's estimate only
    mean_off = float((d0 + i0).mean() / spp)
    mean_on = float((d1 + i1).mean() / spp)
    expect = albedo * L
    assert abs(mean_on - expect) < 0.01 * expect, (mean_on, expect)
    assert mean_off < 0.9 * expect, (mean_off, expect)
    for r in (off, on, on2): r.free()
    scene.free()


@pytest.mark.gpu
def test_env_nee_staged_equals_whole_frame(tmp_path):
    """With env_nee the environment's shadow rays accumulate into a second half of `direct` that is folded in once per frame; the
    staged API folds at the same point (after the last TRACE), so a frame driven stage by stage leaves the whole frame's bits."""
    path = scenegen.sponza(str(tmp_path / "sp"), res=(160, 90), scale=0.1, tex_size=64)[0]
    scene = ptb.Scene(path)
    a = ptb.Renderer(scene, env_nee=True)
    a.render(1)
    da, ia = a.accumulators()
    b = ptb.Renderer(scene, env_nee=True)
    b.stage(capi.STAGE_RAYGEN, 1)
    for _ in range(scene.info.trace_depth):
        b.stage(capi.STAGE_TRACE); b.stage(capi.STAGE_LOGIC); b.stage(capi.STAGE_SHADE)
    db, ib = b.accumulators()
    assert da.max() > 0 and _bits(ia, ib) and _bits(da, db)
    a.free(); b.free(); scene.free()


# ---- the megakernel path (ptb_config.megakernel; reference: pathtrace_megakernel::Pathtrace, src/pathtrace.cu:645-706) ----
def _mega_ref(scene_json, d, iters):
    par.run_refdump(scene_json, d, iters=iters, mega=True)
    img = np.fromfile(os.path.join(d, "mega_image.bin"), dtype="<f4").reshape(-1, 3)
    cnt = np.fromfile(os.path.join(d, "mega_samples.bin"), dtype="<i4")
    return img, cnt


@pytest.mark.gpu
@pytest.mark.parametrize("which,iters", [("cornell", 1), ("cornell", 3), ("fire_small", 2), ("sponza_small", 2)])
def test_megakernel_vs_reference(which, iters, fire_small, sponza_small, tmp_path):
    """The float image of the reference's own megakernel (d_image after `iters` samples, dumped by refdump --mega) against
    the direct accumulator of a megakernel renderer: bit for bit.  Cornell: diffuse + light; fire_small: the four material
    models, MIS + NEE; sponza_small: MTL materials, textures, normal maps and the (unweighted) HDR environment."""
    _need_ref()
    scene_json = {"cornell": lambda: scenegen.cornell(str(tmp_path / "c"), res=(200, 160)), "fire_small": lambda: fire_small,
                  "sponza_small": lambda: sponza_small}[which]()
    ref, cnt = _mega_ref(scene_json, str(tmp_path / "ref"), iters)
    assert (cnt == iters).all()           # the reference's per-pixel sample count is the iteration (ptb_resolve divides by it)
    r = ptb.Renderer(ptb.Scene(scene_json), megakernel=True)
    r.render(1, iters)
    d, i = r.accumulators()
    assert (i == 0).all()                 # everything lands in the direct accumulator (the reference's single d_image)
    assert par.count_bit_mismatch(d, ref) == 0
    mean = r.resolve(iters)
    assert _bits(mean, d / np.float32(iters))
    st = r.stats()
    assert st.frames == iters and st.extension_rays >= iters * ref.shape[0] and st.kernel_launches > 0
    r.free()


@pytest.mark.gpu
def test_megakernel_full_size_vs_reference(fire_full, tmp_path):
    """BASELINE config[1] through the megakernel path: 143,173 triangles, 1920x1080, 8 bounces, 2 spp."""
    _need_ref()
    ref, cnt = _mega_ref(fire_full, str(tmp_path / "ref"), 2)
    r = ptb.Renderer(ptb.Scene(fire_full), megakernel=True)
    r.render(1, 2)
    d, _ = r.accumulators()
    rel = np.abs(d - ref) / np.maximum(np.abs(ref), 1e-6)
    assert rel.max() <= 1e-4                                 # north_star tolerance
    assert par.count_bit_mismatch(d, ref) == 0               # what is actually met
    r.free()


@pytest.mark.gpu
def test_megakernel_tile_partition_and_modes(fire_small):
    """Tiles of a megakernel frame are bit-identical to the full frame (global pixel indices seed the paths); the graph and
    plain-launch forms agree; the options the reference's megakernel does not have are rejected."""
    s = ptb.Scene(fire_small)
    full = ptb.Renderer(s, megakernel=True)
    full.render(1, 2)
    ref, _ = full.accumulators()
    out = np.zeros_like(ref)
    for rank in range(3):
        t = ptb.Renderer(s, megakernel=True, rank=rank, world=3, stripe_rows=8)
        t.render(1, 2)
        d, _ = t.accumulators()
        out[t.owned_pixel_indices()] = d
        t.free()
    assert _bits(out, ref)
    plain = ptb.Renderer(s, megakernel=True, use_graph=False)
    plain.render(1, 2)
    assert _bits(plain.accumulators()[0], ref)
    with pytest.raises(capi.PtbError) as e:
        plain.stage(capi.STAGE_RAYGEN, 1)
    assert e.value.code == capi.ERR_STATE
    for kw in ({"gbuffer": True}, {"denoiser": True}, {"env_nee": True}):
        with pytest.raises(capi.PtbError) as e:
            ptb.Renderer(s, megakernel=True, **kw)
        assert e.value.code == capi.ERR_UNSUPPORTED
    full.free(); plain.free()


# ---- the 4-wide quantised tree (wide.cuh) against the binary walk ------------------------------------------------------
def _render_with(env, make, iters=2):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)          # the traversal knobs are read once, in ptb_init
    try:
        r = make()
        r.render(1, iters)
        d, i = r.accumulators()
        r.free()
        return d, i
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["fire_small", "sponza_small", "heightfield_2m"])
def test_wide_tree_matches_binary_walk(which, fire_small, sponza_small):
    """Shadow rays (the default), extension rays and the megakernel's visibility rays on the 4-wide 15-bit-quantised collapse of
    the LBVH give the accumulators of the binary walk bit for bit: inner boxes are only conservative, a hit is accepted through
    the reference's test on the leaf's fp32 box, near-ties are re-traced in reference order.  The
    height field is the hard case for the quantisation: 2 M triangles, a triangle is 23 grid steps wide, +-2 steps of margin."""
    if which == "heightfield_2m":
        pos, nor, idx, mat = scenegen.heightfield_arrays(2_000_000)
        scene = ptb.Scene.from_arrays(pos, nor, idx, mat, scenegen.stress_materials(), scenegen.stress_camera((640, 360)))
    else:
        scene = ptb.Scene({"fire_small": fire_small, "sponza_small": sponza_small}[which])
    ref = _render_with({"PTB_WIDE": "0"}, lambda: ptb.Renderer(scene))
    for bits in ("1", "2", "3"):
        got = _render_with({"PTB_WIDE": bits}, lambda: ptb.Renderer(scene))
        assert _bits(got[0], ref[0]) and _bits(got[1], ref[1]), f"PTB_WIDE={bits}"
    mref = _render_with({"PTB_WIDE": "0"}, lambda: ptb.Renderer(scene, megakernel=True))
    mgot = _render_with({"PTB_WIDE": "4"}, lambda: ptb.Renderer(scene, megakernel=True))
    assert _bits(mgot[0], mref[0])


# ---- ragged sizes: resolutions that are not multiples of the 8x4 queue tile, short and long bounce budgets -------------
@pytest.mark.gpu
@pytest.mark.parametrize("res,depth", [((201, 163), 3), ((67, 5), 1), ((8, 4), 8), ((333, 2), 12)])
def test_ragged_resolution_and_depth_vs_reference(res, depth, tmp_path):
    """The first queue is laid out in 8x4 pixel tiles with a right strip and bottom rows in linear order (tiled_slot); the frame
    graph has 4 * depth - 2 kernel nodes.  Both render paths against the reference's own builds on resolutions that exercise every
    branch of that layout and on trace depths 1, 3, 8 and 12: bit for bit."""
    _need_ref()
    scene_json = scenegen.cornell(str(tmp_path / "c"), res=res, depth=depth)
    d = str(tmp_path / "ref")
    par.run_refdump(scene_json, d, iters=2, stages=False)
    ref = par.RefDump(d)
    r = ptb.Renderer(ptb.Scene(scene_json))
    r.render(1, 2)
    dn, inn = r.accumulators()
    assert par.count_bit_mismatch(dn, ref.load("direct", "<f4", 3)) == 0 and par.count_bit_mismatch(inn, ref.load("indirect", "<f4", 3)) == 0
    r.free()
    mref, cnt = _mega_ref(scene_json, str(tmp_path / "mref"), 2)
    m = ptb.Renderer(ptb.Scene(scene_json), megakernel=True)
    m.render(1, 2)
    assert par.count_bit_mismatch(m.accumulators()[0], mref) == 0
    m.free()


# ---- multi-GPU inside the C ABI: peer-store gather, hybrid tile x sample partition (include/ptb200.h "multi-GPU") -------
def _devices(n):
    have = capi.lib.ptb_device_count()
    return [k % have for k in range(n)]          # one GPU: the renderers share it (the gather is then a local store)


@pytest.mark.gpu
def test_multi_renderer_tile_and_hybrid(fire_small):
    """ptb_multi_* through ctypes only (no torch in this path): a frame cut into 4 stripe sets is bit-identical to one renderer
    in every output format; 2 tile groups x 2 sample groups agree with one renderer's 4 spp to summation order, twice the same
    bits; the no-P2P copy route gives the peer-store route's bits."""
    s = ptb.Scene(fire_small)
    one = ptb.Renderer(s)
    one.render(1, 3)
    ref_sum, ref_mean, ref_rgba = one.resolve(3, capi.IMG_SUM_RGB32F), one.resolve(3), one.resolve(3, capi.IMG_RGBA8)
    m = ptb.MultiRenderer(s, _devices(4))
    assert (m.tile_groups, m.sample_groups) == (4, 1)
    m.render(1, 3)
    assert _bits(m.resolve(capi.IMG_SUM_RGB32F), ref_sum)
    assert _bits(m.resolve(capi.IMG_MEAN_RGB32F), ref_mean)
    assert np.array_equal(m.resolve(capi.IMG_RGBA8), ref_rgba)
    info = m.info
    assert info.samples == 3 and info.n_devices == 4 and info.peer_stores == 4
    assert sum(m.stats(k).extension_rays for k in range(4)) == one.stats().extension_rays
    m.render(1, 1)                                  # step 1 restarts the accumulation
    one.render(1, 1)
    assert _bits(m.resolve(capi.IMG_SUM_RGB32F), one.resolve(1, capi.IMG_SUM_RGB32F))
    m.free()
    one.render(1, 4)
    ref4 = one.resolve(4, capi.IMG_SUM_RGB32F)
    h = ptb.MultiRenderer(s, _devices(4), tile_groups=2)
    assert (h.tile_groups, h.sample_groups) == (2, 2)
    h.render(1, 2)                                  # 2 steps x 2 sample groups = iterations 1..4
    a = h.resolve(capi.IMG_SUM_RGB32F)
    err = np.abs(a - ref4) / np.maximum(np.abs(ref4), 1e-3)
    assert err.max() <= 1e-5, err.max()
    h.render(1, 2)
    assert _bits(h.resolve(capi.IMG_SUM_RGB32F), a)
    assert h.info.samples == 4
    h.free()
    os.environ["PTB_MULTI_NO_P2P"] = "1"
    try:
        c = ptb.MultiRenderer(s, _devices(4), tile_groups=2)
        c.render(1, 2)
        assert c.info.peer_stores == 1              # only the root itself
        assert _bits(c.resolve(capi.IMG_SUM_RGB32F), a)
        c.free()
    finally:
        del os.environ["PTB_MULTI_NO_P2P"]
    with pytest.raises(capi.PtbError):
        ptb.MultiRenderer(s, _devices(4), tile_groups=3)
    one.free(); s.free()


_IPC_CHILD = """
import sys
sys.path.insert(0, {root!r})
import cuda_path_tracer_b200 as ptb
s = ptb.Scene({scene!r})
r = ptb.Renderer(s, rank=1, world=2, stripe_rows=8)
r.render(1, 2); 
g = ptb.Gather.open(0, bytes.fromhex({handle!r}), r.width, r.height, 1)
g.resolve_from(r, 0)
r.sync()
print("CHILD_OK")
g.free(); r.free()
"""


@pytest.mark.gpu
def test_gather_across_processes_ipc(fire_small):
    """One process per GPU: the root exports its frame buffer (ptb_gather_export, a CUDA IPC handle), another PROCESS opens it
    and its resolve kernel stores its stripes into the root's memory; the combined frame is the one-renderer frame, bit for bit."""
    import subprocess
    import sys
    s = ptb.Scene(fire_small)
    one = ptb.Renderer(s)
    one.render(1, 2)
    ref = one.resolve(2)
    r0 = ptb.Renderer(s, rank=0, world=2, stripe_rows=8)
    g = ptb.Gather(0, r0.width, r0.height, 1)
    code = _IPC_CHILD.format(root=os.path.dirname(os.path.dirname(os.path.abspath(__file__))), scene=fire_small, handle=g.export().hex())
    p = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert "CHILD_OK" in p.stdout, p.stdout[-1000:] + p.stderr[-2000:]
    r0.render(1, 2)
    g.resolve_from(r0, 0)
    r0.sync()
    assert _bits(g.combine(2), ref)
    g.free(); r0.free(); one.free(); s.free()
