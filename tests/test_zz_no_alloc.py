"""GPU: ptb_trace_closest_noalloc (static device scratch, no allocation; include/ptb200.h) against ptb_trace_closest, through the
C ABI.  No counterpart in the reference.  Collected last on
purpose; it builds its own scene and renderer, so it passes alone in a fresh process and after
the other tests.  Tolerance: the one the RayCast tests use -- hit ids equal, t / u / v bit for bit."""
import numpy as np
import pytest

import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen

pytestmark = pytest.mark.gpu


def _bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def _rays(n, seed):
    rng = np.random.RandomState(seed)
    ori = np.zeros((n, 4), np.float32)
    ori[:, :3] = rng.uniform([-4, 0.2, -4], [4, 4.5, 4], (n, 3))
    d = rng.normal(size=(n, 4)).astype(np.float32)
    d[:, :3] /= np.linalg.norm(d[:, :3], axis=1, keepdims=True)
    d[:50, 1] = 0.0                              # axis-parallel rays: inf / NaN slabs
    return ori, d


def test_trace_closest_noalloc_matches_trace_closest(tmp_path):
    scene = ptb.Scene(scenegen.fireplace(str(tmp_path / "fire"), res=(64, 36), scale=0.25)[0])
    r = ptb.Renderer(scene)
    # two calls in a row on the shared static array: a large batch, then a smaller one with other rays -- nothing of the first
    # call may leak into the second (the cursor is cleared per call, results are rewritten before they are read back)
    for n, seed in ((200_000, 1), (70_001, 2), (capi.NOALLOC_MAX_RAYS, 3)):
        ori, d = _rays(n, seed)
        hit, t, u, v = r.trace_closest(ori, d)
        h2, t2, u2, v2 = r.trace_closest_noalloc(ori, d)
        assert (hit >= 0).any() and (hit < 0).any()
        assert np.array_equal(hit, h2), (n, int((hit != h2).sum()))
        assert _bits(t, t2)
        m = hit >= 0
        assert _bits(u[m], u2[m]) and _bits(v[m], v2[m])
    # the same call twice gives the same bits; an empty batch is a no-op
    ori, d = _rays(33, 4)
    a = r.trace_closest_noalloc(ori, d)
    b = r.trace_closest_noalloc(ori, d)
    assert all(_bits(x, y) for x, y in zip(a, b))
    assert r.trace_closest_noalloc(ori[:0], d[:0])[0].size == 0
    # beyond the static array: a clear error, not a silent overflow
    ori, d = _rays(capi.NOALLOC_MAX_RAYS + 1, 5)
    with pytest.raises(capi.PtbError) as e:
        r.trace_closest_noalloc(ori, d)
    assert e.value.code == capi.ERR_INVALID and "PTB_NOALLOC_MAX_RAYS" in str(e.value)
    # a second renderer on the same GPU shares the array; the frame path is untouched by it
    r2 = ptb.Renderer(scene)
    ori, d = _rays(5000, 6)
    assert np.array_equal(r2.trace_closest_noalloc(ori, d)[0], r.trace_closest(ori, d)[0])
    r.render(1, 1)
    assert np.isfinite(r.resolve(1)).all()
    r2.free(); r.free(); scene.free()
