"""GPU: the loader's vertex de-duplication and repack on the device (ptb_scene_load_device, csrc/cuda/dedup.cu) against the host
threads' (ptb_scene_load), through the C ABI.  The host path is itself pinned against the reference's loader
(tests/test_host.py: vertex numbering of src/scene.cpp:525-532, upload arrays of src/wavefront_internal.cu:118-167), and where
the reference harness is present the reference's own loader checksum is compared directly.  Bar: every array bit for bit --
vertex ids are integers, and the floats (NaNs of normalised zero vectors included) are the host's bits."""
import json
import os
import subprocess

import numpy as np
import pytest

import cuda_path_tracer_b200 as ptb
from cuda_path_tracer_b200 import capi, scenegen
from cuda_path_tracer_b200.wavefront import SA

from _parity import REFDUMP, ROOT, have_refdump

pytestmark = pytest.mark.gpu


def _bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))


def _same_scene(host, dev, tmp_path):
    for name in SA:
        assert _bits(host.array(name), dev.array(name)), name
    for f, _ in capi.SceneInfo._fields_:
        a, b = getattr(host.info, f), getattr(dev.info, f)
        assert (list(a) == list(b)) if hasattr(a, "__len__") else (a == b), f
    # the cache file holds the whole container, the pre-normalisation normals / tangents included
    host.save_cache(str(tmp_path / "host.ptbscene"))
    dev.save_cache(str(tmp_path / "dev.ptbscene"))
    assert open(tmp_path / "host.ptbscene", "rb").read() == open(tmp_path / "dev.ptbscene", "rb").read()


HOSTILE_OBJ = """mtllib hostile.mtl
# -0 and +0 are the same vertex; the first occurrence's bits are the ones kept
v 0 0 0
v -0 0 -0
v 1 0 0
v 0 1 0
v 1 1 0
v 0.5 0.5 1e-39
# a zero normal normalises to NaN: such corners equal nothing, not even each other
vn 0 0 0
vn 0 0 1
vn 0 0 2
vt 0 0
vt 1 0
vt 0 1
vt -0 0
usemtl red
f 1/1/2 3/2/2 4/3/2
f 2/4/2 3/2/2 4/3/2
f 1/1/1 3/2/1 4/3/1
f 1/1/1 3/2/1 4/3/1
f 1/1/3 3/2/3 4/3/3
f 3/2/2 5/3/2 4/3/2
f 3//2 5//2 4//2
f 3 5 4
f 6/1/2 3/2/2 4/3/2
usemtl blue
f 1/1/2 3/2/2 4/3/2
f 4/3/2 1/1/2 3/2/2
"""


def _write_hostile(d):
    os.makedirs(d, exist_ok=True)
    open(os.path.join(d, "hostile.obj"), "w").write(HOSTILE_OBJ)
    open(os.path.join(d, "hostile.mtl"), "w").write("newmtl red\nKd 1 0 0\nnewmtl blue\nKd 0 0 1\n")
    cam = {"RES": [8, 8], "FOVY": 45.0, "ITERATIONS": 1, "DEPTH": 2, "FILE": "hostile", "EYE": [0.0, 0.0, 5.0], "LOOKAT": [0.0, 0.0, 0.0], "UP": [0.0, 1.0, 0.0]}
    path = os.path.join(d, "scene.json")
    # the same file three times: the second copy is moved, so its vertices are all new although the file is the same (vertices
    # are de-duplicated ACROSS objects by value); the third lies on the first and repeats its keys exactly (none new, NaN corners excepted)
    objs = [scenegen._obj_entry(os.path.join(d, "hostile.obj")) for _ in range(3)]
    objs[1]["TRANS"] = [3.0, 0.0, 0.0]
    scenegen.write_scene(path, cam, {"grey": scenegen._mat([0.5, 0.5, 0.5])}, objs)
    return path


def test_hostile_corners_match_host_path(tmp_path):
    path = _write_hostile(str(tmp_path / "h"))
    host = ptb.Scene(path)
    dev = ptb.Scene(path, device=0)
    _same_scene(host, dev, tmp_path)
    im = host.array("indices_matid")
    pos = host.array("pos")
    assert im.shape[0] == 33
    # triangles 0 and 1 differ only in the sign of zeros: same three vertices, the bits of the first occurrence
    assert list(im[0][:3]) == list(im[1][:3]) and not np.signbit(pos[im[0][0]][:3]).any()
    # triangles 2 and 3 carry NaN normals: six vertices of their own
    assert len(set(im[2][:3]) | set(im[3][:3])) == 6 and np.isnan(host.array("nor")[im[2][0]][:3]).all()
    # vn (0,0,2) is normalised BEFORE it becomes part of the key (src/scene.cpp:414-417): the same vertices as with vn (0,0,1)
    assert list(im[4][:3]) == list(im[0][:3])
    # the third object repeats the first: only its NaN corners are new
    assert list(im[22][:3]) == list(im[0][:3]) and set(im[24][:3]).isdisjoint(im[2][:3])
    # tangents of untextured triangles are zero vectors: NaN after normalisation, the host's negative default NaN
    tn = dev.array("tangent").view(np.uint32)
    assert (tn[im[7][0]][:3] == 0xFFC00000).all()
    host.free(); dev.free()


@pytest.mark.parametrize("gen", ["cornell", "sponza", "fireplace"])
def test_generated_scenes_match_host_path(tmp_path, gen):
    if gen == "cornell":
        path = scenegen.cornell(str(tmp_path / "c"), res=(64, 64))
    elif gen == "sponza":
        path = scenegen.sponza(str(tmp_path / "s"), res=(64, 36), scale=0.3, tex_size=32)[0]
    else:
        path = scenegen.fireplace(str(tmp_path / "f"), res=(64, 36))[0]          # BASELINE config 2's 143,173 triangles
    host = ptb.Scene(path)
    dev = ptb.Scene(path, device=0)
    assert host.info.num_triangles > 0 and dev.info.num_vertices < 3 * dev.info.num_triangles      # something was shared
    _same_scene(host, dev, tmp_path)
    host.free(); dev.free()


def test_two_million_triangles_match_host_and_reference_loader(tmp_path, monkeypatch):
    """Several thousand tiles of the scan and a table larger than L2; the reference's own loader (oracle/_ref/refdump --load-only)
    where it is present: triangle / vertex counts and the position-weighted checksum over positions, uvs, index triples and
    materials."""
    path, ntri = scenegen.stress_scene(str(tmp_path / "st"), num_triangles=2_000_000, res=(64, 36))
    monkeypatch.setenv("PTB_LOADER_VERBOSE", "1")
    dev = ptb.Scene(path, device=0)
    host = ptb.Scene(path)
    assert dev.info.num_triangles == ntri
    _same_scene(host, dev, tmp_path)
    if have_refdump():
        import importlib.util
        spec = importlib.util.spec_from_file_location("measure_ingest", os.path.join(ROOT, "tools", "measure_ingest.py"))
        mi = importlib.util.module_from_spec(spec); spec.loader.exec_module(mi)
        q = subprocess.run([REFDUMP, path, str(tmp_path / "ref"), "--load-only"], capture_output=True, text=True, timeout=600)
        line = [l for l in q.stdout.splitlines() if l.startswith("REFLOAD")]
        assert line, q.stdout + q.stderr
        r = json.loads(line[0][8:])
        assert (r["triangles"], r["vertices"]) == (dev.info.num_triangles, dev.info.num_vertices)
        assert r["checksum"] == mi.checksum(dev)
    host.free(); dev.free()


def test_environment_switch_and_errors(tmp_path, monkeypatch):
    path = scenegen.cornell(str(tmp_path / "c"), res=(32, 32))
    host = ptb.Scene(path)
    monkeypatch.setenv("PTB_LOADER_GPU", "0")
    monkeypatch.setenv("PTB_LOADER_VERBOSE", "1")
    via_env = ptb.Scene(path)                      # ptb_scene_load itself, routed by the environment (what the shim would use)
    _same_scene(host, via_env, tmp_path)
    monkeypatch.setenv("PTB_LOADER_GPU", "zero")
    with pytest.raises(capi.PtbError) as e:
        ptb.Scene(path)
    assert e.value.code == capi.ERR_INVALID
    monkeypatch.delenv("PTB_LOADER_GPU")
    with pytest.raises(capi.PtbError) as e:        # a device that does not exist: an error, not the host path
        ptb.Scene(path, device=capi.lib.ptb_device_count() + 7)
    assert e.value.code == capi.ERR_CUDA
    # the scene renders like the host-loaded one
    a = ptb.Renderer(host); b = ptb.Renderer(via_env)
    a.render(1, 2); b.render(1, 2)
    (da, ia), (db, ib) = a.accumulators(), b.accumulators()
    assert da.any() and _bits(da, db) and _bits(ia, ib)
    a.free(); b.free(); host.free(); via_env.free()
