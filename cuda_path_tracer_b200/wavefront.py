"""Host-side mirror of the reference's render-path interface, over the C ABI.

Reference interface (C++):            here:
  Scene(filename)                      Scene(filename) / Scene.from_arrays(...)
  pathtrace_wavefront::PathtraceInit   PathtraceInit(scene) -> module-level renderer (one per process,
  pathtrace_wavefront::Pathtrace         like the reference's static pState) or Renderer(scene, ...) handles
  pathtrace_wavefront::PathtraceFree   PathtraceFree()
  BuildLBVH(LBVHData&, MeshData)       build_lbvh(pos4, indices_matid4)
(reference: src/pathtrace_wavefront.h:6-12, src/scene.h:51-92, src/bvh.h:7.)
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import check, lib, ptr

SA = dict(pos=(0, "<f4", 4), nor=(1, "<f4", 4), tangent=(2, "<f4", 4), uv=(3, "<f4", 2), indices_matid=(4, "<i4", 4),
          nor_geom=(5, "<f4", 4), materials=(6, capi.MATERIAL_DTYPE, None), light_tri_idx=(7, "<i4", None),
          light_cdf=(8, "<f4", None), env_probs=(9, "<f4", None), env_aliases=(10, "<i4", None), env_pdf=(11, "<f4", None))
LBVH = dict(aabb_min=(0, "<f4", 4), aabb_max=(1, "<f4", 4), centroid=(2, "<f4", 4), primitive_indices=(3, "<i4", None),
            morton_codes=(4, "<u8", None), child_nodes=(5, "<i4", 2), parent=(6, "<i4", None),
            escape_indices=(7, "<i4", None), world_aabb=(8, "<f4", 4))
PATH_FIELDS = dict(ray_ori=(0, "<f4", 4), ray_dir_dist=(1, "<f4", 4), hit_uv=(2, "<f4", 4), throughput_pdf=(3, "<f4", 4),
                   hit_geom_id=(4, "<i4", None), material_id=(5, "<i4", None), remaining_bounces=(6, "<i4", None),
                   rng_state=(7, "<u4", None))
QUEUES = dict(extension=0, shadow=1, pbr=2, diffuse=3, reflection=4, refraction=5)


def _fetch(getter, handle, spec):
    which, dtype, width = spec
    size = check(getter(handle, which, None, 0))
    dt = np.dtype(dtype)
    out = np.empty(size // dt.itemsize, dtype=dt)
    if size:
        check(getter(handle, which, ptr(out), size))
    return out.reshape(-1, width) if width else out


class Scene:
    """`class Scene` of the reference (src/scene.h:51-92): same scene.json schema and semantics."""

    def __init__(self, filename=None, _handle=None, device=None):
        """device: a GPU ordinal runs the vertex de-duplication and the repack into upload layout there
        (ptb_scene_load_device); None keeps them on the host threads.  Same scene, bit for bit."""
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
        elif device is not None:
            check(lib.ptb_scene_load_device(str(filename).encode(), int(device), C.byref(self._h)))
        else:
            check(lib.ptb_scene_load(str(filename).encode(), C.byref(self._h)))

    @classmethod
    def load_cache(cls, filename):
        """A scene written by save_cache(): no parsing, no de-duplication, bit-identical arrays."""
        h = C.c_void_p()
        check(lib.ptb_scene_load_cache(str(filename).encode(), C.byref(h)))
        return cls(_handle=h)

    def save_cache(self, filename):
        check(lib.ptb_scene_save_cache(self._h, str(filename).encode()))

    @classmethod
    def from_arrays(cls, pos, nor, indices, material_ids, materials, camera, uv=None, tangent=None, trace_depth=8):
        """camera: dict(resolution=(w,h), fovy=deg, eye=, look_at=, up=)."""
        pos = np.ascontiguousarray(pos, dtype=np.float32).reshape(-1, 3)
        nor = np.ascontiguousarray(nor, dtype=np.float32).reshape(-1, 3)
        indices = np.ascontiguousarray(indices, dtype=np.int32).reshape(-1, 3)
        material_ids = np.ascontiguousarray(material_ids, dtype=np.int32).reshape(-1)
        materials = np.ascontiguousarray(materials, dtype=capi.MATERIAL_DTYPE)
        uv = None if uv is None else np.ascontiguousarray(uv, dtype=np.float32).reshape(-1, 2)
        tangent = None if tangent is None else np.ascontiguousarray(tangent, dtype=np.float32).reshape(-1, 3)
        cam = capi.Camera()
        cam.resolution[0], cam.resolution[1] = camera["resolution"]
        h = C.c_void_p()
        check(lib.ptb_scene_from_arrays(ptr(pos), ptr(nor), ptr(uv), ptr(tangent), pos.shape[0], ptr(indices),
                                        ptr(material_ids), indices.shape[0], ptr(materials), materials.shape[0],
                                        C.byref(cam), trace_depth, C.byref(h)))
        s = cls(_handle=h)
        s.set_camera(camera["resolution"], camera["fovy"], camera["eye"], camera["look_at"], camera["up"], trace_depth)
        return s

    def set_camera(self, resolution, fovy, eye, look_at, up, trace_depth=0):
        res = (C.c_int32 * 2)(*resolution)
        e, l, u = (C.c_float * 3)(*eye), (C.c_float * 3)(*look_at), (C.c_float * 3)(*up)
        check(lib.ptb_scene_set_camera(self._h, C.byref(res), float(fovy), C.byref(e), C.byref(l), C.byref(u), trace_depth))

    def set_environment(self, rgb):
        rgb = np.ascontiguousarray(rgb, dtype=np.float32)
        h, w, _ = rgb.shape
        check(lib.ptb_scene_set_environment(self._h, ptr(rgb), w, h))

    @property
    def info(self):
        i = capi.SceneInfo()
        check(lib.ptb_scene_get_info(self._h, C.byref(i)))
        return i

    @property
    def camera(self):
        c = capi.Camera()
        check(lib.ptb_scene_get_camera(self._h, C.byref(c)))
        return c

    def array(self, name):
        return _fetch(lib.ptb_scene_get_array, self._h, SA[name])

    def texture(self, tex_id):
        """Decoded texture `tex_id` as it is uploaded: uint8 (h, w, 4), or float32 (h, w) / (h, w, 4) for the environment maps."""
        w, h, fmt = C.c_int32(), C.c_int32(), C.c_int32()
        size = check(lib.ptb_scene_get_texture(self._h, tex_id, None, 0, C.byref(w), C.byref(h), C.byref(fmt)))
        out = np.empty(size, np.uint8)
        check(lib.ptb_scene_get_texture(self._h, tex_id, ptr(out), size, None, None, None))
        if fmt.value == 0:
            return out.reshape(h.value, w.value, 4)
        f = out.view(np.float32)
        return f.reshape(h.value, w.value) if fmt.value == 1 else f.reshape(h.value, w.value, 4)

    def free(self):
        if self._h:
            lib.ptb_scene_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Renderer:
    """An opaque ptb_renderer: PathtraceInit / Pathtrace / PathtraceFree on a handle."""

    def __init__(self, scene, device=0, trace_depth=0, rank=0, world=1, stripe_rows=8, use_graph=True, gbuffer=False, denoiser=False, env_nee=False, megakernel=False):
        cfg = capi.Config()
        lib.ptb_config_default(C.byref(cfg))
        cfg.device, cfg.trace_depth, cfg.rank, cfg.world, cfg.stripe_rows = device, trace_depth, rank, world, stripe_rows
        cfg.use_graph = 1 if use_graph else 0
        cfg.gbuffer = 1 if gbuffer else 0
        cfg.denoiser = 1 if denoiser else 0
        cfg.env_nee = 1 if env_nee else 0
        cfg.megakernel = 1 if megakernel else 0   # pathtrace_megakernel::Pathtrace (src/pathtrace.cu:645-706)
        self._h = C.c_void_p()
        check(lib.ptb_init(scene._h, C.byref(cfg), C.byref(self._h)))
        t = capi.TileInfo()
        check(lib.ptb_get_tile_info(self._h, C.byref(t)))
        self.width, self.height, self.owned_pixels = t.width, t.height, t.owned_pixels
        self.rank, self.world, self.stripe_rows = t.rank, t.world, t.stripe_rows

    # -- Pathtrace(pbo, frame, iter) --------------------------------------------------------------
    def render(self, iteration, count=1):
        check(lib.ptb_render_many(self._h, iteration, count))

    def sample_environment(self, rnd):
        """sampleEnvironmentMap for rnd[n, 2] -> (dir[n, 3], pdf[n], texel[n])."""
        rnd = np.ascontiguousarray(rnd, dtype=np.float32).reshape(-1, 2)
        n = rnd.shape[0]
        d, p, t = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        check(lib.ptb_sample_environment(self._h, ptr(rnd), n, ptr(d), ptr(p), ptr(t)))
        return d, p, t

    def clear(self):
        """Zero the accumulators (what iteration 1 does implicitly); for sample-index partitions."""
        check(lib.ptb_clear_accumulators(self._h))

    def sync(self):
        check(lib.ptb_sync(self._h))

    def resolve(self, iteration, fmt=capi.IMG_MEAN_RGB32F, out=None):
        """(direct + indirect) / iteration for the owned pixels, into a host array."""
        if fmt in (capi.IMG_RGBA8, capi.IMG_DENOISED_RGBA8) or capi.IMG_VIEW_NORMALS <= fmt <= capi.IMG_VIEW_MOTION:
            out = np.empty((self.owned_pixels, 4), dtype=np.uint8) if out is None else out
        else:
            out = np.empty((self.owned_pixels, 3), dtype=np.float32) if out is None else out
        check(lib.ptb_resolve(self._h, ptr(out), fmt, iteration, 0))
        return out

    def gbuffer(self, which):
        """First-bounce G-buffer of the last frame: 'depth' (P,), 'normal_matid' (P,4), 'albedo' (P,4), 'motion' (P,2)."""
        idx, shape = {"depth": (capi.GB_DEPTH, (self.owned_pixels,)), "normal_matid": (capi.GB_NORMAL_MATID, (self.owned_pixels, 4)),
                      "albedo": (capi.GB_ALBEDO, (self.owned_pixels, 4)), "motion": (capi.GB_MOTION, (self.owned_pixels, 2))}[which]
        out = np.empty(shape, dtype=np.float32)
        check(lib.ptb_get_gbuffer(self._h, idx, ptr(out), out.nbytes))
        return out

    def visualize_lbvh(self, steps=False):
        """LBVH traversal-cost heat map (VisualizeLBVH): float32 (P,3) colours, or int32 (P,) node visits."""
        out = np.empty((self.owned_pixels,), dtype=np.int32) if steps else np.empty((self.owned_pixels, 3), dtype=np.float32)
        check(lib.ptb_visualize_lbvh(self._h, ptr(out), capi.IMG_STEPS_I32 if steps else capi.IMG_SUM_RGB32F, 0))
        return out

    def resolve_device(self, device_ptr, iteration, fmt=capi.IMG_MEAN_RGB32F):
        check(lib.ptb_resolve(self._h, C.c_void_p(device_ptr), fmt, iteration, 1))

    def set_stream(self, cuda_stream):
        check(lib.ptb_set_stream(self._h, C.c_void_p(cuda_stream)))

    def accumulators(self):
        d = np.empty((self.owned_pixels, 3), dtype=np.float32)
        i = np.empty((self.owned_pixels, 3), dtype=np.float32)
        check(lib.ptb_get_accumulators(self._h, ptr(d), ptr(i)))
        return d, i

    def owned_pixel_indices(self):
        out = np.empty(self.owned_pixels, dtype=np.int32)
        check(lib.ptb_get_owned_pixels(self._h, ptr(out), out.size))
        return out

    # -- parity views -----------------------------------------------------------------------------
    def lbvh(self, name):
        return _fetch(lib.ptb_get_lbvh, self._h, LBVH[name])

    def stage(self, stage, iteration=0):
        check(lib.ptb_stage_run(self._h, stage, iteration))

    def queue_count(self, name):
        n = C.c_int32()
        check(lib.ptb_get_queue_count(self._h, QUEUES[name], C.byref(n)))
        return n.value

    def queue(self, name):
        n = self.queue_count(name)
        out = np.empty(max(n, 1), dtype=np.int32)
        check(lib.ptb_get_queue(self._h, QUEUES[name], ptr(out), n))
        return out[:n]

    def shadow_queue(self):
        n = self.queue_count("shadow")
        m = max(n, 1)
        o, d, r = (np.empty((m, 4), dtype=np.float32) for _ in range(3))
        p = np.empty(m, dtype=np.int32)
        check(lib.ptb_get_shadow_queue(self._h, ptr(o), ptr(d), ptr(r), ptr(p), n))
        return o[:n], d[:n], r[:n], p[:n]

    def path_field(self, name):
        which, dtype, width = PATH_FIELDS[name]
        n = self.width * self.height
        out = np.zeros((n, width) if width else n, dtype=dtype)
        check(lib.ptb_get_path_field(self._h, which, ptr(out), out.nbytes))
        return out

    # -- batched RayCast --------------------------------------------------------------------------
    def trace_closest(self, ori4, dir4):
        ori4 = np.ascontiguousarray(ori4, dtype=np.float32).reshape(-1, 4)
        dir4 = np.ascontiguousarray(dir4, dtype=np.float32).reshape(-1, 4)
        n = ori4.shape[0]
        hit = np.empty(n, dtype=np.int32)
        t, u, v = (np.empty(n, dtype=np.float32) for _ in range(3))
        check(lib.ptb_trace_closest(self._h, ptr(ori4), ptr(dir4), n, ptr(hit), ptr(t), ptr(u), ptr(v)))
        return hit, t, u, v

    def trace_closest_noalloc(self, ori4, dir4):
        """trace_closest through ptb_trace_closest_noalloc: static device scratch, at most capi.NOALLOC_MAX_RAYS rays."""
        ori4 = np.ascontiguousarray(ori4, dtype=np.float32).reshape(-1, 4)
        dir4 = np.ascontiguousarray(dir4, dtype=np.float32).reshape(-1, 4)
        n = ori4.shape[0]
        hit = np.empty(n, dtype=np.int32)
        t, u, v = (np.empty(n, dtype=np.float32) for _ in range(3))
        check(lib.ptb_trace_closest_noalloc(self._h, ptr(ori4), ptr(dir4), n, ptr(hit), ptr(t), ptr(u), ptr(v)))
        return hit, t, u, v

    def trace_occluded(self, ori_tmax4, dir4):
        ori_tmax4 = np.ascontiguousarray(ori_tmax4, dtype=np.float32).reshape(-1, 4)
        dir4 = np.ascontiguousarray(dir4, dtype=np.float32).reshape(-1, 4)
        n = ori_tmax4.shape[0]
        occ = np.empty(n, dtype=np.int32)
        check(lib.ptb_trace_occluded(self._h, ptr(ori_tmax4), ptr(dir4), n, ptr(occ)))
        return occ

    def time_raycast(self, count):
        ms, rays = C.c_float(), C.c_int32()
        check(lib.ptb_time_raycast(self._h, count, C.byref(ms), C.byref(rays)))
        return ms.value, rays.value

    def time_kernel(self, which, count, cold=False):
        """(avg ms, rays) of `count` launches of the extension ('ext') or shadow ('shadow') RayCast kernel of the current bounce."""
        ms, rays = C.c_float(), C.c_int32()
        check(lib.ptb_time_kernel(self._h, {"ext": 0, "shadow": 1}[which], count, int(cold), C.byref(ms), C.byref(rays)))
        return ms.value, rays.value

    def stats(self):
        s = capi.Stats()
        check(lib.ptb_get_stats(self._h, C.byref(s)))
        return s

    def stats_reset(self):
        check(lib.ptb_stats_reset(self._h))

    def free(self):
        if self._h:
            check(lib.ptb_free(self._h))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class MultiRenderer:
    """ptb_multi: one frame over several GPUs of ONE process (include/ptb200.h, multi-GPU).  `devices[0]` is the root; the
    resolve kernel of every GPU stores its pixels into the root's frame buffer over NVLink.  tile_groups G cuts the frame into
    G interleaved stripe sets, the remaining len(devices) / G factor splits the sample indices."""

    def __init__(self, scene, devices, tile_groups=0, stripe_rows=8, trace_depth=0, use_graph=True, megakernel=False, env_nee=False):
        cfg = capi.MultiConfig()
        lib.ptb_multi_config_default(C.byref(cfg))
        cfg.n_devices = len(devices)
        for k, d in enumerate(devices):
            cfg.devices[k] = d
        cfg.tile_groups, cfg.stripe_rows, cfg.trace_depth = tile_groups, stripe_rows, trace_depth
        cfg.use_graph, cfg.megakernel, cfg.env_nee = int(use_graph), int(megakernel), int(env_nee)
        self._h = C.c_void_p()
        check(lib.ptb_multi_init(scene._h, C.byref(cfg), C.byref(self._h)))
        i = self.info
        self.width, self.height, self.tile_groups, self.sample_groups = i.width, i.height, i.tile_groups, i.sample_groups

    @property
    def info(self):
        i = capi.MultiInfo()
        check(lib.ptb_multi_get_info(self._h, C.byref(i)))
        return i

    def render(self, first_step, count=1):
        check(lib.ptb_multi_render(self._h, first_step, count))

    def resolve(self, fmt=capi.IMG_MEAN_RGB32F, out=None):
        n = self.width * self.height
        if out is None:
            out = np.empty((n, 4), np.uint8) if fmt == capi.IMG_RGBA8 else np.empty((n, 3), np.float32)
        check(lib.ptb_multi_resolve(self._h, ptr(out), fmt, 0))
        return out

    def resolve_device(self, device_ptr, fmt=capi.IMG_MEAN_RGB32F):
        check(lib.ptb_multi_resolve(self._h, C.c_void_p(device_ptr), fmt, 1))

    def sync(self):
        check(lib.ptb_multi_sync(self._h))

    def stats(self, index):
        h = C.c_void_p()
        check(lib.ptb_multi_get_renderer(self._h, index, C.byref(h)))
        s = capi.Stats()
        check(lib.ptb_get_stats(h, C.byref(s)))
        return s

    def free(self):
        if self._h:
            check(lib.ptb_multi_free(self._h))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Gather:
    """ptb_gather: the root's full-frame buffer other ranks' resolve kernels store into (one process per GPU: export() on the
    root, Gather.open(handle) elsewhere)."""

    def __init__(self, device, width, height, slices=1, _handle=None, _remote=False):
        self.device, self.width, self.height, self.slices, self.remote = device, width, height, slices, _remote
        self._h = C.c_void_p()
        if _handle is not None:
            self._h = _handle
        else:
            check(lib.ptb_gather_create(device, width, height, slices, C.byref(self._h)))

    def export(self):
        buf = C.create_string_buffer(capi.GATHER_HANDLE_BYTES)
        check(lib.ptb_gather_export(self._h, buf, capi.GATHER_HANDLE_BYTES))
        return buf.raw

    @classmethod
    def open(cls, device, handle, width, height, slices=1):
        h = C.c_void_p()
        buf = C.create_string_buffer(bytes(handle), capi.GATHER_HANDLE_BYTES)
        check(lib.ptb_gather_open(device, buf, width, height, slices, C.byref(h)))
        return cls(device, width, height, slices, _handle=h, _remote=True)

    def resolve_from(self, renderer, slice_index=0):
        check(lib.ptb_resolve_into(renderer._h, self._h, slice_index))

    def combine(self, samples, fmt=capi.IMG_MEAN_RGB32F, out=None, stream=0):
        n = self.width * self.height
        if out is None:
            out = np.empty((n, 4), np.uint8) if fmt == capi.IMG_RGBA8 else np.empty((n, 3), np.float32)
        check(lib.ptb_gather_combine(self._h, ptr(out), fmt, samples, 0, C.c_void_p(stream)))
        return out

    def combine_device(self, device_ptr, samples, fmt=capi.IMG_MEAN_RGB32F, stream=0):
        check(lib.ptb_gather_combine(self._h, C.c_void_p(device_ptr), fmt, samples, 1, C.c_void_p(stream)))

    def free(self):
        if self._h:
            check(lib.ptb_gather_free(self._h))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def save_image(path, rgb, width, height, scale=1.0, gamma=False, flip_x=False):
    """Image::savePNG / saveHDR (+ main.cpp's saveImage transform): rgb float32 [height*width, 3]; .png or .hdr."""
    rgb = np.ascontiguousarray(rgb, dtype=np.float32).reshape(height * width, 3)
    flags = (capi.SAVE_GAMMA if gamma else 0) | (capi.SAVE_FLIP_X if flip_x else 0)
    check(lib.ptb_save_image(path.encode(), ptr(rgb), width, height, float(scale), flags))


def build_lbvh(pos4, indices_matid4, device=0):
    """BuildLBVH (src/bvh.h:7) on host arrays; returns the reference-layout arrays and the build time."""
    pos4 = np.ascontiguousarray(pos4, dtype=np.float32).reshape(-1, 4)
    im = np.ascontiguousarray(indices_matid4, dtype=np.int32).reshape(-1, 4)
    V, N = pos4.shape[0], im.shape[0]
    out = dict(morton_codes=np.empty(max(N, 1), np.uint64), primitive_indices=np.empty(2 * max(N, 1), np.int32),
               child_nodes=np.empty((max(N - 1, 1), 2), np.int32), parent=np.empty(2 * max(N, 1), np.int32),
               escape_indices=np.empty(2 * max(N, 1), np.int32), aabb_min=np.empty((2 * max(N, 1), 4), np.float32),
               aabb_max=np.empty((2 * max(N, 1), 4), np.float32), world_aabb=np.empty((2, 4), np.float32))
    ms = C.c_float()
    check(lib.ptb_build_lbvh(device, ptr(pos4), V, ptr(im), N, ptr(out["morton_codes"]), ptr(out["primitive_indices"]),
                             ptr(out["child_nodes"]), ptr(out["parent"]), ptr(out["escape_indices"]), ptr(out["aabb_min"]),
                             ptr(out["aabb_max"]), ptr(out["world_aabb"]), C.cast(C.byref(ms), C.c_void_p)))
    out["build_ms"] = ms.value
    return out


# ---- the reference's process-global front-end (src/pathtrace_wavefront.cu:118-181,205) ----------
_state = {"renderer": None, "scene": None}


def PathtraceInit(scene, **kw):
    if _state["renderer"] is None:
        _state["renderer"] = Renderer(scene, **kw)
        _state["scene"] = scene
    return _state["renderer"]


def Pathtrace(pbo=None, frame=0, iteration=1):
    """One sample per pixel; `pbo`, if given, is a host uint8 (P,4) array that receives the tonemapped frame."""
    r = _state["renderer"]
    if r is None:
        print("[Wavefront] Error: Render called before Init!")  # src/pathtrace_wavefront.cu:207-210
        return
    r.render(iteration)
    if pbo is not None:
        r.resolve(iteration, capi.IMG_RGBA8, out=pbo)


def PathtraceFree():
    if _state["renderer"] is not None:
        _state["renderer"].free()
    _state["renderer"] = None
    _state["scene"] = None
