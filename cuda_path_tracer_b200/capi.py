"""ctypes binding of libptb200.so (include/ptb200.h).

This is the only place Python touches the product: every call below goes through the C ABI.
There is no Python/NumPy fallback -- if the shared library is missing, importing fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PTB200_LIB") or os.path.join(_HERE, "libptb200.so")   # PTB200_LIB: A/B builds during development


class PtbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ptb200 error {code}: {msg}")
        self.code = code


OK, ERR_INVALID, ERR_IO, ERR_PARSE, ERR_CUDA, ERR_UNSUPPORTED, ERR_STATE = 0, -1, -2, -3, -4, -5, -6
IMG_SUM_RGB32F, IMG_MEAN_RGB32F, IMG_RGBA8 = 0, 1, 2
IMG_VIEW_NORMALS, IMG_VIEW_DEPTH, IMG_VIEW_ALBEDO, IMG_VIEW_MOTION, IMG_STEPS_I32 = 3, 4, 5, 6, 7
IMG_DENOISED_RGB32F, IMG_DENOISED_RGBA8 = 8, 9
GB_DEPTH, GB_NORMAL_MATID, GB_ALBEDO, GB_MOTION = 0, 1, 2, 3
SAVE_GAMMA, SAVE_FLIP_X = 1, 2
STAGE_RAYGEN, STAGE_TRACE, STAGE_LOGIC, STAGE_SHADE = 0, 1, 2, 3
Q_EXTENSION, Q_SHADOW, Q_PBR, Q_DIFFUSE, Q_REFLECTION, Q_REFRACTION = range(6)


class Material(C.Structure):  # ptb_material == reference `Material` (src/scene_structs.h:52-65)
    _fields_ = [("basecolor", C.c_float * 3), ("metallic", C.c_float), ("roughness", C.c_float),
                ("emittance", C.c_float), ("ior", C.c_float), ("type", C.c_int32),
                ("diffuse_tex_id", C.c_int32), ("normal_tex_id", C.c_int32),
                ("metallic_roughness_tex_id", C.c_int32), ("_pad", C.c_int32)]


MATERIAL_DTYPE = np.dtype([("basecolor", "<f4", 3), ("metallic", "<f4"), ("roughness", "<f4"), ("emittance", "<f4"),
                           ("ior", "<f4"), ("type", "<i4"), ("diffuse_tex_id", "<i4"), ("normal_tex_id", "<i4"),
                           ("metallic_roughness_tex_id", "<i4"), ("_pad", "<i4")])
assert MATERIAL_DTYPE.itemsize == 48 == C.sizeof(Material)


class Camera(C.Structure):  # ptb_camera == reference `Camera` (src/scene_structs.h:16-26)
    _fields_ = [("resolution", C.c_int32 * 2), ("position", C.c_float * 3), ("look_at", C.c_float * 3),
                ("view", C.c_float * 3), ("up", C.c_float * 3), ("right", C.c_float * 3),
                ("fov", C.c_float * 2), ("pixel_length", C.c_float * 2)]


class SceneInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("num_vertices", "num_triangles", "num_materials", "num_lights", "num_textures",
                                         "width", "height", "trace_depth", "iterations", "env_width", "env_height",
                                         "env_tex_id", "pdf_map_id")] + [("light_total_area", C.c_float)]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("trace_depth", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32),
                ("stripe_rows", C.c_int32), ("use_graph", C.c_int32), ("gbuffer", C.c_int32), ("denoiser", C.c_int32), ("env_nee", C.c_int32), ("megakernel", C.c_int32), ("reserved", C.c_int32 * 4)]


class TileInfo(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("width", "height", "owned_pixels", "rank", "world", "stripe_rows")]


class MultiConfig(C.Structure):  # ptb_multi_config
    _fields_ = [("n_devices", C.c_int32), ("devices", C.c_int32 * 16), ("tile_groups", C.c_int32), ("stripe_rows", C.c_int32),
                ("trace_depth", C.c_int32), ("use_graph", C.c_int32), ("megakernel", C.c_int32), ("env_nee", C.c_int32), ("reserved", C.c_int32 * 6)]


class MultiInfo(C.Structure):  # ptb_multi_info
    _fields_ = [(n, C.c_int32) for n in ("n_devices", "tile_groups", "sample_groups", "width", "height", "samples", "peer_stores", "reserved")]


GATHER_HANDLE_BYTES = 64
NOALLOC_MAX_RAYS = 262144


class Stats(C.Structure):
    _fields_ = [("extension_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("frames", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("lbvh_build_ms", C.c_float), ("lbvh_max_depth", C.c_int32),
                ("num_triangles", C.c_int32), ("last_batch_ms", C.c_float)]


# name -> (restype, argtypes); the list doubles as the symbol inventory the CPU test checks
P = C.c_void_p
SIGNATURES = {
    "ptb_abi_version": (C.c_int, []),
    "ptb_last_error": (C.c_char_p, []),
    "ptb_device_count": (C.c_int, []),
    "ptb_scene_load": (C.c_int, [C.c_char_p, C.POINTER(P)]),
    "ptb_scene_load_device": (C.c_int, [C.c_char_p, C.c_int32, C.POINTER(P)]),
    "ptb_scene_from_arrays": (C.c_int, [P, P, P, P, C.c_int32, P, P, C.c_int32, P, C.c_int32, C.POINTER(Camera), C.c_int32, C.POINTER(P)]),
    "ptb_scene_set_environment": (C.c_int, [P, P, C.c_int32, C.c_int32]),
    "ptb_scene_add_texture": (C.c_int, [P, P, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
    "ptb_scene_get_texture": (C.c_int64, [P, C.c_int32, P, C.c_size_t, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ptb_scene_set_environment_table": (C.c_int, [P, C.c_int32, C.c_int32, P, P, C.c_int32, C.c_int32]),
    "ptb_scene_set_camera": (C.c_int, [P, C.POINTER(C.c_int32 * 2), C.c_float, C.POINTER(C.c_float * 3), C.POINTER(C.c_float * 3), C.POINTER(C.c_float * 3), C.c_int32]),
    "ptb_scene_save_cache": (C.c_int, [P, C.c_char_p]),
    "ptb_scene_load_cache": (C.c_int, [C.c_char_p, C.POINTER(P)]),
    "ptb_scene_free": (None, [P]),
    "ptb_scene_get_info": (C.c_int, [P, C.POINTER(SceneInfo)]),
    "ptb_scene_get_camera": (C.c_int, [P, C.POINTER(Camera)]),
    "ptb_scene_get_array": (C.c_int64, [P, C.c_int, P, C.c_size_t]),
    "ptb_config_default": (None, [C.POINTER(Config)]),
    "ptb_init": (C.c_int, [P, C.POINTER(Config), C.POINTER(P)]),
    "ptb_free": (C.c_int, [P]),
    "ptb_render": (C.c_int, [P, C.c_int32]),
    "ptb_render_many": (C.c_int, [P, C.c_int32, C.c_int32]),
    "ptb_clear_accumulators": (C.c_int, [P]),
    "ptb_save_image": (C.c_int, [C.c_char_p, P, C.c_int32, C.c_int32, C.c_float, C.c_int32]),
    "ptb_get_gbuffer": (C.c_int, [P, C.c_int32, P, C.c_size_t]),
    "ptb_visualize_lbvh": (C.c_int, [P, P, C.c_int32, C.c_int32]),
    "ptb_sample_environment": (C.c_int, [P, P, C.c_int32, P, P, P]),
    "ptb_sync": (C.c_int, [P]),
    "ptb_set_stream": (C.c_int, [P, P]),
    "ptb_resolve": (C.c_int, [P, P, C.c_int32, C.c_int32, C.c_int32]),
    "ptb_get_accumulators": (C.c_int, [P, P, P]),
    "ptb_get_tile_info": (C.c_int, [P, C.POINTER(TileInfo)]),
    "ptb_get_owned_pixels": (C.c_int, [P, P, C.c_int32]),
    "ptb_get_lbvh": (C.c_int64, [P, C.c_int, P, C.c_size_t]),
    "ptb_build_lbvh": (C.c_int, [C.c_int32, P, C.c_int32, P, C.c_int32, P, P, P, P, P, P, P, P, P]),
    "ptb_trace_closest": (C.c_int, [P, P, P, C.c_int32, P, P, P, P]),
    "ptb_trace_occluded": (C.c_int, [P, P, P, C.c_int32, P]),
    "ptb_trace_closest_noalloc": (C.c_int, [P, P, P, C.c_int32, P, P, P, P]),
    "ptb_stage_run": (C.c_int, [P, C.c_int32, C.c_int32]),
    "ptb_get_queue_count": (C.c_int, [P, C.c_int32, C.POINTER(C.c_int32)]),
    "ptb_get_queue": (C.c_int, [P, C.c_int32, P, C.c_int32]),
    "ptb_get_shadow_queue": (C.c_int, [P, P, P, P, P, C.c_int32]),
    "ptb_get_path_field": (C.c_int, [P, C.c_int32, P, C.c_size_t]),
    "ptb_get_stats": (C.c_int, [P, C.POINTER(Stats)]),
    "ptb_stats_reset": (C.c_int, [P]),
    "ptb_time_raycast": (C.c_int, [P, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "ptb_time_kernel": (C.c_int, [P, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "ptb_multi_config_default": (None, [C.POINTER(MultiConfig)]),
    "ptb_multi_init": (C.c_int, [P, C.POINTER(MultiConfig), C.POINTER(P)]),
    "ptb_multi_free": (C.c_int, [P]),
    "ptb_multi_render": (C.c_int, [P, C.c_int32, C.c_int32]),
    "ptb_multi_resolve": (C.c_int, [P, P, C.c_int32, C.c_int32]),
    "ptb_multi_sync": (C.c_int, [P]),
    "ptb_multi_get_renderer": (C.c_int, [P, C.c_int32, C.POINTER(P)]),
    "ptb_multi_get_info": (C.c_int, [P, C.POINTER(MultiInfo)]),
    "ptb_gather_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(P)]),
    "ptb_gather_export": (C.c_int, [P, P, C.c_size_t]),
    "ptb_gather_open": (C.c_int, [C.c_int32, P, C.c_int32, C.c_int32, C.c_int32, C.POINTER(P)]),
    "ptb_gather_free": (C.c_int, [P]),
    "ptb_resolve_into": (C.c_int, [P, P, C.c_int32]),
    "ptb_gather_combine": (C.c_int, [P, P, C.c_int32, C.c_int32, C.c_int32, P]),
}

if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (or __graft_entry__.build()); "
                      "there is no fallback implementation")
lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def check(rc):
    if rc is not None and rc < 0:
        raise PtbError(rc, lib.ptb_last_error().decode("utf-8", "replace"))
    return rc


def ptr(a):
    """Pointer to a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.c_void_p)
