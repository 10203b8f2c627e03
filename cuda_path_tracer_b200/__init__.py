"""ptb200 -- B200-native wavefront path-tracing core (C-ABI library + thin ctypes host mirror)."""
from . import capi, scenegen, wavefront  # noqa: F401  (importing capi fails loudly if libptb200.so is missing)
from .wavefront import Gather, MultiRenderer, Pathtrace, PathtraceFree, PathtraceInit, Renderer, Scene, build_lbvh, save_image  # noqa: F401
