"""libyafaray_b200 — B200-native path-tracing hot path behind libYafaRay's integrator surface.

Only what the hot path needs lives here: `csrc/` (hand-written sm_100a CUDA kernels + the C ABI of include/yafgpu.h,
built in-tree into libyafgpu.so), `capi` (ctypes binding), `interface` (host mirror of yafaray4::Interface) and
`render` (host mirror of TiledIntegrator::render's pass loop).  Importing the package loads libyafgpu.so and fails
loudly when it is missing — there is no CPU fallback.
"""
import os as _os

from . import capi  # noqa: F401
if _os.environ.get("LIBYAFARAY_B200_HOST_ONLY") != "1":
    capi.lib.load()  # dlopens libyafgpu.so, raises ImportError when it is not built (host-only processes defer this to the first device call)
from .interface import Interface  # noqa: F401
from .render import Context, Group, render_scene, pin_host_arrays, unpin_host_arrays, load_film, film_to_image  # noqa: F401
from . import film_io  # noqa: F401

__all__ = ["capi", "Interface", "Context", "Group", "render_scene", "pin_host_arrays", "unpin_host_arrays", "load_film", "film_to_image", "film_io"]
