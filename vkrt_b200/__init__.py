"""vkrt_b200 — B200-native (sm_100a CUDA) ray-tracing hot path of jparimaa/vkrt.

The product is the C-ABI shared library ``vkrt_b200/csrc/libvkrt_b200.so`` (see ``include/vkrt.h``); the C++
host classes that mirror the reference's ``Model`` / ``Camera`` / ``Raytracer`` live in ``vkrt_b200/host``.
This Python package is a thin ctypes binding used by the tests and by ``bench.py``; it has no compute path of
its own and raises if the CUDA library is missing.
"""
from .binding import (  # noqa: F401
    LIB_PATH,
    Blas,
    Context,
    Frame,
    Tlas,
    VkrtError,
    load_library,
    AOV_BGRA8,
    AOV_RGB32F,
    AOV_TRIANGLE_ID,
    AOV_SHADOW_MASK,
    BUILD_DEFAULT,
    BUILD_PREFER_FAST_BUILD,
    TRACE_DEFAULT,
    TRACE_NO_AOV,
)
from .camera import Camera, make_uniforms, LIGHT_POSITIONS  # noqa: F401
from .raytracer import Raytracer  # noqa: F401
