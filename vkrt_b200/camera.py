"""Camera: ctypes view of the host C++ `Camera` class (vkrt_b200/host/Camera.cpp), which restates the reference's
src/Camera.cpp.  Tests and bench.py drive the same host code as the headless C++ application."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .binding import Uniforms

HOST_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "host", "libvkrt_host.so")

# src/Raytracer.cpp:26-31
LIGHT_POSITIONS = np.array([[6, 6, 0, 0], [2, 5, 0, 0], [-2, 4, 0, 0], [-6, 3, 0, 0]], dtype=np.float32)

_host = None


def load_host():
    global _host
    if _host is None:
        if not os.path.exists(HOST_LIB_PATH):
            raise FileNotFoundError(f"{HOST_LIB_PATH} not found: run __graft_entry__.build()")
        lib = C.CDLL(HOST_LIB_PATH)
        f, vp = C.c_float, C.c_void_p
        lib.vkrt_host_camera_create.restype = vp
        lib.vkrt_host_camera_create.argtypes = [f]
        lib.vkrt_host_camera_destroy.argtypes = [vp]
        lib.vkrt_host_camera_set_position.argtypes = [vp, f, f, f]
        lib.vkrt_host_camera_set_rotation.argtypes = [vp, f, f, f]
        lib.vkrt_host_camera_translate.argtypes = [vp, f, f, f]
        lib.vkrt_host_camera_rotate.argtypes = [vp, f, f, f, f]
        lib.vkrt_host_camera_set_aspect.argtypes = [vp, f]
        lib.vkrt_host_camera_axes.argtypes = [vp, vp]
        lib.vkrt_host_camera_matrices.argtypes = [vp, vp]
        lib.vkrt_host_camera_wvp.argtypes = [vp, f, vp]
        lib.vkrt_host_fill_uniforms.argtypes = [vp, vp, C.POINTER(Uniforms)]
        _host = lib
    return _host


class Camera:
    """Same methods as the reference's Camera (src/Camera.hpp:6-33); aspect ratio is a runtime argument."""

    def __init__(self, aspect: float = 1600.0 / 1200.0):
        self._lib = load_host()
        self._h = self._lib.vkrt_host_camera_create(np.float32(aspect))

    def __del__(self):
        try:
            if self._h:
                self._lib.vkrt_host_camera_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def setPosition(self, pos):
        self._lib.vkrt_host_camera_set_position(self._h, *[np.float32(v) for v in pos])

    def setRotation(self, rot):
        self._lib.vkrt_host_camera_set_rotation(self._h, *[np.float32(v) for v in rot])

    def translate(self, t):
        self._lib.vkrt_host_camera_translate(self._h, *[np.float32(v) for v in t])

    def rotate(self, axis, amount):
        self._lib.vkrt_host_camera_rotate(self._h, *[np.float32(v) for v in axis], np.float32(amount))

    def setAspectRatio(self, aspect):
        self._lib.vkrt_host_camera_set_aspect(self._h, np.float32(aspect))

    def _axes(self):
        out = np.zeros(12, dtype=np.float32)
        self._lib.vkrt_host_camera_axes(self._h, out.ctypes.data)
        return out.reshape(4, 3)

    def getPosition(self):
        return self._axes()[0]

    def getForward(self):
        return self._axes()[1]

    def getLeft(self):
        return self._axes()[2]

    def getUp(self):
        return self._axes()[3]

    def _matrices(self):
        out = np.zeros(32, dtype=np.float32)
        self._lib.vkrt_host_camera_matrices(self._h, out.ctypes.data)
        return out

    def getViewMatrix(self):
        """Column-major 16 floats, as glm stores it."""
        return self._matrices()[:16].copy()

    def getProjectionMatrix(self):
        return self._matrices()[16:].copy()

    def wvp(self, scale: float = 0.01) -> np.ndarray:
        """proj * view * scale(s), column-major 16 floats: the rasteriser's uniform (src/Rasterizer.cpp:189-191)."""
        out = np.zeros(16, dtype=np.float32)
        self._lib.vkrt_host_camera_wvp(self._h, np.float32(scale), out.ctypes.data)
        return out

    def uniforms(self, lights: np.ndarray = LIGHT_POSITIONS) -> Uniforms:
        """The 256-byte block Raytracer::update fills (src/Raytracer.cpp:224-235)."""
        u = Uniforms()
        l = np.ascontiguousarray(lights, dtype=np.float32).reshape(16)
        self._lib.vkrt_host_fill_uniforms(self._h, l.ctypes.data, C.byref(u))
        return u


def make_uniforms(width: int, height: int, position=(6.3, 4.5, -0.7), rotation=(0.0, 1.57, 0.0), lights=LIGHT_POSITIONS) -> Uniforms:
    """Uniforms for the reference's default ray-tracing pose (src/Raytracer.cpp:267-271) at a given frame size."""
    cam = Camera(np.float32(width) / np.float32(height))
    cam.setPosition(position)
    cam.setRotation(rotation)
    return cam.uniforms(lights)


def uniforms_to_array(u: Uniforms) -> np.ndarray:
    return np.frombuffer(bytes(u), dtype=np.float32).copy()
