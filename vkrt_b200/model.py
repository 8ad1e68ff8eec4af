"""Model: ctypes view of the host C++ glTF loader (vkrt_b200/host/Model.cpp), the drop-in counterpart of the
reference's `Model` (src/Model.hpp:8-54).  Returns the same dictionary layout as vkrt_b200.scenes."""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict

import numpy as np

from .camera import load_host


def fnv1a(array: np.ndarray) -> str:
    """FNV-1a 64 of an array's bytes (computed by the host library), as printed by oracle/ref_driver/ref_model_dump."""
    lib = _lib()
    a = np.ascontiguousarray(array)
    return f"{lib.vkrt_host_fnv1a(a.ctypes.data, a.nbytes):016x}"


def _lib():
    lib = load_host()
    if not getattr(lib, "_model_bound", False):
        vp, u32 = C.c_void_p, C.c_uint32
        lib.vkrt_host_model_create.restype = vp
        lib.vkrt_host_model_create.argtypes = [C.c_char_p, C.c_char_p]
        lib.vkrt_host_model_destroy.argtypes = [vp]
        lib.vkrt_host_model_submesh_count.restype = u32
        lib.vkrt_host_model_submesh_count.argtypes = [vp]
        lib.vkrt_host_model_submesh_info.argtypes = [vp, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(C.c_int)]
        lib.vkrt_host_model_submesh_copy.argtypes = [vp, u32, vp, vp]
        lib.vkrt_host_model_material_count.restype = u32
        lib.vkrt_host_model_material_count.argtypes = [vp]
        lib.vkrt_host_model_material.argtypes = [vp, u32, vp]
        lib.vkrt_host_model_image_count.restype = u32
        lib.vkrt_host_model_image_count.argtypes = [vp]
        lib.vkrt_host_model_image_info.argtypes = [vp, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]
        lib.vkrt_host_model_image_copy.argtypes = [vp, u32, vp]
        lib.vkrt_host_decode_image.restype = C.c_int
        lib.vkrt_host_decode_image.argtypes = [vp, C.c_uint64, C.POINTER(u32), C.POINTER(u32), vp, C.c_uint64]
        lib.vkrt_host_fnv1a.restype = C.c_uint64
        lib.vkrt_host_fnv1a.argtypes = [vp, C.c_uint64]
        lib._model_bound = True
    return lib


def load_model(models_folder: str, filename: str = "sponza/Sponza.gltf") -> Dict:
    """Loads a glTF through the host C++ `Model` (aborts the process on a missing file, like the reference)."""
    lib = _lib()
    h = lib.vkrt_host_model_create(os.path.join(models_folder, "").encode(), filename.encode())
    try:
        subs = []
        for i in range(lib.vkrt_host_model_submesh_count(h)):
            nv, ni, mat = C.c_uint32(), C.c_uint32(), C.c_int()
            lib.vkrt_host_model_submesh_info(h, i, C.byref(nv), C.byref(ni), C.byref(mat))
            v = np.zeros((nv.value, 16), dtype=np.float32)
            idx = np.zeros(ni.value, dtype=np.uint32)
            lib.vkrt_host_model_submesh_copy(h, i, v.ctypes.data, idx.ctypes.data)
            subs.append({"vertices": v, "indices": idx, "material": mat.value})
        mats = []
        for i in range(lib.vkrt_host_model_material_count(h)):
            m = np.zeros(3, dtype=np.int32)
            lib.vkrt_host_model_material(h, i, m.ctypes.data)
            mats.append(tuple(int(x) for x in m))
        images = []
        for i in range(lib.vkrt_host_model_image_count(h)):
            w, hh, c, b = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
            lib.vkrt_host_model_image_info(h, i, C.byref(w), C.byref(hh), C.byref(c), C.byref(b))
            im = np.zeros((hh.value, w.value, c.value), dtype=np.uint8)
            lib.vkrt_host_model_image_copy(h, i, im.ctypes.data)
            images.append(im)
    finally:
        lib.vkrt_host_model_destroy(h)
    return {"submeshes": subs, "materials": mats, "images": images}


def decode_image(data: bytes) -> np.ndarray:
    lib = _lib()
    w, h = C.c_uint32(), C.c_uint32()
    buf = np.frombuffer(data, dtype=np.uint8)
    if lib.vkrt_host_decode_image(buf.ctypes.data, len(data), C.byref(w), C.byref(h), None, 0) != 0:
        raise ValueError("unsupported or corrupt image")
    out = np.zeros((h.value, w.value, 4), dtype=np.uint8)
    lib.vkrt_host_decode_image(buf.ctypes.data, len(data), C.byref(w), C.byref(h), out.ctypes.data, out.nbytes)
    return out
