"""ctypes binding of include/vkrt.h.  No fallbacks: a missing library or a missing CUDA device raises."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

# VKRT_B200_LIB: developer override to time experimental builds of the same library; there is still no CPU fallback
LIB_PATH = os.environ.get("VKRT_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libvkrt_b200.so")

AOV_BGRA8, AOV_RGB32F, AOV_TRIANGLE_ID, AOV_SHADOW_MASK = 0, 1, 2, 3
BUILD_DEFAULT, BUILD_PREFER_FAST_BUILD, BUILD_INPUT_ON_DEVICE = 0, 1, 2
BUILD_PREFER_FAST_TRACE = 8  # reinsertion passes after PLOC: slower build, fewer node visits per ray
TRACE_DEFAULT, TRACE_NO_AOV = 0, 1
TEXTURE_LOD_ZERO, TEXTURE_LOD_RAY_CONES = 0, 1
TILE = 64
UNIQUE_ID_BYTES = 128


class VkrtError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"vkrt error {code}: {text}")
        self.code = code


class GeometryRange(C.Structure):
    _fields_ = [("primitiveCount", C.c_uint32), ("maxVertex", C.c_uint32), ("indexByteOffset", C.c_uint64)]


class BuildSizes(C.Structure):
    _fields_ = [("accelerationStructureSize", C.c_uint64), ("buildScratchSize", C.c_uint64), ("geometryCopySize", C.c_uint64)]


class BlasInfo(C.Structure):
    _fields_ = [
        ("triangleCount", C.c_uint64),
        ("nodeCount", C.c_uint64),
        ("nodeBytes", C.c_uint64),
        ("triangleBytes", C.c_uint64),
        ("boundsMin", C.c_float * 3),
        ("boundsMax", C.c_float * 3),
        ("sahCost", C.c_float),
        ("buildMilliseconds", C.c_float),
    ]


class TlasInfo(C.Structure):
    _fields_ = [
        ("instanceCount", C.c_uint32),
        ("nodeCount", C.c_uint32),
        ("levels", C.c_uint32),
        ("reserved", C.c_uint32),
        ("boundsMin", C.c_float * 3),
        ("boundsMax", C.c_float * 3),
        ("sahCost", C.c_float),
        ("buildMilliseconds", C.c_float),
    ]


class Instance(C.Structure):
    _fields_ = [
        ("transform", C.c_float * 12),
        ("instanceCustomIndex", C.c_uint32, 24),
        ("mask", C.c_uint32, 8),
        ("sbtOffset", C.c_uint32, 24),
        ("flags", C.c_uint32, 8),
        ("accelerationStructureReference", C.c_uint64),
    ]


class SubmeshInfo(C.Structure):
    _fields_ = [
        ("baseColorTextureIndex", C.c_int32),
        ("metallicRoughnessTextureIndex", C.c_int32),
        ("normalTextureIndex", C.c_int32),
        ("indexBufferOffset", C.c_int32),
    ]


class Image(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("rgba8", C.c_void_p)]


class Uniforms(C.Structure):
    _fields_ = [
        ("viewInverse", C.c_float * 16),
        ("projInverse", C.c_float * 16),
        ("position", C.c_float * 4),
        ("right", C.c_float * 4),
        ("up", C.c_float * 4),
        ("forward", C.c_float * 4),
        ("lightPositions", (C.c_float * 4) * 4),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("raysPrimary", C.c_uint64),
        ("raysShadow", C.c_uint64),
        ("raysReflection", C.c_uint64),
        ("nodesVisited", C.c_uint64),
        ("trianglesTested", C.c_uint64),
        ("nodesVisitedAny", C.c_uint64),
        ("trianglesTestedAny", C.c_uint64),
        ("kernelLaunches", C.c_uint64),
        ("msRaygen", C.c_float),
        ("msTraceClosest", C.c_float),
        ("msShade", C.c_float),
        ("msTraceAny", C.c_float),
        ("msResolve", C.c_float),
        ("msTotal", C.c_float),
        ("samplesPerWave", C.c_uint32),
        ("traversalOverflows", C.c_uint32),
        ("msStage", C.c_float * 10),
        ("raysShadowBounce", C.c_uint64),
    ]


RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("direction", "<f4", 3), ("tmax", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("instance", "<u4"), ("geometry", "<u4"), ("primitive", "<u4")])
NODE_DTYPE = np.dtype(
    [
        ("p", "<f4", 3),
        ("sx", "<u2"),
        ("sy", "<u2"),
        ("childBase", "<u4"),
        ("primBase", "<u4"),
        ("leafBits", "<u4"),
        ("sz", "<u2"),
        ("imask", "u1"),
        ("reserved", "u1"),
        ("qlox", "u1", 8),
        ("qloy", "u1", 8),
        ("qloz", "u1", 8),
        ("qhix", "u1", 8),
        ("qhiy", "u1", 8),
        ("qhiz", "u1", 8),
    ]
)
TRI_DTYPE = np.dtype([("v0", "<f4", 3), ("v1", "<f4", 3), ("v2", "<f4", 3), ("geometry", "<u4"), ("primitive", "<u4"), ("buildIndex", "<u4")])
assert NODE_DTYPE.itemsize == 80 and TRI_DTYPE.itemsize == 48 and RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 24

# every symbol include/vkrt.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "vkrt_version",
    "vkrt_context_create",
    "vkrt_context_destroy",
    "vkrt_last_error",
    "vkrt_context_synchronize",
    "vkrt_context_stream",
    "vkrt_trace_set_lane_refill",
    "vkrt_rasterize",
    "vkrt_trace_set_texture_lod",
    "vkrt_textures_generate_mips",
    "vkrt_texture_info",
    "vkrt_texture_read",
    "vkrt_blas_sizes",
    "vkrt_blas_build",
    "vkrt_blas_rebuild",
    "vkrt_blas_destroy",
    "vkrt_blas_reference",
    "vkrt_blas_info",
    "vkrt_blas_read",
    "vkrt_tlas_build",
    "vkrt_tlas_destroy",
    "vkrt_tlas_info",
    "vkrt_tlas_read",
    "vkrt_materials_set",
    "vkrt_textures_set",
    "vkrt_frame_create",
    "vkrt_frame_destroy",
    "vkrt_trace",
    "vkrt_frame_read",
    "vkrt_frame_present",
    "vkrt_frame_present_wait",
    "vkrt_host_alloc",
    "vkrt_host_free",
    "vkrt_frame_device_bgra8",
    "vkrt_trace_rays",
    "vkrt_trace_rays_static",
    "vkrt_stats_enable",
    "vkrt_stats_reset",
    "vkrt_stats_get",
    "vkrt_comm_unique_id",
    "vkrt_comm_init",
    "vkrt_comm_emulate",
    "vkrt_comm_destroy",
    "vkrt_tile_owner",
    "vkrt_hash_fnv1a64",
    "vkrt_tiles_gather",
    "vkrt_textures_broadcast",
    "vkrt_tiles_gather_wait",
]

STATIC_MAX_RAYS = 65536  # VKRT_STATIC_MAX_RAYS

_lib = None


def load_library(path: Optional[str] = None):
    """Loads libvkrt_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise FileNotFoundError(f"{p} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                                "(there is no CPU fallback)")
    lib = C.CDLL(p)
    vp, u32, u64, i32, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t
    sig = {
        "vkrt_version": (i32, []),
        "vkrt_context_create": (i32, [i32, C.POINTER(vp)]),
        "vkrt_context_destroy": (None, [vp]),
        "vkrt_last_error": (C.c_char_p, [vp]),
        "vkrt_context_synchronize": (i32, [vp]),
        "vkrt_context_stream": (vp, [vp]),
        "vkrt_trace_set_lane_refill": (i32, [vp, u32]),
        "vkrt_blas_sizes": (i32, [vp, C.POINTER(GeometryRange), u32, u64, u64, C.POINTER(BuildSizes)]),
        "vkrt_blas_build": (i32, [vp, vp, sz, u32, vp, sz, C.POINTER(GeometryRange), u32, u32, C.POINTER(vp)]),
        "vkrt_blas_rebuild": (i32, [vp, vp, vp, sz, u32]),
        "vkrt_blas_destroy": (None, [vp, vp]),
        "vkrt_blas_reference": (u64, [vp]),
        "vkrt_blas_info": (i32, [vp, vp, C.POINTER(BlasInfo)]),
        "vkrt_blas_read": (i32, [vp, vp, vp, sz, vp, sz]),
        "vkrt_tlas_build": (i32, [vp, C.POINTER(Instance), u32, C.POINTER(vp)]),
        "vkrt_tlas_destroy": (None, [vp, vp]),
        "vkrt_tlas_info": (i32, [vp, vp, C.POINTER(TlasInfo)]),
        "vkrt_tlas_read": (i32, [vp, vp, vp, sz, vp, sz]),
        "vkrt_materials_set": (i32, [vp, C.POINTER(SubmeshInfo), u32]),
        "vkrt_textures_set": (i32, [vp, C.POINTER(Image), u32]),
        "vkrt_textures_generate_mips": (i32, [vp]),
        "vkrt_texture_info": (i32, [vp, u32, u32, C.POINTER(u32), C.POINTER(u32), C.POINTER(u32)]),
        "vkrt_texture_read": (i32, [vp, u32, u32, vp, sz]),
        "vkrt_trace_set_texture_lod": (i32, [vp, i32]),
        "vkrt_rasterize": (i32, [vp, vp, vp, u32, vp]),
        "vkrt_frame_create": (i32, [vp, u32, u32, C.POINTER(vp)]),
        "vkrt_frame_destroy": (None, [vp, vp]),
        "vkrt_trace": (i32, [vp, vp, C.POINTER(Uniforms), u32, u32, u32, vp]),
        "vkrt_frame_read": (i32, [vp, vp, i32, vp, sz]),
        "vkrt_frame_present": (i32, [vp, vp, vp, sz, C.POINTER(C.c_uint64)]),
        "vkrt_frame_present_wait": (i32, [vp, vp, C.c_uint64]),
        "vkrt_host_alloc": (i32, [vp, sz, C.POINTER(vp)]),
        "vkrt_host_free": (None, [vp, vp]),
        "vkrt_frame_device_bgra8": (vp, [vp]),
        "vkrt_trace_rays": (i32, [vp, vp, vp, u64, i32, vp]),
        "vkrt_trace_rays_static": (i32, [vp, vp, vp, u32, i32, vp]),
        "vkrt_stats_enable": (i32, [vp, i32, i32]),
        "vkrt_stats_reset": (i32, [vp]),
        "vkrt_stats_get": (i32, [vp, C.POINTER(Stats)]),
        "vkrt_comm_unique_id": (i32, [vp]),
        "vkrt_comm_init": (i32, [vp, vp, i32, i32]),
        "vkrt_comm_emulate": (i32, [vp, i32, i32]),
        "vkrt_comm_destroy": (i32, [vp]),
        "vkrt_tile_owner": (i32, [u32, u32, u32, i32]),
        "vkrt_hash_fnv1a64": (u64, [vp, sz]),
        "vkrt_tiles_gather": (i32, [vp, vp, i32]),
        "vkrt_textures_broadcast": (i32, [vp, i32]),
        "vkrt_tiles_gather_wait": (i32, [vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One CUDA device + stream (headless stand-in for the device half of the reference's Context)."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        h = C.c_void_p()
        r = self.lib.vkrt_context_create(device, C.byref(h))
        if r != 0:
            raise VkrtError(r, (self.lib.vkrt_last_error(None) or b"").decode())
        self.handle = h
        self._keep = []
        self._pinned = {}

    def check(self, r: int):
        if r != 0:
            raise VkrtError(r, (self.lib.vkrt_last_error(self.handle) or b"").decode())

    def close(self):
        if self.handle:
            for p in list(getattr(self, "_pinned", {}).values()):
                self.lib.vkrt_host_free(self.handle, C.c_void_p(p))
            self._pinned = {}
            self.lib.vkrt_context_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def synchronize(self):
        self.check(self.lib.vkrt_context_synchronize(self.handle))

    @property
    def stream(self) -> int:
        return int(self.lib.vkrt_context_stream(self.handle) or 0)

    def set_materials(self, rows: np.ndarray):
        rows = np.ascontiguousarray(rows, dtype=np.int32).reshape(-1, 4)
        self.check(self.lib.vkrt_materials_set(self.handle, rows.ctypes.data_as(C.POINTER(SubmeshInfo)), rows.shape[0]))

    def set_textures(self, images: Sequence[np.ndarray]):
        arr = (Image * len(images))()
        keep = []
        for i, im in enumerate(images):
            im = np.ascontiguousarray(im, dtype=np.uint8)
            assert im.ndim == 3 and im.shape[2] == 4
            keep.append(im)
            arr[i].width, arr[i].height, arr[i].rgba8 = im.shape[1], im.shape[0], im.ctypes.data
        self.check(self.lib.vkrt_textures_set(self.handle, arr, len(images)))

    def textures_broadcast(self, root: int = 0):
        """Collective over the communicator: ranks other than `root` receive root's uploaded textures (vkrt_textures_broadcast)."""
        self.check(self.lib.vkrt_textures_broadcast(self.handle, root))

    def generate_mips(self):
        """Raytracer::createMipmaps (src/Raytracer.cpp:572-640) for every image, on the device."""
        self.check(self.lib.vkrt_textures_generate_mips(self.handle))

    def texture_info(self, image: int, level: int = 0):
        """(width, height, level count) of one level of one image."""
        w, h, n = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.check(self.lib.vkrt_texture_info(self.handle, image, level, C.byref(w), C.byref(h), C.byref(n)))
        return w.value, h.value, n.value

    def read_texture(self, image: int, level: int = 0) -> np.ndarray:
        w, h, _ = self.texture_info(image, level)
        out = np.empty((h, w, 4), dtype=np.uint8)
        self.check(self.lib.vkrt_texture_read(self.handle, image, level, _ptr(out), out.nbytes))
        return out

    def set_texture_lod(self, mode: int):
        """TEXTURE_LOD_ZERO (the reference's ray-tracing path, default) or TEXTURE_LOD_RAY_CONES."""
        self.check(self.lib.vkrt_trace_set_texture_lod(self.handle, int(mode)))

    def host_array(self, shape, dtype=np.uint8) -> np.ndarray:
        """numpy view of page-locked host memory (vkrt_host_alloc); released with host_free(array) or with the context."""
        nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self.check(self.lib.vkrt_host_alloc(self.handle, nbytes, C.byref(p)))
        buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        self._pinned[arr.ctypes.data] = p.value
        return arr

    def host_free(self, arr: np.ndarray):
        p = self._pinned.pop(arr.ctypes.data, None)
        if p:
            self.lib.vkrt_host_free(self.handle, C.c_void_p(p))

    def stats_enable(self, count_traversal: bool, time_stages: bool = False):
        self.check(self.lib.vkrt_stats_enable(self.handle, int(count_traversal), int(time_stages)))

    def set_lane_refill(self, lanes: int):
        self.check(self.lib.vkrt_trace_set_lane_refill(self.handle, lanes))

    def stats_reset(self):
        self.check(self.lib.vkrt_stats_reset(self.handle))

    def stats(self) -> Stats:
        s = Stats()
        self.check(self.lib.vkrt_stats_get(self.handle, C.byref(s)))
        return s

    def comm_emulate(self, rank: int, size: int):
        self.check(self.lib.vkrt_comm_emulate(self.handle, rank, size))

    def comm_init(self, unique_id: bytes, rank: int, size: int):
        buf = (C.c_uint8 * UNIQUE_ID_BYTES).from_buffer_copy(unique_id)
        self.check(self.lib.vkrt_comm_init(self.handle, buf, rank, size))

    @staticmethod
    def comm_unique_id() -> bytes:
        lib = load_library()
        buf = (C.c_uint8 * UNIQUE_ID_BYTES)()
        r = lib.vkrt_comm_unique_id(buf)
        if r != 0:
            raise VkrtError(r, (lib.vkrt_last_error(None) or b"").decode())
        return bytes(buf)


def _buffer(a, dtype=None):
    """(pointer, bytes, elements, on_device, keepalive) of a numpy array or of a CUDA torch tensor (duck-typed: data_ptr/is_cuda)."""
    if hasattr(a, "data_ptr") and getattr(a, "is_cuda", False):
        a = a.contiguous()
        return C.c_void_p(a.data_ptr()), a.numel() * a.element_size(), a.numel(), True, a
    a = np.ascontiguousarray(a, dtype=dtype)
    return _ptr(a), a.nbytes, a.size, False, a


class Blas:
    def __init__(self, ctx: Context, vertices, indices, ranges: np.ndarray, flags: int = 0, stride: int = 64):
        """vertices: (V,16) float32 (Model::Vertex) or raw bytes with `stride`; indices: global u32; ranges: rows of
        (primitiveCount, maxVertex, indexByteOffset).  numpy arrays (host) or CUDA torch tensors (both on the device:
        VKRT_BUILD_INPUT_ON_DEVICE is added)."""
        self.ctx = ctx
        vp, vbytes, _, vdev, vkeep = _buffer(vertices)
        ip, _, icount, idev, ikeep = _buffer(indices, None if hasattr(indices, "data_ptr") else np.uint32)
        if vdev != idev:
            raise ValueError("vertices and indices must both be host arrays or both be CUDA tensors")
        if vdev:
            flags |= BUILD_INPUT_ON_DEVICE
        rng = (GeometryRange * len(ranges))()
        for i, r in enumerate(ranges):
            rng[i].primitiveCount, rng[i].maxVertex, rng[i].indexByteOffset = int(r[0]), int(r[1]), int(r[2])
        h = C.c_void_p()
        ctx.check(ctx.lib.vkrt_blas_build(ctx.handle, vp, vbytes, stride, ip, icount, rng, len(ranges), flags, C.byref(h)))
        self.handle = h
        self._ranges = rng

    def rebuild(self, vertices, flags: int = 0):
        vp, vbytes, _, vdev, vkeep = _buffer(vertices)
        if vdev:
            flags |= BUILD_INPUT_ON_DEVICE
        self.ctx.check(self.ctx.lib.vkrt_blas_rebuild(self.ctx.handle, self.handle, vp, vbytes, flags))

    @property
    def reference(self) -> int:
        return int(self.ctx.lib.vkrt_blas_reference(self.handle))

    def info(self) -> BlasInfo:
        out = BlasInfo()
        self.ctx.check(self.ctx.lib.vkrt_blas_info(self.ctx.handle, self.handle, C.byref(out)))
        return out

    def read(self):
        info = self.info()
        nodes = np.zeros(int(info.nodeCount), dtype=NODE_DTYPE)
        tris = np.zeros(int(info.triangleCount), dtype=TRI_DTYPE)
        self.ctx.check(self.ctx.lib.vkrt_blas_read(self.ctx.handle, self.handle, _ptr(nodes), nodes.nbytes,
                                                   _ptr(tris) if tris.size else None, tris.nbytes))
        return nodes, tris

    def close(self):
        if self.handle and self.ctx.handle:
            self.ctx.lib.vkrt_blas_destroy(self.ctx.handle, self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


INSTANCE_DTYPE = np.dtype([("transform", "<f4", 12), ("index_mask", "<u4"), ("sbt_flags", "<u4"), ("reference", "<u8")])
assert INSTANCE_DTYPE.itemsize == 64


class Tlas:
    def __init__(self, ctx: Context, blases: Sequence[Blas], transforms: np.ndarray, masks: Optional[np.ndarray] = None):
        """transforms: (N,12) row-major 3x4; blases: one Blas per instance (or a single one, shared); masks: (N,) 8-bit
        VkAccelerationStructureInstanceKHR::mask values (default 0xFF, src/Raytracer.cpp:1174)."""
        self.ctx = ctx
        t = np.ascontiguousarray(transforms, dtype=np.float32).reshape(-1, 12)
        n = t.shape[0]
        # the 64-byte VkAccelerationStructureInstanceKHR records, filled as arrays (thousands of instances)
        inst = np.zeros(n, dtype=INSTANCE_DTYPE)
        inst["transform"] = t
        m = np.full(n, 0xFF, dtype=np.uint32) if masks is None else np.asarray(masks, dtype=np.uint32) & 0xFF
        inst["index_mask"] = m << 24  # instanceCustomIndex 0 : 24, mask : 8
        inst["sbt_flags"] = np.uint32(0x1 << 24)  # sbtOffset 0, VK_GEOMETRY_INSTANCE_TRIANGLE_FACING_CULL_DISABLE_BIT_KHR
        if len(blases) > 1:
            inst["reference"] = np.array([b.reference for b in blases], dtype=np.uint64)
        else:
            inst["reference"] = np.uint64(blases[0].reference)
        h = C.c_void_p()
        ctx.check(ctx.lib.vkrt_tlas_build(ctx.handle, inst.ctypes.data_as(C.POINTER(Instance)), n, C.byref(h)))
        self.handle = h
        self._blases = list(blases)

    def info(self) -> "TlasInfo":
        out = TlasInfo()
        self.ctx.check(self.ctx.lib.vkrt_tlas_info(self.ctx.handle, self.handle, C.byref(out)))
        return out

    def read(self):
        """(top-level nodes, instance index per top-level leaf); both empty for a single instance."""
        info = self.info()
        nodes = np.zeros(int(info.nodeCount), dtype=NODE_DTYPE)
        leaves = np.zeros(int(info.instanceCount) if info.nodeCount else 0, dtype=np.uint32)
        self.ctx.check(self.ctx.lib.vkrt_tlas_read(self.ctx.handle, self.handle, _ptr(nodes) if nodes.size else None, nodes.nbytes,
                                                   _ptr(leaves) if leaves.size else None, leaves.nbytes))
        return nodes, leaves

    def trace_rays(self, rays: np.ndarray, any_hit: bool = False) -> np.ndarray:
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        self.ctx.check(self.ctx.lib.vkrt_trace_rays(self.ctx.handle, self.handle, _ptr(rays), rays.shape[0], int(any_hit), _ptr(hits)))
        return hits

    def trace_rays_static(self, rays: np.ndarray, any_hit: bool = False) -> np.ndarray:
        """vkrt_trace_rays_static: the same batch through the library's file-scope scratch array, no allocation on the
        device; at most STATIC_MAX_RAYS rays per call."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        self.ctx.check(self.ctx.lib.vkrt_trace_rays_static(self.ctx.handle, self.handle, _ptr(rays), rays.shape[0], int(any_hit),
                                                           _ptr(hits)))
        return hits

    def close(self):
        if self.handle and self.ctx.handle:
            self.ctx.lib.vkrt_tlas_destroy(self.ctx.handle, self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Frame:
    def __init__(self, ctx: Context, width: int, height: int):
        self.ctx, self.width, self.height = ctx, width, height
        h = C.c_void_p()
        ctx.check(ctx.lib.vkrt_frame_create(ctx.handle, width, height, C.byref(h)))
        self.handle = h

    def trace(self, tlas: Tlas, uniforms: Uniforms, spp: int = 1, frame_index: int = 0, flags: int = 0):
        self.ctx.check(self.ctx.lib.vkrt_trace(self.ctx.handle, tlas.handle, C.byref(uniforms), spp, frame_index, flags, self.handle))

    def rasterize(self, blas: "Blas", wvp: np.ndarray, flags: int = 0):
        """One frame of the reference's Rasterizer (src/Rasterizer.cpp:119-147) through the column-major wvp matrix."""
        m = np.ascontiguousarray(wvp, dtype=np.float32).reshape(16)
        self.ctx.check(self.ctx.lib.vkrt_rasterize(self.ctx.handle, blas.handle, _ptr(m), flags, self.handle))

    def gather(self, root: int = 0):
        self.ctx.check(self.ctx.lib.vkrt_tiles_gather(self.ctx.handle, self.handle, root))

    def gather_wait(self):
        """Stream-ordered: the context stream waits for the outstanding tile gather of this frame."""
        self.ctx.check(self.ctx.lib.vkrt_tiles_gather_wait(self.ctx.handle, self.handle))

    def read(self, aov: int, out: Optional[np.ndarray] = None) -> np.ndarray:
        shape, dt = {
            AOV_BGRA8: ((self.height, self.width, 4), np.uint8),
            AOV_RGB32F: ((self.height, self.width, 3), np.float32),
            AOV_TRIANGLE_ID: ((self.height, self.width), np.uint32),
            AOV_SHADOW_MASK: ((self.height, self.width), np.uint8),
        }[aov]
        if out is None:
            out = np.empty(shape, dtype=dt)
        self.ctx.check(self.ctx.lib.vkrt_frame_read(self.ctx.handle, self.handle, aov, _ptr(out), 0))
        return out

    def present(self, out: np.ndarray) -> int:
        """Queues the BGRA8 image copy into `out` (page-locked host memory: Context.host_array or a pinned torch tensor)
        and returns its ticket at once; the next trace may be queued immediately (vkrt_frame_present)."""
        assert out.nbytes >= self.width * self.height * 4 and out.flags["C_CONTIGUOUS"]
        t = C.c_uint64()
        self.ctx.check(self.ctx.lib.vkrt_frame_present(self.ctx.handle, self.handle, _ptr(out), 0, C.byref(t)))
        return int(t.value)

    def present_wait(self, ticket: int):
        self.ctx.check(self.ctx.lib.vkrt_frame_present_wait(self.ctx.handle, self.handle, ticket))

    @property
    def device_bgra8(self) -> int:
        return int(self.ctx.lib.vkrt_frame_device_bgra8(self.handle) or 0)

    def close(self):
        if self.handle and self.ctx.handle:
            self.ctx.lib.vkrt_frame_destroy(self.ctx.handle, self.handle)
        self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
