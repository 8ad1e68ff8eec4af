"""Test-side bindings: the CPU oracle (oracle/libvkrt_oracle.so) and the host emulation of the kernel logic
(tests/emu/libvkrt_emu.so).  Only tests, smoke() and bench.py's CPU-baseline legs load these."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "libvkrt_oracle.so")
EMU_LIB = os.path.join(ROOT, "tests", "emu", "libvkrt_emu.so")

import sys

if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from vkrt_b200 import binding as B  # noqa: E402


def _ensure(path: str, makedir: str):
    # always ask make: the harnesses include the library's headers, a stale build would test yesterday's logic
    try:
        subprocess.check_call(["make", "-C", makedir], stdout=subprocess.DEVNULL)
    except (OSError, subprocess.CalledProcessError):
        if not os.path.exists(path):
            raise
    return path


class OrcGeometryRange(C.Structure):
    _fields_ = [("primitiveCount", C.c_uint32), ("maxVertex", C.c_uint32), ("indexByteOffset", C.c_uint64)]


class OrcImage(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("rgba8", C.c_void_p)]


class OrcStats(C.Structure):
    _fields_ = [("raysPrimary", C.c_uint64), ("raysShadow", C.c_uint64), ("raysReflection", C.c_uint64),
                ("nodesVisited", C.c_uint64), ("trianglesTested", C.c_uint64), ("seconds", C.c_double)]


MODE_BRUTE, MODE_BVH = 0, 1
_oracle = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        lib = C.CDLL(_ensure(ORACLE_LIB, os.path.join(ROOT, "oracle")))
        vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
        lib.orc_scene_create.restype = vp
        lib.orc_scene_create.argtypes = [vp, u64, vp, u64, C.POINTER(OrcGeometryRange), vp, u32, C.POINTER(OrcImage), u32, vp, u32]
        lib.orc_scene_destroy.argtypes = [vp]
        lib.orc_scene_set_instance_masks.argtypes = [vp, vp, u32]
        lib.orc_scene_update_vertices.argtypes = [vp, vp, u64]
        lib.orc_scene_build_seconds.restype = C.c_double
        lib.orc_scene_build_seconds.argtypes = [vp]
        lib.orc_render.argtypes = [vp, vp, u32, u32, u32, i32, i32, u32, u32, u32, u32, u32, u32, vp, vp, vp, vp, C.POINTER(OrcStats)]
        lib.orc_trace_rays.argtypes = [vp, vp, u64, i32, i32, i32, vp, C.POINTER(OrcStats)]
        lib.orc_sample_texture.argtypes = [vp, u32, C.c_float, C.c_float, vp]
        lib.orc_halton23.argtypes = [u32, vp]
        lib.orc_primary_ray.argtypes = [vp, u32, u32, u32, u32, C.c_float, C.c_float, vp]
        lib.orc_affine_inverse.argtypes = [vp, vp]
        lib.orc_rasterize.argtypes = [vp, vp, u32, u32, i32, vp, vp]
        lib.orc_scene_generate_mips.argtypes = [vp]
        lib.orc_scene_set_texture_lod.argtypes = [vp, i32]
        lib.orc_texture_level_count.restype = u32
        lib.orc_texture_level_count.argtypes = [vp, u32]
        lib.orc_texture_level.restype = vp
        lib.orc_texture_level.argtypes = [vp, u32, u32, C.POINTER(u32), C.POINTER(u32)]
        lib.orc_sample_texture_lod.argtypes = [vp, u32, C.c_float, C.c_float, C.c_float, vp]
        _oracle = lib
    return _oracle


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleScene:
    """flat: output of vkrt_b200.raytracer.flatten_model; instances (N,12)."""

    def __init__(self, flat, instances):
        lib = oracle_lib()
        self.lib = lib
        self.v = np.ascontiguousarray(flat["vertices"], dtype=np.float32)
        self.i = np.ascontiguousarray(flat["indices"], dtype=np.uint32)
        rng = (OrcGeometryRange * len(flat["ranges"]))()
        for k, r in enumerate(flat["ranges"]):
            rng[k].primitiveCount, rng[k].maxVertex, rng[k].indexByteOffset = int(r[0]), int(r[1]), int(r[2])
        self.rows = np.ascontiguousarray(flat["rows"], dtype=np.int32)
        self.imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in flat["images"]]
        arr = (OrcImage * len(self.imgs))()
        for k, im in enumerate(self.imgs):
            arr[k].width, arr[k].height, arr[k].rgba8 = im.shape[1], im.shape[0], im.ctypes.data
        inst = np.ascontiguousarray(instances, dtype=np.float32).reshape(-1, 12)
        self.h = lib.orc_scene_create(_p(self.v), self.v.shape[0], _p(self.i), self.i.size, rng, _p(self.rows), len(flat["ranges"]),
                                      arr, len(self.imgs), _p(inst), inst.shape[0])

    def __del__(self):
        try:
            if self.h:
                self.lib.orc_scene_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_instance_masks(self, masks):
        m = np.ascontiguousarray(masks, dtype=np.uint8)
        self.lib.orc_scene_set_instance_masks(self.h, _p(m), m.size)

    def build_seconds(self):
        return self.lib.orc_scene_build_seconds(self.h)

    def render(self, uniforms, width, height, spp=1, mode=MODE_BVH, threads=0, rect=None, stride=(1, 1)):
        x0, y0, x1, y1 = rect or (0, 0, width, height)
        rgb = np.zeros((height, width, 3), dtype=np.float32)
        bgra = np.zeros((height, width, 4), dtype=np.uint8)
        tri = np.full((height, width), 0xFFFFFFFF, dtype=np.uint32)
        mask = np.zeros((height, width), dtype=np.uint8)
        stats = OrcStats()
        self.lib.orc_render(self.h, C.byref(uniforms), width, height, spp, mode, threads, x0, y0, x1, y1, stride[0], stride[1],
                            _p(rgb), _p(bgra), _p(tri), _p(mask), C.byref(stats))
        return {"rgb": rgb, "bgra": bgra, "tri": tri, "mask": mask, "stats": stats}

    def trace_rays(self, rays, any_hit=False, mode=MODE_BVH, threads=0):
        rays = np.ascontiguousarray(rays, dtype=B.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=B.HIT_DTYPE)
        stats = OrcStats()
        self.lib.orc_trace_rays(self.h, _p(rays), rays.shape[0], int(any_hit), mode, threads, _p(hits), C.byref(stats))
        return hits, stats

    def sample_texture(self, index, u, v):
        out = np.zeros(4, dtype=np.float32)
        self.lib.orc_sample_texture(self.h, index, np.float32(u), np.float32(v), _p(out))
        return out

    def generate_mips(self):
        self.lib.orc_scene_generate_mips(self.h)

    def rasterize(self, wvp, width, height, threads=0):
        """The reference's Rasterizer frame (needs generate_mips): {"bgra", "tri"}."""
        wvp = np.ascontiguousarray(wvp, dtype=np.float32).reshape(16)
        bgra = np.zeros((height, width, 4), dtype=np.uint8)
        tri = np.zeros((height, width), dtype=np.uint32)
        self.lib.orc_rasterize(self.h, _p(wvp), width, height, threads, _p(bgra), _p(tri))
        return {"bgra": bgra, "tri": tri}

    def set_texture_lod(self, mode):
        self.lib.orc_scene_set_texture_lod(self.h, int(mode))

    def level_count(self, index):
        return self.lib.orc_texture_level_count(self.h, index)

    def texture_level(self, index, level):
        w, h = C.c_uint32(), C.c_uint32()
        p = self.lib.orc_texture_level(self.h, index, level, C.byref(w), C.byref(h))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h.value, w.value, 4)).copy()

    def sample_texture_lod(self, index, u, v, footprint2):
        out = np.zeros(4, dtype=np.float32)
        self.lib.orc_sample_texture_lod(self.h, index, np.float32(u), np.float32(v), np.float32(footprint2), _p(out))
        return out


_emu = None


def emu_lib():
    global _emu
    if _emu is None:
        lib = C.CDLL(_ensure(EMU_LIB, os.path.join(ROOT, "tests", "emu")))
        vp, u32, u64, i32, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_size_t
        lib.emu_blas_build.restype = vp
        lib.emu_blas_build.argtypes = [vp, sz, u32, vp, sz, C.POINTER(B.GeometryRange), u32]
        lib.emu_blas_build_flags.restype = vp
        lib.emu_blas_build_flags.argtypes = [vp, sz, u32, vp, sz, C.POINTER(B.GeometryRange), u32, u32]
        lib.emu_blas_destroy.argtypes = [vp]
        lib.emu_blas_node_count.restype = u32
        lib.emu_blas_node_count.argtypes = [vp]
        lib.emu_blas_triangle_count.restype = u32
        lib.emu_blas_triangle_count.argtypes = [vp]
        lib.emu_blas_nodes.restype = vp
        lib.emu_blas_nodes.argtypes = [vp]
        lib.emu_blas_triangles.restype = vp
        lib.emu_blas_triangles.argtypes = [vp]
        lib.emu_blas_sah.restype = C.c_float
        lib.emu_blas_sah.argtypes = [vp]
        lib.emu_scene_create.restype = vp
        lib.emu_scene_create.argtypes = [vp, vp, u32, vp, u32, C.POINTER(B.Image), u32]
        lib.emu_scene_destroy.argtypes = [vp]
        lib.emu_trace_rays.argtypes = [vp, vp, u64, i32, vp, C.POINTER(u64), C.POINTER(u64)]
        lib.emu_render.argtypes = [vp, vp, u32, u32, u32, u32, u32, u32, u32, u32, u32, vp, vp, vp, vp]
        lib.emu_rasterize.argtypes = [vp, vp, u32, u32, vp, vp]
        lib.emu_scene_generate_mips.argtypes = [vp]
        lib.emu_scene_set_texture_lod.argtypes = [vp, i32]
        lib.emu_texture_level.restype = vp
        lib.emu_texture_level.argtypes = [vp, u32, u32, C.POINTER(u32), C.POINTER(u32)]
        lib.emu_sample_texture_lod.argtypes = [vp, u32, C.c_float, C.c_float, C.c_float, vp]
        _emu = lib
    return _emu


class EmuScene:
    """Host emulation of libvkrt_b200's kernel logic over the same flat scene."""

    def __init__(self, flat, instances, flags=0):
        lib = emu_lib()
        self.lib = lib
        self.v = np.ascontiguousarray(flat["vertices"], dtype=np.float32)
        self.i = np.ascontiguousarray(flat["indices"], dtype=np.uint32)
        rng = (B.GeometryRange * len(flat["ranges"]))()
        for k, r in enumerate(flat["ranges"]):
            rng[k].primitiveCount, rng[k].maxVertex, rng[k].indexByteOffset = int(r[0]), int(r[1]), int(r[2])
        self.blas = lib.emu_blas_build_flags(_p(self.v), self.v.nbytes, 64, _p(self.i), self.i.size, rng, len(flat["ranges"]), flags)
        assert self.blas, "emu_blas_build failed"
        self.rows = np.ascontiguousarray(flat["rows"], dtype=np.int32)
        self.imgs = [np.ascontiguousarray(im, dtype=np.uint8) for im in flat["images"]]
        arr = (B.Image * len(self.imgs))()
        for k, im in enumerate(self.imgs):
            arr[k].width, arr[k].height, arr[k].rgba8 = im.shape[1], im.shape[0], im.ctypes.data
        inst = np.ascontiguousarray(instances, dtype=np.float32).reshape(-1, 12)
        self.h = lib.emu_scene_create(self.blas, _p(inst), inst.shape[0], _p(self.rows), self.rows.shape[0], arr, len(self.imgs))

    def __del__(self):
        try:
            if self.h:
                self.lib.emu_scene_destroy(self.h)
                self.lib.emu_blas_destroy(self.blas)
                self.h = None
        except Exception:
            pass

    def generate_mips(self):
        self.lib.emu_scene_generate_mips(self.h)

    def rasterize(self, wvp, width, height):
        wvp = np.ascontiguousarray(wvp, dtype=np.float32).reshape(16)
        bgra = np.zeros((height, width, 4), dtype=np.uint8)
        tri = np.zeros((height, width), dtype=np.uint32)
        self.lib.emu_rasterize(self.h, _p(wvp), width, height, _p(bgra), _p(tri))
        return {"bgra": bgra, "tri": tri}

    def set_texture_lod(self, mode):
        self.lib.emu_scene_set_texture_lod(self.h, int(mode))

    def texture_level(self, image, level):
        w, h = C.c_uint32(), C.c_uint32()
        p = self.lib.emu_texture_level(self.h, image, level, C.byref(w), C.byref(h))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(h.value, w.value, 4)).copy()

    def sample_texture_lod(self, image, u, v, footprint2):
        out = np.zeros(4, dtype=np.float32)
        self.lib.emu_sample_texture_lod(self.h, image, np.float32(u), np.float32(v), np.float32(footprint2), _p(out))
        return out

    def read_blas(self):
        n, t = self.lib.emu_blas_node_count(self.blas), self.lib.emu_blas_triangle_count(self.blas)
        nodes = np.ctypeslib.as_array(C.cast(self.lib.emu_blas_nodes(self.blas), C.POINTER(C.c_uint8)), shape=(n * 80,)).copy().view(B.NODE_DTYPE)
        tris = np.ctypeslib.as_array(C.cast(self.lib.emu_blas_triangles(self.blas), C.POINTER(C.c_uint8)), shape=(max(t, 1) * 48,)).copy().view(B.TRI_DTYPE)[:t]
        return nodes, tris

    def sah(self):
        return self.lib.emu_blas_sah(self.blas)

    def trace_rays(self, rays, any_hit=False):
        rays = np.ascontiguousarray(rays, dtype=B.RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=B.HIT_DTYPE)
        nodes, tris = C.c_uint64(), C.c_uint64()
        self.lib.emu_trace_rays(self.h, _p(rays), rays.shape[0], int(any_hit), _p(hits), C.byref(nodes), C.byref(tris))
        return hits, nodes.value, tris.value

    def render(self, uniforms, width, height, spp=1, rect=None, stride=(1, 1)):
        x0, y0, x1, y1 = rect or (0, 0, width, height)
        rgb = np.zeros((height, width, 3), dtype=np.float32)
        bgra = np.zeros((height, width, 4), dtype=np.uint8)
        tri = np.full((height, width), 0xFFFFFFFF, dtype=np.uint32)
        mask = np.zeros((height, width), dtype=np.uint8)
        self.lib.emu_render(self.h, C.byref(uniforms), width, height, spp, x0, y0, x1, y1, stride[0], stride[1], _p(rgb), _p(bgra), _p(tri), _p(mask))
        return {"rgb": rgb, "bgra": bgra, "tri": tri, "mask": mask}


def random_rays(n, seed, lo, hi, tmin=0.001, tmax=1000.0):
    """Rays with origins inside [lo,hi] and random directions (not normalised on purpose)."""
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, dtype=B.RAY_DTYPE)
    lo, hi = np.asarray(lo, dtype=np.float32), np.asarray(hi, dtype=np.float32)
    rays["origin"] = (rng.random((n, 3), dtype=np.float32) * (hi - lo) + lo).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    rays["direction"] = d
    rays["tmin"] = tmin
    rays["tmax"] = tmax
    return rays


def compare_frames(a, b, sample=None):
    """Returns agreement statistics between two render dicts (tri / mask / rgb) over `sample` (boolean mask) pixels."""
    sel = sample if sample is not None else np.ones(a["tri"].shape, dtype=bool)
    n = int(sel.sum())
    tri_ok = (a["tri"] == b["tri"]) & sel
    mask_ok = (a["mask"] == b["mask"]) & sel
    both = tri_ok & mask_ok
    da = a["rgb"][both].astype(np.float64)
    db = b["rgb"][both].astype(np.float64)
    finite = np.isfinite(da).all(axis=1) & np.isfinite(db).all(axis=1)
    diff = np.abs(np.clip(da[finite], 0, 1) - np.clip(db[finite], 0, 1)).max() if finite.any() else 0.0
    bdiff = np.abs(a["bgra"][both].astype(np.int32) - b["bgra"][both].astype(np.int32)).max() if both.any() else 0
    return {"pixels": n, "tri_agree": tri_ok.sum() / max(n, 1), "mask_agree": mask_ok.sum() / max(n, 1),
            "rgb_max_diff": float(diff), "bgra_max_diff": int(bdiff),
            "rgb_bit_equal": bool(np.array_equal(a["rgb"][both].view(np.uint32), b["rgb"][both].view(np.uint32)))}
