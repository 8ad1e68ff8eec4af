"""Raytracer: Python mirror of the reference's `Raytracer` host logic (src/Raytracer.cpp) over the C ABI.

`flatten_model` restates createVertexAndIndexBuffer (src/Raytracer.cpp:642-742) and the SubmeshInfo table of
updateMaterialIndexDescriptorSet (:1412-1427); `Raytracer` owns context, BLAS, TLAS, materials, textures and a
frame, and `render()` is one vkrt_trace (+ tile gather) like one iteration of Raytracer::render (:132-203).
The C++ equivalent used by the headless executable lives in vkrt_b200/host/Raytracer.cpp.
"""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

from . import binding as B
from .camera import Camera, LIGHT_POSITIONS
from .scenes import REFERENCE_INSTANCE


def flatten_model(model: Dict) -> Dict:
    """Concatenate submesh vertices; rebase indices by the running vertex count; per submesh record
    {primitiveCount, maxVertex (local), indexByteOffset}; material rows with cumulative triangle offsets and the
    reference's clamp of missing normal / metallic-roughness indices to image 0 (base colour is not clamped)."""
    subs = model["submeshes"]
    vertices = np.concatenate([s["vertices"] for s in subs], axis=0).astype(np.float32)
    indices = np.empty(sum(s["indices"].size for s in subs), dtype=np.uint32)
    ranges = np.zeros((len(subs), 3), dtype=np.uint64)
    rows = np.zeros((len(subs), 4), dtype=np.int32)
    index_counter_offset = 0
    index_counter = 0
    index_byte_offset = 0
    index_buffer_offset = 0  # triangles
    for i, s in enumerate(subs):
        idx = s["indices"].astype(np.uint32)
        indices[index_counter:index_counter + idx.size] = idx + np.uint32(index_counter_offset)
        ranges[i] = (idx.size // 3, int(idx.max()) if idx.size else 0, index_byte_offset)
        base, mr, nrm = model["materials"][s["material"]]
        rows[i] = (base, max(mr, 0), max(nrm, 0), index_buffer_offset)
        index_counter += idx.size
        index_counter_offset += s["vertices"].shape[0]
        index_byte_offset += 4 * idx.size
        index_buffer_offset += idx.size // 3
    return {"vertices": vertices, "indices": indices, "ranges": ranges, "rows": rows, "images": model.get("images")}


class Raytracer:
    """Headless renderer: scene upload + acceleration-structure build in the constructor, `render()` per frame."""

    def __init__(self, model: Dict, width: int, height: int, device: int = 0, instances: Optional[np.ndarray] = None,
                 build_flags: int = 0, rank: int = 0, world_size: int = 1, unique_id: Optional[bytes] = None,
                 emulate_ranks: bool = False, flat: Optional[Dict] = None, textures_from_root: Optional[int] = None):
        """`flat` (optional) is an already flattened scene whose vertices/indices may be CUDA tensors (config 5)."""
        self.width, self.height = width, height
        self.flat = flat if flat is not None else flatten_model(model)
        self.instances = np.asarray(instances if instances is not None else REFERENCE_INSTANCE, dtype=np.float32).reshape(-1, 12)
        self.ctx = B.Context(device)
        if world_size > 1:
            if emulate_ranks:
                self.ctx.comm_emulate(rank, world_size)
            else:
                self.ctx.comm_init(unique_id, rank, world_size)
        self.rank, self.world_size = rank, world_size
        self.blas = B.Blas(self.ctx, self.flat["vertices"], self.flat["indices"], self.flat["ranges"], flags=build_flags)
        self.tlas = B.Tlas(self.ctx, [self.blas], self.instances)
        self.ctx.set_materials(self.flat["rows"])
        if self.flat.get("images") is not None:
            self.ctx.set_textures(self.flat["images"])
        if world_size > 1 and not emulate_ranks and textures_from_root is not None:
            # scene ingest for N ranks: only `textures_from_root` decoded the image files; the others get the texels over NVLink
            self.ctx.textures_broadcast(textures_from_root)
        self.ctx.generate_mips()  # createTextures -> createMipmaps, src/Raytracer.cpp:562 (sampled only in the ray-cone mode)
        self.frame = B.Frame(self.ctx, width, height)
        self.camera = Camera(np.float32(width) / np.float32(height))
        # Raytracer::setupCamera, src/Raytracer.cpp:267-271
        self.camera.setPosition((6.3, 4.5, -0.7))
        self.camera.setRotation((0.0, 1.57, 0.0))
        self.lights = LIGHT_POSITIONS
        self.frame_index = 0

    def render(self, spp: int = 1, flags: int = 0, gather: bool = True):
        u = self.camera.uniforms(self.lights)
        self.frame.trace(self.tlas, u, spp, self.frame_index, flags)
        if gather and self.world_size > 1 and self.ctx is not None:
            self.frame.gather(0)
        self.frame_index += 1

    def set_texture_lod(self, mode: int):
        """B.TEXTURE_LOD_ZERO: every texture() samples level 0, as the reference's ray-tracing shaders do (default);
        B.TEXTURE_LOD_RAY_CONES: trilinear at the ray-cone level of detail over the generated mip chain."""
        self.ctx.set_texture_lod(mode)

    def rebuild(self, vertices, flags: int = 0):
        """Per-frame full rebuild over moved vertices (BASELINE.json config 5): same storage, new BVH, new TLAS bounds."""
        self.blas.rebuild(vertices, flags)
        self.tlas.close()
        self.tlas = B.Tlas(self.ctx, [self.blas], self.instances)

    def readback(self, aov: int = B.AOV_BGRA8, out: Optional[np.ndarray] = None) -> np.ndarray:
        return self.frame.read(aov, out)

    def present(self, out: np.ndarray) -> int:
        """Asynchronous copy of the finished image into page-locked `out`; mirrors the copy to the acquired swapchain
        image + vkQueuePresentKHR (src/Raytracer.cpp:175-185, src/Context.cpp:154-180).  Returns a ticket."""
        return self.frame.present(out)

    def present_wait(self, ticket: int):
        """The wait acquireNextSwapchainImage implies before a ring image is reused (src/Context.cpp:141-152)."""
        self.frame.present_wait(ticket)

    def close(self):
        for o in (self.frame, self.tlas, self.blas):
            o.close()
        self.ctx.close()


class Rasterizer:
    """Headless mirror of the reference's second renderer (src/Rasterizer.hpp:11-17, src/Rasterizer.cpp): the scene is
    uploaded in the constructor, `render()` draws one frame with shader.vert / shader.frag semantics (base-colour texture
    through its mip chain, alpha test, 8x MSAA, depth test LESS) on the GPU via vkrt_rasterize."""

    def __init__(self, model: Dict, width: int, height: int, device: int = 0, flat: Optional[Dict] = None, rank: int = 0,
                 world_size: int = 1, unique_id: Optional[bytes] = None, emulate_ranks: bool = False,
                 textures_from_root: Optional[int] = None):
        self.width, self.height = width, height
        self.flat = flat if flat is not None else flatten_model(model)
        self.ctx = B.Context(device)
        self.rank, self.world_size = rank, world_size
        if world_size > 1:
            if emulate_ranks:
                self.ctx.comm_emulate(rank, world_size)
            else:
                self.ctx.comm_init(unique_id, rank, world_size)
        # vertex/index buffers (createVertexAndIndexBuffer, src/Rasterizer.cpp:1040-1085); the BLAS object holds the device copy
        self.blas = B.Blas(self.ctx, self.flat["vertices"], self.flat["indices"], self.flat["ranges"], flags=B.BUILD_PREFER_FAST_BUILD)
        self.ctx.set_materials(self.flat["rows"])
        if self.flat.get("images") is not None:
            self.ctx.set_textures(self.flat["images"])
        if world_size > 1 and not emulate_ranks and textures_from_root is not None:
            self.ctx.textures_broadcast(textures_from_root)
        self.ctx.generate_mips()  # createTextures -> createMipmaps, src/Rasterizer.cpp:560-700
        self.frame = B.Frame(self.ctx, width, height)
        self.camera = Camera(np.float32(width) / np.float32(height))
        # Rasterizer::setupCamera, src/Rasterizer.cpp:207-211
        self.camera.setPosition((-4.0, 2.0, -0.2))
        self.camera.setRotation((0.0, 1.51, 0.0))
        self.scale = 0.01  # Rasterizer::update, src/Rasterizer.cpp:189

    def wvp(self) -> np.ndarray:
        return self.camera.wvp(self.scale)

    def render(self, flags: int = 0, gather: bool = True):
        self.frame.rasterize(self.blas, self.wvp(), flags)
        if gather and self.world_size > 1:
            self.frame.gather(0)

    def readback(self, aov: int = B.AOV_BGRA8, out: Optional[np.ndarray] = None) -> np.ndarray:
        return self.frame.read(aov, out)

    def close(self):
        for o in (self.frame, self.blas):
            o.close()
        self.ctx.close()
