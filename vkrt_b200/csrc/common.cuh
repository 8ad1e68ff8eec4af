// common.cuh — shared types of libvkrt_b200 (device layouts, error plumbing, small math).
//
// Everything on the parity path is FP32 with a FIXED evaluation order and NO implicit FMA
// contraction (the library is compiled with -fmad=false); explicit __fmaf_rn is used only where
// results need to be conservative rather than bit-equal to the CPU oracle (box tests).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "../../include/vkrt.h"

namespace vkrt
{

// ---------------------------------------------------------------------------------------------
// Device data layouts
// ---------------------------------------------------------------------------------------------

// Compressed 8-wide BVH node, 80 bytes (Ylitie, Karras, Laine 2017, "Efficient Incoherent Ray Traversal on
// GPUs Through Compressed Wide BVHs").  Loaded as five 16-byte vectors.
struct alignas(16) Node8
{
    float px, py, pz;     // quantisation origin (node box minimum)
    uint16_t sx, sy;      // per-axis scale 2^(e-127) as the upper half of that float: e << 7 (a shift or a mask makes the float)
    uint32_t childBase;   // index of the first internal child node
    uint32_t primBase;    // index of the first primitive (packed triangle or instance) of this node
    uint32_t leafBits;    // bit 3*s+j set: slot s is a leaf holding a j-th triangle (j < 3); its packed triangle is
                          // primBase + popcount(leafBits below that bit): leaves are stored in slot order
    uint16_t sz;          // as sx, sy
    uint8_t imask;        // bit s set: slot s holds an internal child
    uint8_t reserved;     // BLAS nodes: 0.  TLAS nodes: bit s set: slot s is a leaf holding one instance (k_tlas_finalize)
    uint8_t qlox[8], qloy[8], qloz[8];
    uint8_t qhix[8], qhiy[8], qhiz[8];
};
// Packs a scale exponent byte the way Node8 stores it; the builder keeps e <= 239.
inline __host__ __device__ uint16_t packScaleExponent(uint32_t e) { return (uint16_t)(e << 7); }
static_assert(sizeof(Node8) == 80, "Node8 must be 80 bytes");

// Packed triangle, 48 bytes: raw vertex positions (the watertight test needs all three) + Vulkan ids.
struct alignas(16) PackedTri
{
    float v0x, v0y, v0z;
    float v1x, v1y, v1z;
    float v2x, v2y, v2z;
    uint32_t geometry;   // gl_GeometryIndexEXT
    uint32_t primitive;  // gl_PrimitiveID
    uint32_t buildIndex; // position in (geometry, primitive) order; tie-break key for equal-t hits
};
static_assert(sizeof(PackedTri) == 48, "PackedTri must be 48 bytes");

// Per-instance record used by traversal (w2o) and shading (o2w, geometry pointers).  Built on the host from the 64-byte
// VkAccelerationStructureInstanceKHR (src/Raytracer.cpp:1165-1177); the world box comes from k_instance_bounds.
struct alignas(16) InstanceRecord
{
    float w2o[12];
    float o2w[12];
    const Node8* nodes;
    const PackedTri* tris;
    const uint8_t* vertices; // device copy of the vertex buffer (stride bytes per vertex)
    const uint32_t* indices; // device copy of the (global) index buffer
    uint32_t vertexStride;
    uint32_t mask;           // VkAccelerationStructureInstanceKHR::mask (8 bits): visible to a ray iff mask & cullMask
    uint32_t customIndex;    // instanceCustomIndex (24 bits); the reference's shaders never read gl_InstanceCustomIndexEXT
    uint32_t flags;          // VkGeometryInstanceFlagsKHR; no ray of this pipeline culls, so TRIANGLE_FACING_CULL_DISABLE changes nothing
    float blasLo[3], blasHi[3]; // object-space bounds of the BLAS at TLAS build time
    float boxScale;             // 2^(114 - smallest scale exponent of the BLAS's nodes), see traverse.cuh::setBoxRay
    uint32_t pad[1];
};
static_assert(sizeof(InstanceRecord) == 176, "InstanceRecord layout");

struct TextureRecord
{
    const uint8_t* texels; // RGBA8 rows
    uint32_t width, height;
};

// Mip levels 1.. of a texture (vkrt_textures_generate_mips), tightly packed one after the other; a parallel array to
// TextureRecord so that the LOD-0 path reads exactly what it read before mips existed.
struct TextureMips
{
    const uint8_t* texels; // level 1 starts here
    uint32_t levels;       // floor(log2(max(w,h))) + 1, level 0 included
    uint32_t pad;
};

// Hit record written by the closest-hit traversal.
struct HitRecord
{
    float t, b1, b2;
    uint32_t tri;      // index into the instance's packed triangle array, 0xFFFFFFFF = miss
    uint32_t instance; // index into the TLAS instance array
};

constexpr uint32_t kInvalid = 0xFFFFFFFFu;
constexpr uint32_t kMaxInstances = 1u << 24; // as many as instanceCustomIndex can name; the AOV triangle id keeps the low 8 bits
constexpr float kRayTMin = 0.001f;    // shaders/shader.rgen:58, shader.rchit:139
constexpr float kRayTMax = 1000.0f;   // shaders/shader.rgen:60
constexpr int kTile = VKRT_TILE_SIZE; // multi-GPU image tile edge in pixels
constexpr int kTilePixels = kTile * kTile;

// ---------------------------------------------------------------------------------------------
// Error plumbing
// ---------------------------------------------------------------------------------------------
struct Error
{
    int code;
    std::string text;
};

#define VKRT_CUDA_TRY(ctx, expr)                                                                            \
    do                                                                                                      \
    {                                                                                                       \
        cudaError_t err__ = (expr);                                                                         \
        if (err__ != cudaSuccess)                                                                           \
        {                                                                                                   \
            return (ctx)->fail(err__ == cudaErrorMemoryAllocation ? VKRT_ERROR_OUT_OF_MEMORY : VKRT_ERROR_CUDA, \
                               std::string(#expr) + ": " + cudaGetErrorString(err__));                      \
        }                                                                                                   \
    } while (0)

// ---------------------------------------------------------------------------------------------
// Device buffer with grow-only capacity (library-owned scratch, reused across frames/rebuilds)
// ---------------------------------------------------------------------------------------------
struct DeviceBuffer
{
    void* ptr = nullptr;
    size_t capacity = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= capacity)
        {
            return cudaSuccess;
        }
        if (ptr)
        {
            cudaFree(ptr);
            ptr = nullptr;
            capacity = 0;
        }
        cudaError_t e = cudaMalloc(&ptr, bytes);
        if (e == cudaSuccess)
        {
            capacity = bytes;
        }
        return e;
    }
    void release()
    {
        if (ptr)
        {
            cudaFree(ptr);
        }
        ptr = nullptr;
        capacity = 0;
    }
    template<typename T>
    T* as() const
    {
        return reinterpret_cast<T*>(ptr);
    }
};

} // namespace vkrt
