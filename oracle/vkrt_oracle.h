/*
 * vkrt_oracle.h — CPU ORACLE for the vkrt ray-tracing hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-CPU restatement of the reference's GLSL ray-tracing pipeline
 *   /root/reference/shaders/shader.rgen:30-77        (ray generation + bounce loop)
 *   /root/reference/shaders/shader.rchit:78-172      (closest-hit shading)
 *   /root/reference/shaders/shader.rmiss:15-18       (sky)
 *   /root/reference/shaders/shader_shadow.rmiss:6-9  (shadow miss)
 * plus the Vulkan traceRayEXT semantics the driver supplies (closest / any hit over one BLAS
 * instanced through a TLAS; /root/reference/src/Raytracer.cpp:1027-1283).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libvkrt_b200.so) never links, loads or calls it.
 *
 * PARITY UNPINNED: the reference ships no golden vectors for this path and its traversal lives in
 * the Vulkan driver, which cannot run on this box (no graphics stack).  The oracle is pinned only
 * by (1) brute-force == BVH traversal, (2) analytic scenes with hand-computed pixels and (3) the
 * reference's own Camera.cpp / Model.cpp compiled from /root/reference (oracle/_ref).
 */
#ifndef VKRT_ORACLE_H
#define VKRT_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OrcScene OrcScene;

/* One triangle geometry of the single BLAS (mirrors VkAccelerationStructureBuildRangeInfoKHR,
 * /root/reference/src/Raytracer.cpp:1077-1082). */
typedef struct OrcGeometryRange
{
    uint32_t primitiveCount;  /* triangles */
    uint32_t maxVertex;       /* unused by the oracle, kept for layout parity */
    uint64_t indexByteOffset; /* byte offset of the first index of this geometry in `indices` */
} OrcGeometryRange;

/* One row of the reference's SubmeshInfo / GLSL MaterialInfo table
 * (/root/reference/src/Raytracer.cpp:33-39, shaders/shader.rchit:43-49). */
typedef struct OrcSubmeshInfo
{
    int32_t baseColorTextureIndex;
    int32_t metallicRoughnessTextureIndex;
    int32_t normalTextureIndex;
    int32_t indexBufferOffset; /* in TRIANGLES */
} OrcSubmeshInfo;

typedef struct OrcImage
{
    uint32_t width, height;
    const uint8_t* rgba8; /* width*height*4, row-major, R8G8B8A8_UNORM (Raytracer.cpp:467) */
} OrcImage;

/* Per-frame constants, same 256 B layout as UniformBufferInfo (Raytracer.cpp:15-24). */
typedef struct OrcUniforms
{
    float viewInverse[16]; /* column-major */
    float projInverse[16]; /* column-major */
    float position[4], right[4], up[4], forward[4];
    float lightPositions[4][4];
} OrcUniforms;

typedef struct OrcStats
{
    uint64_t raysPrimary, raysShadow, raysReflection;
    uint64_t nodesVisited, trianglesTested;
    double seconds; /* wall time of the last orc_render */
} OrcStats;

typedef struct OrcRay
{
    float origin[3];
    float tmin;
    float direction[3];
    float tmax;
} OrcRay;

typedef struct OrcHit
{
    float t;            /* world-parametrised distance; undefined on miss */
    float b1, b2;       /* barycentric weights of vertex 1 and 2 (Vulkan hitAttribute) */
    uint32_t instance;  /* instance index in the TLAS, 0xFFFFFFFF on miss */
    uint32_t geometry;  /* gl_GeometryIndexEXT */
    uint32_t primitive; /* gl_PrimitiveID */
} OrcHit;

enum
{
    ORC_MODE_BRUTE = 0, /* O(N) loop over all triangles — ground truth */
    ORC_MODE_BVH = 1    /* binary SAH BVH — must agree with ORC_MODE_BRUTE */
};

/*
 * vertices      : vertexCount × 16 floats (position, normal, uv, tangent as vec4 — Model::Vertex)
 * indices       : global (already rebased) u32 indices, Raytracer.cpp:670-694
 * instances     : instanceCount × 12 floats, row-major 3×4 object→world (VkTransformMatrixKHR)
 * All inputs are copied.
 */
OrcScene* orc_scene_create(const float* vertices, uint64_t vertexCount, const uint32_t* indices, uint64_t indexCount,
                           const OrcGeometryRange* ranges, const OrcSubmeshInfo* rows, uint32_t geometryCount,
                           const OrcImage* images, uint32_t imageCount, const float* instances,
                           uint32_t instanceCount);
void orc_scene_destroy(OrcScene* scene);
/* 8-bit VkAccelerationStructureInstanceKHR::mask per instance (default 0xFF, Raytracer.cpp:1174); an instance whose
 * mask & 0xFF (the cullMask of every traceRayEXT in the shaders) is zero is invisible to all rays. */
void orc_scene_set_instance_masks(OrcScene* scene, const uint8_t* masks, uint32_t count);

/* Replace vertex positions/attributes in place (config 5 rebuild) and rebuild the BVH. */
void orc_scene_update_vertices(OrcScene* scene, const float* vertices, uint64_t vertexCount);

/* Seconds spent building the oracle BVH in orc_scene_create / update. */
double orc_scene_build_seconds(const OrcScene* scene);

/*
 * Render the pixel rectangle [x0,x1)×[y0,y1) of a width×height frame, visiting only pixels with
 * ((x - x0) % strideX == 0 && (y - y0) % strideY == 0).  Outputs are full-frame arrays (may be NULL);
 * unvisited pixels are left untouched.
 *   rgb32f     : width*height*3 floats (finalHitValue averaged over spp)
 *   bgra8      : width*height*4 bytes  (B,G,R,A — VK_FORMAT_B8G8R8A8_UNORM, VulkanUtils.hpp:24)
 *   triangleId : width*height u32, (instance<<24) + indexBufferOffset + primitiveID of the primary hit of
 *                sample 0, 0xFFFFFFFF on miss
 *   shadowMask : width*height bytes, bits0-3 light i shadowed, bits4-7 shadow ray i cast (primary hit, sample 0)
 */
void orc_render(OrcScene* scene, const OrcUniforms* uniforms, uint32_t width, uint32_t height, uint32_t spp,
                int mode, int threads, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t strideX,
                uint32_t strideY, float* rgb32f, uint8_t* bgra8, uint32_t* triangleId, uint8_t* shadowMask,
                OrcStats* stats);

/* Ray-level entry: anyHit=0 → closest hit, anyHit=1 → terminate on first hit (hit.instance != ~0 if occluded). */
void orc_trace_rays(OrcScene* scene, const OrcRay* rays, uint64_t count, int anyHit, int mode, int threads,
                    OrcHit* hits, OrcStats* stats);

/* Bilinear REPEAT LOD-0 sample of image `index` (rchit:107,157,162 + sampler Raytracer.cpp:436-458). */
void orc_sample_texture(const OrcScene* scene, uint32_t index, float u, float v, float rgba[4]);

/* Mip chain as Raytracer::createMipmaps builds it (src/Raytracer.cpp:572-640): level k = linear blit of level k-1. */
void orc_scene_generate_mips(OrcScene* scene);
/* 0 = every texture() samples level 0 (the reference's ray-tracing path, default); 1 = ray-cone level of detail
 * (needs orc_scene_generate_mips; not in the reference: checks the product's optional VKRT_TEXTURE_LOD_RAY_CONES). */
void orc_scene_set_texture_lod(OrcScene* scene, int mode);
uint32_t orc_texture_level_count(const OrcScene* scene, uint32_t index);
const uint8_t* orc_texture_level(const OrcScene* scene, uint32_t index, uint32_t level, uint32_t* width, uint32_t* height);
void orc_sample_texture_lod(const OrcScene* scene, uint32_t index, float u, float v, float footprint2, float rgba[4]);

/* The reference's rasteriser (src/Rasterizer.cpp:100-170, shaders/shader.vert, shaders/shader.frag): one frame through
 * the column-major `wvp` matrix; needs orc_scene_generate_mips.  bgra8: width*height*4 (resolved image); triangleId:
 * width*height u32, triangle (draw order) that owns sample 0 of the pixel, 0xFFFFFFFF where the clear colour shows. */
void orc_rasterize(OrcScene* scene, const float* wvp, uint32_t width, uint32_t height, int threads, uint8_t* bgra8,
                   uint32_t* triangleId);

/* Halton(2,3) sample s in [0,1)^2 used instead of +0.5 when spp > 1. */
void orc_halton23(uint32_t s, float out[2]);

/* Primary ray for pixel (x,y) with in-pixel offset (jx,jy) — shader.rgen:32-38. */
void orc_primary_ray(const OrcUniforms* uniforms, uint32_t width, uint32_t height, uint32_t x, uint32_t y, float jx,
                     float jy, OrcRay* ray);

/* Inverse of a row-major 3×4 affine transform (double cofactors, rounded to float). */
void orc_affine_inverse(const float m[12], float inv[12]);

#ifdef __cplusplus
}
#endif
#endif
