/*
 * vkrt.h — C ABI of libvkrt_b200.so: the B200-native (sm_100a CUDA) replacement for the Vulkan
 * KHR ray-tracing calls that jparimaa/vkrt's `Raytracer` class makes.
 *
 * The reference has no FFI layer; the boundary sits where `Raytracer` crosses into the Vulkan
 * driver.  Each entry point below names the reference call site it replaces (paths relative to the
 * reference checkout).  Plain pointers and sizes only; no C++ or torch types.
 *
 * Conventions
 *   - every function returns 0 on success, a VkrtResult error code otherwise; the library never
 *     aborts (the reference's VK_CHECK -> abort() convention, src/VulkanUtils.hpp:28-37, is kept by
 *     the host-side VKRT_CHECK macro in vkrt_b200/host/Utils.hpp);
 *   - host input buffers are copied during the call and may be freed on return;
 *   - a context is single-threaded, owns one CUDA device and one stream; builds are complete on
 *     return; vkrt_trace is stream-ordered; vkrt_frame_read synchronises;
 *   - there is NO CPU fallback: without a CUDA device vkrt_context_create fails.
 */
#ifndef VKRT_H
#define VKRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define VKRT_API __attribute__((visibility("default")))
#else
#define VKRT_API
#endif

typedef struct VkrtContext VkrtContext;
typedef struct VkrtBlas VkrtBlas;
typedef struct VkrtTlas VkrtTlas;
typedef struct VkrtFrame VkrtFrame;

typedef enum VkrtResult
{
    VKRT_SUCCESS = 0,
    VKRT_ERROR_INVALID_ARGUMENT = 1,
    VKRT_ERROR_NO_DEVICE = 2,
    VKRT_ERROR_CUDA = 3,
    VKRT_ERROR_OUT_OF_MEMORY = 4,
    VKRT_ERROR_NCCL = 5,
    VKRT_ERROR_NOT_READY = 6,
    VKRT_ERROR_TRAVERSAL_OVERFLOW = 7 /* a ray needed more traversal-stack entries than the library keeps (48); reported by the
                                         next synchronising call (vkrt_frame_read, vkrt_context_synchronize, vkrt_trace_rays).  The
                                         builders refuse hierarchies deep enough to cause it, so this signals a library defect. */
} VkrtResult;

/* ------------------------------------------------------------------------------------------------
 * Context — replaces the device half of `Context` (src/Context.cpp:280-366 createDevice,
 * :368-380 queue/command pool).  One CUDA device, one stream.
 * ---------------------------------------------------------------------------------------------- */
VKRT_API int vkrt_context_create(int deviceOrdinal, VkrtContext** out);
VKRT_API void vkrt_context_destroy(VkrtContext* ctx);
/* Text of the last error on this context (or of the last failed vkrt_context_create if ctx==NULL). */
VKRT_API const char* vkrt_last_error(const VkrtContext* ctx);
/* Library/ABI version, `major*10000 + minor*100 + patch`. */
VKRT_API int vkrt_version(void);
/* Blocks until all work queued on the context's stream is done (vkQueueWaitIdle, src/VulkanUtils.cpp:202). */
VKRT_API int vkrt_context_synchronize(VkrtContext* ctx);
/* The context's cudaStream_t, for callers that time with CUDA events on the launching stream. */
VKRT_API void* vkrt_context_stream(VkrtContext* ctx);

/* ------------------------------------------------------------------------------------------------
 * Bottom-level acceleration structure — replaces Raytracer::createBLAS (src/Raytracer.cpp:1027-1157):
 * one BLAS with `geometryCount` triangle geometries that share one vertex buffer (positions are the
 * first 3 floats of each `vertexStride`-byte vertex, VK_FORMAT_R32G32B32_SFLOAT, :1060-1062) and one
 * u32 index buffer (:1064).  All geometries are opaque, no culling (:1072, :1176).
 * ---------------------------------------------------------------------------------------------- */
typedef struct VkrtGeometryRange
{
    uint32_t primitiveCount;  /* VkAccelerationStructureBuildRangeInfoKHR::primitiveCount  (:1078) */
    uint32_t maxVertex;       /* VkAccelerationStructureGeometryTrianglesDataKHR::maxVertex (:1063) */
    uint64_t indexByteOffset; /* ...BuildRangeInfoKHR::primitiveOffset, bytes into the index buffer (:1079) */
} VkrtGeometryRange;

typedef enum VkrtBuildFlags
{
    VKRT_BUILD_DEFAULT = 0,
    VKRT_BUILD_PREFER_FAST_BUILD = 1u << 0, /* LBVH + greedy collapse instead of PLOC + cost-optimal collapse: ~2x faster build, ~25 % slower rays */
    VKRT_BUILD_INPUT_ON_DEVICE = 1u << 1,   /* `vertices` and `indices` are CUDA device pointers */
    VKRT_BUILD_KEEP_HOST_INPUT = 1u << 2,   /* reserved */
    /* VK_BUILD_ACCELERATION_STRUCTURE_PREFER_FAST_TRACE_BIT_KHR: after PLOC, parallel reinsertion passes (Meister & Bittner 2018)
     * move subtrees to where they cost the least surface area, large nodes first; several times the build time for
     * fewer node visits per ray.  For hierarchies that are built once and traced many times.  Ignored together with
     * VKRT_BUILD_PREFER_FAST_BUILD; if the passes leave the hierarchy deeper than the traversal stack it is built again
     * without them. */
    VKRT_BUILD_PREFER_FAST_TRACE = 1u << 3
} VkrtBuildFlags;

typedef struct VkrtBuildSizes
{
    uint64_t accelerationStructureSize; /* bytes of compressed nodes + packed triangles */
    uint64_t buildScratchSize;          /* bytes of builder scratch */
    uint64_t geometryCopySize;          /* bytes of the device copy of vertex+index data kept for shading */
} VkrtBuildSizes;

/* Mirror of vkGetAccelerationStructureBuildSizesKHR (src/Raytracer.cpp:1106): upper bounds. */
VKRT_API int vkrt_blas_sizes(VkrtContext* ctx, const VkrtGeometryRange* ranges, uint32_t geometryCount,
                             uint64_t vertexBytes, uint64_t indexCount, VkrtBuildSizes* out);

/* `vertexStride` >= 12 and a multiple of 4 is enough to build and to trace rays (vkrt_trace_rays).  Shading (vkrt_trace)
 * and vkrt_rasterize read the reference's 64-byte Model::Vertex (position, normal, uv, tangent as 4 x vec4,
 * src/Model.hpp:13-19) with 16-byte loads: they need vertexStride >= 64 and a multiple of 16 and fail with
 * VKRT_ERROR_INVALID_ARGUMENT otherwise.  The build fails if the hierarchy comes out deeper than the traversal stack (48 levels). */
VKRT_API int vkrt_blas_build(VkrtContext* ctx, const void* vertices, size_t vertexBytes, uint32_t vertexStride,
                             const uint32_t* indices, size_t indexCount, const VkrtGeometryRange* ranges,
                             uint32_t geometryCount, uint32_t flags, VkrtBlas** out);
/* Re-runs the build over new vertex data of the same size, reusing all storage (config 5: per-frame rebuild). */
VKRT_API int vkrt_blas_rebuild(VkrtContext* ctx, VkrtBlas* blas, const void* vertices, size_t vertexBytes,
                               uint32_t flags);
VKRT_API void vkrt_blas_destroy(VkrtContext* ctx, VkrtBlas* blas);
/* Handle to put in VkrtInstance::accelerationStructureReference
 * (vkGetAccelerationStructureDeviceAddressKHR, src/Raytracer.cpp:1141-1146). */
VKRT_API uint64_t vkrt_blas_reference(const VkrtBlas* blas);

typedef struct VkrtBlasInfo
{
    uint64_t triangleCount;
    uint64_t nodeCount;      /* compressed 8-wide nodes, 80 B each */
    uint64_t nodeBytes;
    uint64_t triangleBytes;  /* 48 B per packed triangle */
    float boundsMin[3];
    float boundsMax[3];
    float sahCost;           /* SAH cost of the wide BVH (node cost 1, triangle cost 1, root-area normalised) */
    float buildMilliseconds; /* device time of the last build, CUDA events */
} VkrtBlasInfo;
VKRT_API int vkrt_blas_info(VkrtContext* ctx, const VkrtBlas* blas, VkrtBlasInfo* out);

/* Debug/readback of the built structure for invariant tests (every primitive referenced once, quantised
 * child boxes contain the exact boxes).  `nodes` receives nodeCount*80 bytes, `triangles` triangleCount*48. */
VKRT_API int vkrt_blas_read(VkrtContext* ctx, const VkrtBlas* blas, void* nodes, size_t nodeBytes, void* triangles,
                            size_t triangleBytes);

/* ------------------------------------------------------------------------------------------------
 * Top-level acceleration structure — replaces Raytracer::createTLAS (src/Raytracer.cpp:1159-1283).
 * VkrtInstance is byte-compatible with the 64-byte VkAccelerationStructureInstanceKHR (:1171-1177).
 * ---------------------------------------------------------------------------------------------- */
typedef struct VkrtInstance
{
    float transform[12];                     /* row-major 3x4 object->world (VkTransformMatrixKHR) */
    uint32_t instanceCustomIndex : 24;
    uint32_t mask : 8;
    uint32_t instanceShaderBindingTableRecordOffset : 24;
    uint32_t flags : 8;
    uint64_t accelerationStructureReference; /* vkrt_blas_reference() */
} VkrtInstance;

/* Builds the top level on the device: world boxes of the instances (k_instance_bounds), then the same builder as for a
 * BLAS over those boxes, one instance per leaf of an 8-wide hierarchy; 1 .. 2^24 instances.  Traversal switches to an
 * instance's BLAS with the ray transformed by the inverse of `transform` (origin and direction, t shared) and enters an
 * instance only when `mask & 0xFF` (the cullMask of every traceRayEXT in the reference's shaders) is non-zero.  `flags`
 * and `instanceCustomIndex` are recorded; nothing in this pipeline culls by facing or reads the custom index.
 * A TLAS is tied to the BLAS builds it saw: after vkrt_blas_rebuild / vkrt_blas_destroy of a referenced BLAS, tracing
 * it returns VKRT_ERROR_NOT_READY until the TLAS is built again (the reference rebuilds both in Raytracer's constructor). */
VKRT_API int vkrt_tlas_build(VkrtContext* ctx, const VkrtInstance* instances, uint32_t instanceCount, VkrtTlas** out);
VKRT_API void vkrt_tlas_destroy(VkrtContext* ctx, VkrtTlas* tlas);

typedef struct VkrtTlasInfo
{
    uint32_t instanceCount;
    uint32_t nodeCount; /* 80-byte top-level nodes; 0 for a single instance (no top level is needed) */
    uint32_t levels;    /* depth of the top-level hierarchy */
    uint32_t reserved;
    float boundsMin[3];
    float boundsMax[3];
    float sahCost;
    float buildMilliseconds; /* device time of the top-level build, CUDA events */
} VkrtTlasInfo;
VKRT_API int vkrt_tlas_info(VkrtContext* ctx, const VkrtTlas* tlas, VkrtTlasInfo* out);
/* Debug/readback for invariant tests: `nodes` receives nodeCount*80 bytes (Node8::reserved = mask of leaf slots, leaf slot s
 * of a node holds instance leafInstances[primBase + popcount(reserved & ((1<<s)-1))]), `leafInstances` instanceCount u32. */
VKRT_API int vkrt_tlas_read(VkrtContext* ctx, const VkrtTlas* tlas, void* nodes, size_t nodeBytes, uint32_t* leafInstances,
                            size_t leafBytes);

/* ------------------------------------------------------------------------------------------------
 * Shading resources
 * ---------------------------------------------------------------------------------------------- */
/* One row per geometry: the reference's SubmeshInfo / GLSL MaterialInfo
 * (src/Raytracer.cpp:33-39, :1412-1427; shaders/shader.rchit:43-49). */
typedef struct VkrtSubmeshInfo
{
    int32_t baseColorTextureIndex;
    int32_t metallicRoughnessTextureIndex;
    int32_t normalTextureIndex;
    int32_t indexBufferOffset; /* cumulative TRIANGLE count */
} VkrtSubmeshInfo;
/* Replaces createMaterialIndexBuffer + updateMaterialIndexDescriptorSet (src/Raytracer.cpp:1004-1012, :1386-1442). */
VKRT_API int vkrt_materials_set(VkrtContext* ctx, const VkrtSubmeshInfo* rows, uint32_t count);

typedef struct VkrtImage
{
    uint32_t width, height;
    const uint8_t* rgba8; /* R8G8B8A8_UNORM, tightly packed rows (src/Raytracer.cpp:467) */
} VkrtImage;
/* Replaces createTextures + createSampler + updateTexturesDescriptorSets (src/Raytracer.cpp:436-570, :1444-1467):
 * bindless array of RGBA8 images sampled LINEAR / REPEAT at LOD 0 (shaders/shader.rchit:76,107,157,162). */
VKRT_API int vkrt_textures_set(VkrtContext* ctx, const VkrtImage* images, uint32_t count);
/* Replaces Raytracer::createMipmaps (src/Raytracer.cpp:572-640): the full mip chain of every image,
 * floor(log2(max(w,h))) + 1 levels (:467-468), level k = max(w>>k,1) x max(h>>k,1), each level a
 * vkCmdBlitImage(VK_FILTER_LINEAR) of the previous one (:610-625).  Built on the device; complete on return.
 * vkrt_textures_set discards the chain. */
VKRT_API int vkrt_textures_generate_mips(VkrtContext* ctx);
/* Extent of `level` of image `image` and its level count (1 until the chain has been generated). */
VKRT_API int vkrt_texture_info(VkrtContext* ctx, uint32_t image, uint32_t level, uint32_t* width, uint32_t* height,
                               uint32_t* levels);
/* Copies one level (tightly packed RGBA8 rows) to host memory; for the parity tests of the chain. */
VKRT_API int vkrt_texture_read(VkrtContext* ctx, uint32_t image, uint32_t level, void* dst, size_t bytes);

typedef enum VkrtTextureLod
{
    VKRT_TEXTURE_LOD_ZERO = 0,     /* every texture() samples level 0: what the reference's ray-tracing shaders get
                                      (no derivatives in rchit; sampler src/Raytracer.cpp:436-458) - default, parity mode */
    VKRT_TEXTURE_LOD_RAY_CONES = 1 /* trilinear at the ray-cone level of detail (Ray Tracing Gems ch. 20): cone width =
                                      pixel spread angle x distance travelled, per-triangle texel density, |n.d| term */
} VkrtTextureLod;
/* Texture level-of-detail mode of the shading kernels (no counterpart in the reference: its mip chain is built but
 * never sampled on the ray-tracing path).  RAY_CONES needs vkrt_textures_generate_mips first. */
VKRT_API int vkrt_trace_set_texture_lod(VkrtContext* ctx, int mode);

/* ------------------------------------------------------------------------------------------------
 * Frames and tracing
 * ---------------------------------------------------------------------------------------------- */
/* 256-byte per-frame constants, identical layout to UniformBufferInfo (src/Raytracer.cpp:15-24) and the GLSL
 * CommonUniformBuffer (shaders/shader.rgen:16-26). */
typedef struct VkrtUniforms
{
    float viewInverse[16]; /* column-major */
    float projInverse[16]; /* column-major */
    float position[4];
    float right[4];
    float up[4];
    float forward[4];
    float lightPositions[4][4];
} VkrtUniforms;

/* Offscreen colour target + AOVs; replaces createColorImage (src/Raytracer.cpp:326-392).  With a communicator
 * attached (vkrt_comm_init) the frame is split into 64x64-pixel tiles, tile i owned by rank i % size. */
VKRT_API int vkrt_frame_create(VkrtContext* ctx, uint32_t width, uint32_t height, VkrtFrame** out);
VKRT_API void vkrt_frame_destroy(VkrtContext* ctx, VkrtFrame* frame);

typedef enum VkrtTraceFlags
{
    VKRT_TRACE_DEFAULT = 0,
    VKRT_TRACE_NO_AOV = 1u << 0 /* skip the triangle-id / shadow-mask / float RGB outputs (timed runs) */
} VkrtTraceFlags;

/* Replaces the per-frame UBO memcpy (src/Raytracer.cpp:218-235) + vkCmdTraceRaysKHR(W,H,1) (:157): runs
 * shader.rgen/.rchit/.rmiss/shader_shadow.rmiss semantics for every owned pixel, `spp` samples per pixel
 * (spp==1: pixel centre +0.5 as shaders/shader.rgen:32; spp>1: Halton(2,3) offsets), and resolves to B8G8R8A8.
 * Asynchronous on the context stream. */
VKRT_API int vkrt_trace(VkrtContext* ctx, const VkrtTlas* tlas, const VkrtUniforms* uniforms, uint32_t spp,
                        uint32_t frameIndex, uint32_t flags, VkrtFrame* frame);

typedef enum VkrtAov
{
    VKRT_AOV_BGRA8 = 0,           /* 4 B/pixel, B,G,R,A (VK_FORMAT_B8G8R8A8_UNORM, src/VulkanUtils.hpp:24) */
    VKRT_AOV_RGB32F = 1,          /* 12 B/pixel, finalHitValue averaged over spp */
    VKRT_AOV_TRIANGLE_ID_U32 = 2, /* (instance<<24) + indexBufferOffset + gl_PrimitiveID of the primary hit, ~0 = miss */
    VKRT_AOV_SHADOW_MASK_U8 = 3   /* bits 0-3: light i shadowed, bits 4-7: shadow ray i cast (primary hit, sample 0) */
} VkrtAov;
/* Copies one output plane to host memory (`pitch` = bytes per row, 0 = tight). Replaces the copy to the
 * swapchain image (src/Raytracer.cpp:175-185) in the headless mode.  Synchronises the stream.  With a
 * communicator, valid on the root rank after vkrt_tiles_gather (BGRA8) and on every rank for owned tiles. */
VKRT_API int vkrt_frame_read(VkrtContext* ctx, VkrtFrame* frame, int aov, void* dst, size_t pitch);
/* Presentation ring, the headless form of the reference's swapchain (acquire / submit / present with 3 images in
 * flight, src/Context.cpp:141-180; image copy src/Raytracer.cpp:175-185).  vkrt_frame_present queues a copy of the
 * BGRA8 image into `dst` (`pitch` = bytes per row, 0 = tight) on the library's copy stream, ordered after everything
 * queued on the context so far, and returns without waiting; `*ticket` identifies the copy.  `dst` should be page-locked
 * (vkrt_host_alloc) or the copy degrades to a synchronous one.  The next vkrt_trace / vkrt_tiles_gather into the same
 * frame may be queued at once: only its final image write waits (on the device) for outstanding copies, so the copy
 * overlaps the next frame's traversal.  vkrt_frame_present_wait blocks the host until the copy `ticket` has landed
 * (the wait a swapchain image needs before it is reused).  At most 8 copies per frame are kept in flight. */
VKRT_API int vkrt_frame_present(VkrtContext* ctx, VkrtFrame* frame, void* dst, size_t pitch, uint64_t* ticket);
VKRT_API int vkrt_frame_present_wait(VkrtContext* ctx, VkrtFrame* frame, uint64_t ticket);
/* Page-locked host memory for vkrt_frame_present / vkrt_frame_read destinations. */
VKRT_API int vkrt_host_alloc(VkrtContext* ctx, size_t bytes, void** out);
VKRT_API void vkrt_host_free(VkrtContext* ctx, void* ptr);
/* Device pointer of the assembled BGRA8 image (width*height*4), valid until the frame is destroyed. */
VKRT_API void* vkrt_frame_device_bgra8(VkrtFrame* frame);

/* ------------------------------------------------------------------------------------------------
 * Rasteriser — the reference's second renderer, `Rasterizer` (src/Rasterizer.cpp; `main` instantiates the ray tracer,
 * src/main.cpp:8).  One call replaces the render pass of Rasterizer::render (:119-147): clear to (0,0,0.2,1) / depth 1,
 * one indexed draw per geometry of `geometry` with shaders/shader.vert + shader.frag (base-colour texture with its mip
 * chain, alpha test at 0.1), back-face culling with counter-clockwise front faces, depth test LESS on D24, 8x MSAA
 * without sample shading, resolve into the frame's B8G8R8A8 image (pipeline state :779-829).
 * `geometry` supplies the vertex and index data (positions + uv of the 64-byte Model::Vertex, draw order = geometry order,
 * :1054-1075); `wvp` is the column-major proj * view * world matrix of Rasterizer::update (:189-191).  Materials, textures
 * and the mip chain must have been set.  Stream-ordered like vkrt_trace; with AOVs, VKRT_AOV_TRIANGLE_ID_U32 holds the
 * triangle (cumulative index over the geometries) seen by sample 0 of every pixel, ~0 where the clear colour shows.
 * With a communicator attached the frame is split like vkrt_trace's: each rank draws the pixels of the 64x64 tiles it owns
 * (every rank runs the vertex stage over all triangles), and vkrt_tiles_gather assembles the image on the root.
 * The ImGui overlay (:149-159) is not drawn.
 * ---------------------------------------------------------------------------------------------- */
VKRT_API int vkrt_rasterize(VkrtContext* ctx, const VkrtBlas* geometry, const float* wvp, uint32_t flags, VkrtFrame* frame);

/* Scheduling knob of the persistent trace kernels (no counterpart in the reference; results do not depend on it).
 * lanes == 0 (default): a warp takes 32 new rays once all of its rays have finished - fastest on coherent scenes.
 * lanes in 1..32: lanes whose ray has finished are re-armed as soon as fewer than `lanes` lanes of the warp are still
 * traversing - pays on incoherent scenes (the config-5 triangle soup). */
VKRT_API int vkrt_trace_set_lane_refill(VkrtContext* ctx, uint32_t lanes);

/* Ray-level entry (ray-query style; used by the parity tests and per-ray-type benchmarks).  `rays` are
 * 32-byte records {origin[3], tmin, direction[3], tmax}; `hits` 24-byte records {t, b1, b2, instance, geometry,
 * primitive}; instance == 0xFFFFFFFF on miss.  anyHit != 0 → gl_RayFlagsTerminateOnFirstHitEXT semantics
 * (shaders/shader.rchit:114-117): only hit/miss is meaningful.  Host pointers. */
typedef struct VkrtRay
{
    float origin[3];
    float tmin;
    float direction[3];
    float tmax;
} VkrtRay;
typedef struct VkrtHit
{
    float t, b1, b2;
    uint32_t instance, geometry, primitive;
} VkrtHit;
VKRT_API int vkrt_trace_rays(VkrtContext* ctx, const VkrtTlas* tlas, const VkrtRay* rays, uint64_t count, int anyHit,
                             VkrtHit* hits);

/* vkrt_trace_rays without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same records and results as vkrt_trace_rays, bit for bit.  The staged rays, the hits, the ray-pool
 * cursor and the overflow flag live in ONE file-scope device array of the library (about 3.5 MiB: 256 bytes +
 * VKRT_STATIC_MAX_RAYS x (32 + 24) bytes), which every GPU that loads the library holds until the process exits; the
 * call makes no cudaMalloc / cudaFree / cudaMallocHost and no allocating library call.  `rays` and `hits` are the
 * caller's buffers and may be host or device pointers (copied with cudaMemcpyDefault on the context's stream); the
 * call returns after the stream has drained.  count > VKRT_STATIC_MAX_RAYS → VKRT_ERROR_INVALID_ARGUMENT ("... at most
 * 65536 rays per call"), nothing is launched.  All calls on one GPU share the array and it keeps the last call's
 * contents (the call clears the cursor and the flag itself): two calls that overlap on the same GPU - from two
 * threads or two contexts - race; serialise them.  The first call on each device resolves the array's device address
 * (cudaGetSymbolAddress); make that call outside stream capture. */
#define VKRT_STATIC_MAX_RAYS 65536u
VKRT_API int vkrt_trace_rays_static(VkrtContext* ctx, const VkrtTlas* tlas, const VkrtRay* rays, uint32_t count, int anyHit,
                                    VkrtHit* hits);

/* ------------------------------------------------------------------------------------------------
 * Statistics for the roofline: counted since the last vkrt_stats_reset.
 * ---------------------------------------------------------------------------------------------- */
typedef struct VkrtStats
{
    uint64_t raysPrimary, raysShadow, raysReflection;
    uint64_t nodesVisited, trianglesTested; /* closest-hit + any-hit; only counted when enabled with vkrt_stats_enable */
    uint64_t nodesVisitedAny, trianglesTestedAny; /* the any-hit (shadow ray) share of the two totals above */
    uint64_t kernelLaunches;                /* kernels launched by this library */
    float msRaygen, msTraceClosest, msShade, msTraceAny, msResolve, msTotal; /* first wave of the last vkrt_trace, if timing enabled */
    uint32_t samplesPerWave; /* samples per pixel that went through the pipeline together in the last vkrt_trace */
    uint32_t traversalOverflows; /* times VKRT_ERROR_TRAVERSAL_OVERFLOW was raised since the last reset */
    /* the ten stages of the first wave one by one (same events as the sums above): raygen, closest (primary), surface,
     * any (shadow), resolve, closest (reflection), surface, any, resolve, sample resolve */
    float msStage[10];
    uint64_t raysShadowBounce; /* the part of raysShadow cast from reflection hits (the second k_trace_any launch of a wave) */
} VkrtStats;
VKRT_API int vkrt_stats_enable(VkrtContext* ctx, int countTraversal, int timeStages);
VKRT_API int vkrt_stats_reset(VkrtContext* ctx);
VKRT_API int vkrt_stats_get(VkrtContext* ctx, VkrtStats* out);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU: one context per GPU/process, NCCL over NVLink.  The scene is replicated; frames are
 * partitioned into interleaved 64x64 tiles and gathered to `root`.  (No counterpart in the reference.)
 * ---------------------------------------------------------------------------------------------- */
#define VKRT_TILE_SIZE 64
#define VKRT_UNIQUE_ID_BYTES 128
/* Rank 0 creates the id; the caller ships the 128 bytes to the other ranks (MPI, torch.distributed store, file). */
VKRT_API int vkrt_comm_unique_id(uint8_t id[VKRT_UNIQUE_ID_BYTES]);
VKRT_API int vkrt_comm_init(VkrtContext* ctx, const uint8_t id[VKRT_UNIQUE_ID_BYTES], int rank, int size);
/* Tile partition without NCCL (tests, emulation): subsequent frames own tiles with index % size == rank. */
VKRT_API int vkrt_comm_emulate(VkrtContext* ctx, int rank, int size);
VKRT_API int vkrt_comm_destroy(VkrtContext* ctx);
/* FNV-1a (64-bit) of a host buffer: the frame checksum bench.py prints, so that runs on 1, 2, 4 and 8 GPUs can be seen to
 * produce the same image.  Plain host code. */
VKRT_API uint64_t vkrt_hash_fnv1a64(const void* data, size_t bytes);
/* Owner rank of the tile containing pixel (x,y) for a frame of this width. */
VKRT_API int vkrt_tile_owner(uint32_t width, uint32_t x, uint32_t y, int size);
/* Scene ingest for several ranks (the reference decodes and uploads every image in Raytracer::createTextures,
 * src/Raytracer.cpp:460-570, once per process): `root` has called vkrt_textures_set; every other rank calls this INSTEAD
 * of vkrt_textures_set and receives the device-resident texel pool over NVLink (one ncclBroadcast), so the image files
 * are decoded once per box, not once per GPU.  Collective: every rank of the communicator must call it.  Complete on
 * return; the mip chain is not transferred (vkrt_textures_generate_mips rebuilds it in 0.2 ms). */
VKRT_API int vkrt_textures_broadcast(VkrtContext* ctx, int root);
/* Gathers the BGRA8 tiles of all ranks into root's frame: a grouped NCCL send/recv exchange plus one scatter kernel on the
 * library's communication stream, ordered after everything queued on the context so far.  Returns without waiting; the
 * next vkrt_trace into the same frame may be queued at once (only its final image write waits for the gather, so the
 * exchange overlaps the next frame's traversal).  vkrt_frame_read, vkrt_frame_present and vkrt_context_synchronize wait
 * for it; vkrt_tiles_gather_wait makes the context stream wait for it explicitly (stream-ordered, for callers that time
 * with events on vkrt_context_stream). */
VKRT_API int vkrt_tiles_gather(VkrtContext* ctx, VkrtFrame* frame, int root);
VKRT_API int vkrt_tiles_gather_wait(VkrtContext* ctx, VkrtFrame* frame);

#ifdef __cplusplus
}
#endif
#endif /* VKRT_H */
