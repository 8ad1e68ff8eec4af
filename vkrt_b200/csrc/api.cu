// api.cu — the extern "C" entry points declared in include/vkrt.h.
#include "context.hpp"
#include "device_math.cuh"
#include "traverse.cuh"

#include <algorithm>
#include <cstring>
#include <new>

#include <nvtx3/nvToolsExt.h>

using namespace vkrt;

namespace
{
// NVTX ranges named after the reference's debug labels (DebugMarker::beginLabel, src/DebugMarker.cpp; "Render" at
// src/Raytracer.cpp:151).  Header-only NVTX v3: a no-op unless a profiler is attached.
struct NvtxRange
{
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

std::string g_createError;

Context* C(VkrtContext* c) { return static_cast<Context*>(c); }
const Context* C(const VkrtContext* c) { return static_cast<const Context*>(c); }
Blas* B(VkrtBlas* b) { return static_cast<Blas*>(b); }
const Blas* B(const VkrtBlas* b) { return static_cast<const Blas*>(b); }
Tlas* T(VkrtTlas* t) { return static_cast<Tlas*>(t); }
const Tlas* T(const VkrtTlas* t) { return static_cast<const Tlas*>(t); }
Frame* F(VkrtFrame* f) { return static_cast<Frame*>(f); }

int uploadGeometry(Context* ctx, Blas* blas, const void* vertices, size_t vertexBytes, const uint32_t* indices, size_t indexCount,
                   bool onDevice)
{
    const cudaMemcpyKind kind = onDevice ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    VKRT_CUDA_TRY(ctx, blas->vertices.reserve(std::max<size_t>(vertexBytes, 16)));
    if (vertexBytes)
    {
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(blas->vertices.ptr, vertices, vertexBytes, kind, ctx->stream));
    }
    if (indices)
    {
        VKRT_CUDA_TRY(ctx, blas->indices.reserve(std::max<size_t>(indexCount * 4, 16)));
        if (indexCount)
        {
            VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(blas->indices.ptr, indices, indexCount * 4, kind, ctx->stream));
        }
    }
    return VKRT_SUCCESS;
}
} // namespace

namespace vkrt
{
// The device copy of the material table with every texture index made safe to dereference.  The reference clamps only
// the normal and metallic-roughness indices (src/Raytracer.cpp:1424-1426) and binds the base colour unchecked
// (-1 = "material has no base-colour texture" would index textures[-1] in shader.rchit:157, undefined in Vulkan).
// Here indices < 0 or >= the image count sample image 0, the rule the rasteriser path (raster.cu) and the oracle follow.
int syncMaterials(Context* ctx)
{
    if (!ctx->materialsDirty || ctx->materialCount == 0)
    {
        return VKRT_SUCCESS;
    }
    std::vector<VkrtSubmeshInfo> safe = ctx->materialsHost;
    if (ctx->textureCount)
    {
        auto fix = [&](int32_t& index) {
            if (index < 0 || (uint32_t)index >= ctx->textureCount)
            {
                index = 0;
            }
        };
        for (VkrtSubmeshInfo& r : safe)
        {
            fix(r.baseColorTextureIndex), fix(r.metallicRoughnessTextureIndex), fix(r.normalTextureIndex);
        }
    }
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaMemcpy(ctx->materials.ptr, safe.data(), safe.size() * sizeof(VkrtSubmeshInfo), cudaMemcpyHostToDevice));
    ctx->materialsDirty = ctx->textureCount == 0; // checked again once textures exist
    return VKRT_SUCCESS;
}

// A TLAS holds device pointers into its BLASes and their bounds: it is stale once one of them has been rebuilt
// (vkrt_blas_rebuild moves the bounds and may move the node storage) or destroyed.
int checkTlasCurrent(Context* ctx, const Tlas* tlas, const char* who)
{
    for (const auto& ref : tlas->blasRefs)
    {
        if (std::find(ctx->liveBlas.begin(), ctx->liveBlas.end(), ref.first) == ctx->liveBlas.end())
        {
            return ctx->fail(VKRT_ERROR_NOT_READY, std::string(who) + ": the TLAS references a destroyed BLAS");
        }
        if (ref.first->generation != ref.second)
        {
            return ctx->fail(VKRT_ERROR_NOT_READY, std::string(who) + ": a BLAS was rebuilt after this TLAS was built; build the TLAS again");
        }
    }
    return VKRT_SUCCESS;
}

int checkSceneForShading(Context* ctx, const Tlas* tlas, const char* who)
{
    if (int r = checkTlasCurrent(ctx, tlas, who))
    {
        return r;
    }
    if (ctx->materialCount == 0 || ctx->textureCount == 0)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, std::string(who) + ": materials and textures must be set first");
    }
    if (!tlas->shadingLayout)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, std::string(who) + ": shading reads the 64-byte Model::Vertex (position, normal, uv, tangent "
                                                                          "as 4 x vec4): every BLAS needs vertexStride >= 64 and a multiple of 16");
    }
    if (ctx->materialCount < tlas->maxGeometryCount)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, std::string(who) + ": the material table has " + std::to_string(ctx->materialCount) +
                                                          " rows but a BLAS has " + std::to_string(tlas->maxGeometryCount) + " geometries");
    }
    return syncMaterials(ctx);
}

// Device-side error flags ([0] = a traversal stack overflowed and a subtree was dropped).  Read where the host already
// waits for the stream; a set flag is cleared and reported once.
int checkDeviceErrors(Context* ctx, const char* who)
{
    uint32_t flags[4] = {0, 0, 0, 0};
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(flags, ctx->errorFlags.ptr, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (flags[0])
    {
        cudaMemsetAsync(ctx->errorFlags.ptr, 0, sizeof(flags), ctx->stream);
        ctx->stats.traversalOverflows += 1;
        return ctx->fail(VKRT_ERROR_TRAVERSAL_OVERFLOW, std::string(who) + ": a ray needed more than " + std::to_string(kStackSize) +
                                                             " traversal stack entries; the frame is incomplete");
    }
    return VKRT_SUCCESS;
}
} // namespace vkrt

extern "C" {

int vkrt_version(void)
{
    return 300; // 0.3.0: + device-built TLAS hierarchy, instance masks, vkrt_tlas_info / _read, error surfacing
}

int vkrt_context_create(int deviceOrdinal, VkrtContext** out)
{
    if (!out)
    {
        g_createError = "vkrt_context_create: out is NULL";
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
    {
        g_createError = std::string("vkrt_context_create: no CUDA device (") + cudaGetErrorString(e) +
                        "); libvkrt_b200 has no CPU fallback";
        return VKRT_ERROR_NO_DEVICE;
    }
    if (deviceOrdinal < 0 || deviceOrdinal >= count)
    {
        g_createError = "vkrt_context_create: device ordinal out of range";
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    e = cudaSetDevice(deviceOrdinal);
    if (e != cudaSuccess)
    {
        g_createError = std::string("cudaSetDevice: ") + cudaGetErrorString(e);
        return VKRT_ERROR_CUDA;
    }
    Context* ctx = new (std::nothrow) Context();
    if (!ctx)
    {
        return VKRT_ERROR_OUT_OF_MEMORY;
    }
    ctx->device = deviceOrdinal;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, deviceOrdinal) == cudaSuccess)
    {
        ctx->smCount = prop.multiProcessorCount;
    }
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess)
    {
        e = ctx->statCounters.reserve(128);
    }
    if (e == cudaSuccess)
    {
        e = cudaMemsetAsync(ctx->statCounters.ptr, 0, 128, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        e = ctx->errorFlags.reserve(16);
    }
    if (e == cudaSuccess)
    {
        e = cudaMemsetAsync(ctx->errorFlags.ptr, 0, 16, ctx->stream);
    }
    if (e != cudaSuccess)
    {
        g_createError = std::string("vkrt_context_create: ") + cudaGetErrorString(e);
        delete ctx;
        return VKRT_ERROR_CUDA;
    }
    *out = ctx;
    return VKRT_SUCCESS;
}

void vkrt_context_destroy(VkrtContext* c)
{
    if (!c)
    {
        return;
    }
    Context* ctx = C(c);
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    commDestroy(ctx);
    for (void* p : ctx->textureAllocs)
    {
        cudaFree(p);
    }
    ctx->materials.release();
    ctx->textureRecords.release();
    ctx->textureMips.release();
    ctx->mipPool.release();
    ctx->uniforms.release();
    ctx->buildScratch.release();
    ctx->wave.release();
    ctx->raster.release();
    ctx->statCounters.release();
    ctx->errorFlags.release();
    ctx->gatherStaging.release();
    if (ctx->eventsCreated)
    {
        for (auto& ev : ctx->stageEvents)
        {
            cudaEventDestroy(ev);
        }
    }
    if (ctx->copyStream)
    {
        cudaStreamSynchronize(ctx->copyStream);
        cudaStreamDestroy(ctx->copyStream);
    }
    if (ctx->buildReadback)
    {
        cudaFreeHost(ctx->buildReadback);
        ctx->buildReadback = nullptr;
    }
    if (ctx->commStream)
    {
        cudaStreamSynchronize(ctx->commStream);
        cudaStreamDestroy(ctx->commStream);
        cudaEventDestroy(ctx->commReady);
    }
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* vkrt_last_error(const VkrtContext* c)
{
    return c ? C(c)->lastError.c_str() : g_createError.c_str();
}

int vkrt_context_synchronize(VkrtContext* c)
{
    Context* ctx = C(c);
    cudaSetDevice(ctx->device);
    if (ctx->commStream)
    {
        VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->commStream));
    }
    return checkDeviceErrors(ctx, "vkrt_context_synchronize");
}

void* vkrt_context_stream(VkrtContext* c)
{
    return c ? (void*)C(c)->stream : nullptr;
}

int vkrt_blas_sizes(VkrtContext* c, const VkrtGeometryRange* ranges, uint32_t geometryCount, uint64_t vertexBytes, uint64_t indexCount,
                    VkrtBuildSizes* out)
{
    Context* ctx = C(c);
    if (!out || (!ranges && geometryCount))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_sizes: NULL argument");
    }
    uint64_t n = 0;
    for (uint32_t g = 0; g < geometryCount; ++g)
    {
        n += ranges[g].primitiveCount;
    }
    const uint64_t nodeCap = std::max<uint64_t>(1, n / 2 + 8);
    out->accelerationStructureSize = nodeCap * sizeof(Node8) + std::max<uint64_t>(1, n) * sizeof(PackedTri);
    // triangle boxes 32, build-order packed triangles 48, keys/vals x2 16, binary nodes 64, flags 4, taskRef 2-4,
    // primSource 4, onesweep look-back words 1, PLOC cluster boxes 64 (two ping-pong float4 pairs)
    out->buildScratchSize = std::max<uint64_t>(1, n) * 252 + 16384;
    out->geometryCopySize = vertexBytes + indexCount * 4;
    return VKRT_SUCCESS;
}

int vkrt_blas_build(VkrtContext* c, const void* vertices, size_t vertexBytes, uint32_t vertexStride, const uint32_t* indices,
                    size_t indexCount, const VkrtGeometryRange* ranges, uint32_t geometryCount, uint32_t flags, VkrtBlas** out)
{
    Context* ctx = C(c);
    if (!out)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: out is NULL");
    }
    *out = nullptr;
    if (vertexStride < 12 || (vertexStride & 3u))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: vertexStride must be a multiple of 4 and >= 12");
    }
    if ((!ranges && geometryCount) || (!vertices && vertexBytes) || (!indices && indexCount))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: NULL input");
    }
    std::vector<uint32_t> firstTri(geometryCount + 1, 0);
    std::vector<uint64_t> indexBase(std::max<uint32_t>(geometryCount, 1), 0);
    uint64_t n = 0;
    for (uint32_t g = 0; g < geometryCount; ++g)
    {
        if (ranges[g].indexByteOffset & 3u)
        {
            return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: indexByteOffset must be a multiple of 4");
        }
        indexBase[g] = ranges[g].indexByteOffset / 4;
        if (indexBase[g] + 3ull * ranges[g].primitiveCount > indexCount)
        {
            return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: geometry " + std::to_string(g) + " reads past the index buffer");
        }
        firstTri[g] = (uint32_t)n;
        n += ranges[g].primitiveCount;
    }
    if (n >= 0x40000000ull) // the sort's look-back words carry 30-bit counts
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: at most 2^30-1 triangles per BLAS");
    }
    firstTri[geometryCount] = (uint32_t)n;
    const bool onDevice = (flags & VKRT_BUILD_INPUT_ON_DEVICE) != 0;
    if (!onDevice && n > 0)
    {
        // host-side validation of the index range (the reference relies on maxVertex, Raytracer.cpp:1063)
        const uint64_t vertexCount = vertexBytes / vertexStride;
        for (uint32_t g = 0; g < geometryCount; ++g)
        {
            const uint32_t* idx = indices + indexBase[g];
            for (uint64_t k = 0; k < 3ull * ranges[g].primitiveCount; ++k)
            {
                if (idx[k] >= vertexCount)
                {
                    return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: index out of range in geometry " + std::to_string(g));
                }
            }
        }
    }

    cudaSetDevice(ctx->device);
    Blas* blas = new (std::nothrow) Blas();
    if (!blas)
    {
        return ctx->fail(VKRT_ERROR_OUT_OF_MEMORY, "vkrt_blas_build: host allocation failed");
    }
    blas->vertexStride = vertexStride;
    blas->shadingLayout = vertexStride >= 64 && (vertexStride & 15u) == 0; // shading reads 4 x float4 per vertex (Model::Vertex)
    blas->vertexBytes = vertexBytes;
    blas->indexCount = indexCount;
    blas->geometryCount = geometryCount;
    blas->triangleCount = (uint32_t)n;
    blas->flags = flags;
    auto cleanup = [&](int code) {
        vkrt_blas_destroy(c, blas);
        return code;
    };
    NvtxRange range("createBLAS"); // src/Raytracer.cpp:1027
    int r = uploadGeometry(ctx, blas, vertices, vertexBytes, indices, indexCount, onDevice);
    if (r != VKRT_SUCCESS)
    {
        return cleanup(r);
    }
    cudaError_t e = blas->geomFirstTri.reserve(firstTri.size() * 4);
    if (e == cudaSuccess)
    {
        e = blas->geomIndexBase.reserve(indexBase.size() * 8);
    }
    if (e == cudaSuccess)
    {
        e = cudaMemcpyAsync(blas->geomFirstTri.ptr, firstTri.data(), firstTri.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        e = cudaMemcpyAsync(blas->geomIndexBase.ptr, indexBase.data(), indexBase.size() * 8, cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        e = cudaStreamSynchronize(ctx->stream); // the host vectors above go out of scope
    }
    if (e != cudaSuccess)
    {
        return cleanup(ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_blas_build: ") + cudaGetErrorString(e)));
    }
    r = buildBlasOnDevice(ctx, blas);
    if (r != VKRT_SUCCESS)
    {
        return cleanup(r);
    }
    ctx->liveBlas.push_back(blas);
    *out = blas;
    return VKRT_SUCCESS;
}

int vkrt_blas_rebuild(VkrtContext* c, VkrtBlas* b, const void* vertices, size_t vertexBytes, uint32_t flags)
{
    Context* ctx = C(c);
    Blas* blas = B(b);
    if (!blas || (!vertices && vertexBytes))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_rebuild: NULL argument");
    }
    if (vertexBytes != blas->vertexBytes)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_rebuild: vertex data size differs from the original build");
    }
    cudaSetDevice(ctx->device);
    NvtxRange range("createBLAS (rebuild)");
    blas->flags = flags; // PLOC (default) or LBVH (VKRT_BUILD_PREFER_FAST_BUILD) may change between rebuilds
    int r = uploadGeometry(ctx, blas, vertices, vertexBytes, nullptr, 0, (flags & VKRT_BUILD_INPUT_ON_DEVICE) != 0);
    if (r != VKRT_SUCCESS)
    {
        return r;
    }
    return buildBlasOnDevice(ctx, blas);
}

void vkrt_blas_destroy(VkrtContext* c, VkrtBlas* b)
{
    if (!b)
    {
        return;
    }
    Blas* blas = B(b);
    if (c)
    {
        cudaSetDevice(C(c)->device);
        cudaStreamSynchronize(C(c)->stream);
        auto& live = C(c)->liveBlas;
        live.erase(std::remove(live.begin(), live.end(), (const Blas*)blas), live.end());
    }
    if (blas->buildGraph)
    {
        cudaGraphExecDestroy(blas->buildGraph);
        blas->buildGraph = nullptr;
    }
    blas->vertices.release(), blas->indices.release();
    blas->geomFirstTri.release(), blas->geomIndexBase.release();
    blas->nodes.release(), blas->tris.release();
    delete blas;
}

uint64_t vkrt_blas_reference(const VkrtBlas* b)
{
    return (uint64_t)(uintptr_t)b;
}

int vkrt_blas_info(VkrtContext* c, const VkrtBlas* b, VkrtBlasInfo* out)
{
    Context* ctx = C(c);
    if (!b || !out)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_info: NULL argument");
    }
    const Blas* blas = B(b);
    out->triangleCount = blas->triangleCount;
    out->nodeCount = blas->nodeCount;
    out->nodeBytes = (uint64_t)blas->nodeCount * sizeof(Node8);
    out->triangleBytes = (uint64_t)blas->triangleCount * sizeof(PackedTri);
    for (int k = 0; k < 3; ++k)
    {
        out->boundsMin[k] = blas->boundsMin[k];
        out->boundsMax[k] = blas->boundsMax[k];
    }
    out->sahCost = blas->sahCost;
    out->buildMilliseconds = blas->buildMs;
    return VKRT_SUCCESS;
}

int vkrt_blas_read(VkrtContext* c, const VkrtBlas* b, void* nodes, size_t nodeBytes, void* triangles, size_t triangleBytes)
{
    Context* ctx = C(c);
    const Blas* blas = B(b);
    if (!blas)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_read: NULL blas");
    }
    cudaSetDevice(ctx->device);
    if (nodes)
    {
        const size_t bytes = std::min<size_t>(nodeBytes, (size_t)blas->nodeCount * sizeof(Node8));
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(nodes, blas->nodes.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (triangles)
    {
        const size_t bytes = std::min<size_t>(triangleBytes, (size_t)blas->triangleCount * sizeof(PackedTri));
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(triangles, blas->tris.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return VKRT_SUCCESS;
}

int vkrt_tlas_build(VkrtContext* c, const VkrtInstance* instances, uint32_t instanceCount, VkrtTlas** out)
{
    Context* ctx = C(c);
    if (!out || (!instances && instanceCount))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tlas_build: NULL argument");
    }
    *out = nullptr;
    if (instanceCount == 0 || instanceCount > kMaxInstances)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tlas_build: instanceCount must be in 1.." + std::to_string(kMaxInstances));
    }
    cudaSetDevice(ctx->device);
    Tlas* tlas = new (std::nothrow) Tlas();
    if (!tlas)
    {
        return ctx->fail(VKRT_ERROR_OUT_OF_MEMORY, "vkrt_tlas_build: host allocation failed");
    }
    auto cleanup = [&](int code) {
        vkrt_tlas_destroy(c, tlas);
        return code;
    };
    NvtxRange range("createTLAS"); // src/Raytracer.cpp:1159
    std::vector<InstanceRecord> recs(instanceCount);
    uint32_t deepest = 0;
    const Blas* lastSeen = nullptr;
    for (uint32_t i = 0; i < instanceCount; ++i)
    {
        const Blas* blas = reinterpret_cast<const Blas*>((uintptr_t)instances[i].accelerationStructureReference);
        if (blas != lastSeen) // thousands of instances usually share a few BLASes
        {
            if (!blas || std::find(ctx->liveBlas.begin(), ctx->liveBlas.end(), blas) == ctx->liveBlas.end())
            {
                return cleanup(ctx->fail(VKRT_ERROR_INVALID_ARGUMENT,
                                         "vkrt_tlas_build: instance " + std::to_string(i) + " does not reference a live BLAS of this context"));
            }
            bool seen = false;
            for (const auto& ref : tlas->blasRefs)
            {
                seen = seen || ref.first == blas;
            }
            if (!seen)
            {
                tlas->blasRefs.emplace_back(blas, blas->generation);
            }
            deepest = std::max(deepest, blas->levels);
            tlas->maxGeometryCount = std::max(tlas->maxGeometryCount, blas->geometryCount);
            tlas->shadingLayout = tlas->shadingLayout && blas->shadingLayout;
            lastSeen = blas;
        }
        InstanceRecord& r = recs[i];
        std::memset(&r, 0, sizeof(r));
        std::memcpy(r.o2w, instances[i].transform, sizeof(float) * 12);
        affineInverseHost(r.o2w, r.w2o);
        r.nodes = blas->nodes.as<Node8>();
        r.tris = blas->tris.as<PackedTri>();
        r.vertices = blas->vertices.as<uint8_t>();
        r.indices = blas->indices.as<uint32_t>();
        r.vertexStride = blas->vertexStride;
        r.mask = instances[i].mask;
        r.customIndex = instances[i].instanceCustomIndex;
        r.flags = instances[i].flags;
        r.boxScale = blas->boxScale;
        for (int k = 0; k < 3; ++k)
        {
            r.blasLo[k] = blas->boundsMin[k];
            r.blasHi[k] = blas->boundsMax[k];
        }
    }
    tlas->instanceCount = instanceCount;
    cudaError_t e = tlas->instances.reserve(recs.size() * sizeof(InstanceRecord));
    if (e == cudaSuccess)
    {
        e = cudaMemcpyAsync(tlas->instances.ptr, recs.data(), recs.size() * sizeof(InstanceRecord), cudaMemcpyHostToDevice, ctx->stream);
    }
    if (e == cudaSuccess)
    {
        e = cudaStreamSynchronize(ctx->stream); // `recs` goes out of scope
    }
    if (e != cudaSuccess)
    {
        return cleanup(ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_tlas_build: ") + cudaGetErrorString(e)));
    }
    if (int r = buildTlasOnDevice(ctx, tlas))
    {
        return cleanup(r);
    }
    // one stack serves both levels (traverse.cuh): top-level path + the two groups parked on entering an instance + BLAS path
    if (tlas->nodeCount && tlas->levels + 2u + deepest > (uint32_t)kStackSize)
    {
        return cleanup(ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tlas_build: " + std::to_string(tlas->levels) + " top levels + " +
                                                                  std::to_string(deepest) + " bottom levels exceed the traversal stack"));
    }
    *out = tlas;
    return VKRT_SUCCESS;
}

void vkrt_tlas_destroy(VkrtContext* c, VkrtTlas* t)
{
    if (!t)
    {
        return;
    }
    if (c)
    {
        cudaSetDevice(C(c)->device);
        cudaStreamSynchronize(C(c)->stream);
    }
    Tlas* tlas = T(t);
    tlas->instances.release();
    tlas->nodes.release();
    tlas->leafInstance.release();
    delete tlas;
}

int vkrt_tlas_info(VkrtContext* c, const VkrtTlas* t, VkrtTlasInfo* out)
{
    Context* ctx = C(c);
    if (!t || !out)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tlas_info: NULL argument");
    }
    const Tlas* tlas = T(t);
    out->instanceCount = tlas->instanceCount;
    out->nodeCount = tlas->nodeCount;
    out->levels = tlas->levels;
    out->reserved = 0;
    for (int k = 0; k < 3; ++k)
    {
        out->boundsMin[k] = tlas->boundsMin[k];
        out->boundsMax[k] = tlas->boundsMax[k];
    }
    out->sahCost = tlas->sahCost;
    out->buildMilliseconds = tlas->buildMs;
    return VKRT_SUCCESS;
}

int vkrt_tlas_read(VkrtContext* c, const VkrtTlas* t, void* nodes, size_t nodeBytes, uint32_t* leafInstances, size_t leafBytes)
{
    Context* ctx = C(c);
    const Tlas* tlas = T(t);
    if (!tlas)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tlas_read: NULL tlas");
    }
    cudaSetDevice(ctx->device);
    if (nodes && tlas->nodeCount)
    {
        const size_t bytes = std::min<size_t>(nodeBytes, (size_t)tlas->nodeCount * sizeof(Node8));
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(nodes, tlas->nodes.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (leafInstances && tlas->nodeCount)
    {
        const size_t bytes = std::min<size_t>(leafBytes, (size_t)tlas->instanceCount * 4);
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(leafInstances, tlas->leafInstance.ptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    }
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return VKRT_SUCCESS;
}

int vkrt_materials_set(VkrtContext* c, const VkrtSubmeshInfo* rows, uint32_t count)
{
    Context* ctx = C(c);
    if (!rows || count == 0)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_materials_set: empty table");
    }
    cudaSetDevice(ctx->device);
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    VKRT_CUDA_TRY(ctx, ctx->materials.reserve((size_t)count * sizeof(VkrtSubmeshInfo)));
    ctx->materialsHost.assign(rows, rows + count);
    ctx->materialCount = count;
    ctx->materialsDirty = true;
    return syncMaterials(ctx);
}

int vkrt_textures_set(VkrtContext* c, const VkrtImage* images, uint32_t count)
{
    Context* ctx = C(c);
    if (!images || count == 0)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_textures_set: empty array");
    }
    for (uint32_t i = 0; i < count; ++i)
    {
        if (!images[i].rgba8 || images[i].width == 0 || images[i].height == 0)
        {
            return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_textures_set: image " + std::to_string(i) + " is empty");
        }
    }
    cudaSetDevice(ctx->device);
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    for (void* p : ctx->textureAllocs)
    {
        cudaFree(p);
    }
    ctx->textureAllocs.clear();
    ctx->textureCount = 0;
    ctx->textureHost.clear();
    ctx->mipsHost.clear();
    ctx->mipsReady = false; // a new texture set invalidates the mip chain
    // one allocation for all texels (256-byte aligned sub-ranges)
    std::vector<size_t> offsets(count);
    size_t total = 0;
    for (uint32_t i = 0; i < count; ++i)
    {
        offsets[i] = total;
        total += (((size_t)images[i].width * images[i].height * 4) + 255) & ~(size_t)255;
    }
    void* pool = nullptr;
    VKRT_CUDA_TRY(ctx, cudaMalloc(&pool, total));
    ctx->textureAllocs.push_back(pool);
    std::vector<TextureRecord> recs(count);
    for (uint32_t i = 0; i < count; ++i)
    {
        uint8_t* dst = static_cast<uint8_t*>(pool) + offsets[i];
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(dst, images[i].rgba8, (size_t)images[i].width * images[i].height * 4, cudaMemcpyHostToDevice, ctx->stream));
        recs[i].texels = dst;
        recs[i].width = images[i].width;
        recs[i].height = images[i].height;
    }
    VKRT_CUDA_TRY(ctx, ctx->textureRecords.reserve(recs.size() * sizeof(TextureRecord)));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(ctx->textureRecords.ptr, recs.data(), recs.size() * sizeof(TextureRecord), cudaMemcpyHostToDevice, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->textureCount = count;
    ctx->textureHost = recs;
    ctx->materialsDirty = true; // texture indices of the material rows are checked against the new image count
    ctx->textureGeneration += 1;
    return VKRT_SUCCESS;
}

int vkrt_textures_generate_mips(VkrtContext* c)
{
    Context* ctx = C(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (ctx->textureCount == 0)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_textures_generate_mips: vkrt_textures_set must come first");
    }
    cudaSetDevice(ctx->device);
    return generateMips(ctx);
}

int vkrt_texture_info(VkrtContext* c, uint32_t image, uint32_t level, uint32_t* width, uint32_t* height, uint32_t* levels)
{
    Context* ctx = C(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (image >= ctx->textureCount)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_texture_info: no image " + std::to_string(image));
    }
    const TextureRecord& t = ctx->textureHost[image];
    const uint32_t n = ctx->mipsReady ? ctx->mipsHost[image].levels : 1u;
    if (level >= n)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_texture_info: image " + std::to_string(image) + " has " + std::to_string(n) + " level(s)");
    }
    if (width)
    {
        *width = std::max(t.width >> level, 1u);
    }
    if (height)
    {
        *height = std::max(t.height >> level, 1u);
    }
    if (levels)
    {
        *levels = n;
    }
    return VKRT_SUCCESS;
}

int vkrt_texture_read(VkrtContext* c, uint32_t image, uint32_t level, void* dst, size_t bytes)
{
    Context* ctx = C(c);
    uint32_t w = 0, h = 0;
    if (int r = vkrt_texture_info(c, image, level, &w, &h, nullptr))
    {
        return r;
    }
    const size_t need = (size_t)w * h * 4;
    if (!dst || bytes < need)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_texture_read: destination smaller than " + std::to_string(need) + " bytes");
    }
    const TextureRecord& t = ctx->textureHost[image];
    const uint8_t* src = t.texels;
    if (level > 0)
    {
        src = ctx->mipsHost[image].texels;
        for (uint32_t k = 1; k < level; ++k)
        {
            src += (size_t)std::max(t.width >> k, 1u) * std::max(t.height >> k, 1u) * 4;
        }
    }
    cudaSetDevice(ctx->device);
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(dst, src, need, cudaMemcpyDeviceToHost, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return VKRT_SUCCESS;
}

int vkrt_trace_set_texture_lod(VkrtContext* c, int mode)
{
    Context* ctx = C(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (mode != VKRT_TEXTURE_LOD_ZERO && mode != VKRT_TEXTURE_LOD_RAY_CONES)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_set_texture_lod: unknown mode " + std::to_string(mode));
    }
    if (mode == VKRT_TEXTURE_LOD_RAY_CONES && !ctx->mipsReady)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_trace_set_texture_lod: vkrt_textures_generate_mips must come first");
    }
    ctx->textureLod = mode;
    return VKRT_SUCCESS;
}

int vkrt_frame_create(VkrtContext* c, uint32_t width, uint32_t height, VkrtFrame** out)
{
    Context* ctx = C(c);
    if (!out || width == 0 || height == 0 || width > 65536 || height > 65536)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_create: bad size");
    }
    *out = nullptr;
    cudaSetDevice(ctx->device);
    Frame* f = new (std::nothrow) Frame();
    if (!f)
    {
        return ctx->fail(VKRT_ERROR_OUT_OF_MEMORY, "vkrt_frame_create: host allocation failed");
    }
    f->width = width, f->height = height;
    f->tilesX = (width + kTile - 1) / kTile;
    f->tilesY = (height + kTile - 1) / kTile;
    f->rank = ctx->rank, f->size = ctx->size;
    const uint32_t tiles = f->tilesX * f->tilesY;
    f->ownedTiles = tiles > (uint32_t)f->rank ? (tiles - (uint32_t)f->rank + (uint32_t)f->size - 1) / (uint32_t)f->size : 0;
    f->paths = f->ownedTiles * (uint32_t)kTilePixels;
    cudaError_t e = f->bgra8.reserve((size_t)width * height * 4);
    if (e == cudaSuccess)
    {
        e = cudaMemsetAsync(f->bgra8.ptr, 0, (size_t)width * height * 4, ctx->stream);
    }
    if (e == cudaSuccess && f->size > 1)
    {
        e = f->tileBgra8.reserve(std::max<size_t>(1, f->ownedTiles) * kTilePixels * 4);
    }
    if (e != cudaSuccess)
    {
        vkrt_frame_destroy(c, f);
        return ctx->fail(e == cudaErrorMemoryAllocation ? VKRT_ERROR_OUT_OF_MEMORY : VKRT_ERROR_CUDA,
                         std::string("vkrt_frame_create: ") + cudaGetErrorString(e));
    }
    *out = f;
    return VKRT_SUCCESS;
}

void vkrt_frame_destroy(VkrtContext* c, VkrtFrame* fr)
{
    if (!fr)
    {
        return;
    }
    if (c)
    {
        cudaSetDevice(C(c)->device);
        cudaStreamSynchronize(C(c)->stream);
    }
    Frame* f = F(fr);
    if (c && C(c)->copyStream)
    {
        cudaStreamSynchronize(C(c)->copyStream);
    }
    if (c && C(c)->commStream)
    {
        cudaStreamSynchronize(C(c)->commStream);
    }
    if (f->gatherDone)
    {
        cudaEventDestroy(f->gatherDone);
    }
    if (f->presentReady)
    {
        cudaEventDestroy(f->presentReady);
    }
    for (cudaEvent_t e : f->presentDone)
    {
        if (e)
        {
            cudaEventDestroy(e);
        }
    }
    f->bgra8.release(), f->tileBgra8.release(), f->rgb32f.release(), f->triangleId.release(), f->shadowMask.release(), f->visibility.release();
    delete f;
}

int vkrt_trace(VkrtContext* c, const VkrtTlas* tlas, const VkrtUniforms* uniforms, uint32_t spp, uint32_t frameIndex, uint32_t flags,
               VkrtFrame* frame)
{
    Context* ctx = C(c);
    if (!tlas || !uniforms || !frame)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace: NULL argument");
    }
    cudaSetDevice(ctx->device);
    NvtxRange range("Render"); // DebugMarker::beginLabel(cb, "Render"), src/Raytracer.cpp:151
    return traceFrame(ctx, T(tlas), uniforms, spp, frameIndex, flags, F(frame));
}

int vkrt_frame_read(VkrtContext* c, VkrtFrame* fr, int aov, void* dst, size_t pitch)
{
    Context* ctx = C(c);
    Frame* f = F(fr);
    if (!f || !dst)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_read: NULL argument");
    }
    cudaSetDevice(ctx->device);
    const void* src = nullptr;
    size_t bpp = 0;
    switch (aov)
    {
    case VKRT_AOV_BGRA8:
        src = f->bgra8.ptr, bpp = 4;
        break;
    case VKRT_AOV_RGB32F:
        src = f->rgb32f.ptr, bpp = 12;
        break;
    case VKRT_AOV_TRIANGLE_ID_U32:
        src = f->triangleId.ptr, bpp = 4;
        break;
    case VKRT_AOV_SHADOW_MASK_U8:
        src = f->shadowMask.ptr, bpp = 1;
        break;
    default:
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_read: unknown AOV");
    }
    if (!src)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_frame_read: this output was not rendered (VKRT_TRACE_NO_AOV?)");
    }
    if (int r = joinGather(ctx, f))
    {
        return r;
    }
    const size_t row = (size_t)f->width * bpp;
    if (pitch == 0)
    {
        pitch = row;
    }
    if (pitch < row)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_read: pitch smaller than a row");
    }
    VKRT_CUDA_TRY(ctx, cudaMemcpy2DAsync(dst, pitch, src, row, row, f->height, cudaMemcpyDeviceToHost, ctx->stream));
    return checkDeviceErrors(ctx, "vkrt_frame_read");
}

int vkrt_frame_present(VkrtContext* c, VkrtFrame* fr, void* dst, size_t pitch, uint64_t* ticket)
{
    Context* ctx = C(c);
    Frame* f = F(fr);
    if (!f || !dst || !ticket)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_present: NULL argument");
    }
    if (!f->bgra8.ptr)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_frame_present: no image");
    }
    const size_t row = (size_t)f->width * 4;
    if (pitch == 0)
    {
        pitch = row;
    }
    if (pitch < row)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_present: pitch smaller than a row");
    }
    cudaSetDevice(ctx->device);
    if (!ctx->copyStream)
    {
        VKRT_CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
    }
    if (!f->presentReady)
    {
        VKRT_CUDA_TRY(ctx, cudaEventCreateWithFlags(&f->presentReady, cudaEventDisableTiming));
    }
    if (int r = joinGather(ctx, f))
    {
        return r;
    }
    const uint64_t n = f->presentNext;
    cudaEvent_t& done = f->presentDone[n % Frame::kPresentRing];
    if (!done)
    {
        VKRT_CUDA_TRY(ctx, cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
    }
    else if (n >= (uint64_t)Frame::kPresentRing && f->presentWaited + Frame::kPresentRing <= n)
    {
        // the event is about to be reused: the copy it stood for (ticket n - ring) has to be over
        VKRT_CUDA_TRY(ctx, cudaEventSynchronize(done));
        f->presentWaited = n - Frame::kPresentRing + 1;
    }
    VKRT_CUDA_TRY(ctx, cudaEventRecord(f->presentReady, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copyStream, f->presentReady, 0));
    VKRT_CUDA_TRY(ctx, cudaMemcpy2DAsync(dst, pitch, f->bgra8.ptr, row, row, f->height, cudaMemcpyDeviceToHost, ctx->copyStream));
    VKRT_CUDA_TRY(ctx, cudaEventRecord(done, ctx->copyStream));
    *ticket = n;
    f->presentNext = n + 1;
    return VKRT_SUCCESS;
}

int vkrt_frame_present_wait(VkrtContext* c, VkrtFrame* fr, uint64_t ticket)
{
    Context* ctx = C(c);
    Frame* f = F(fr);
    if (!f)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_present_wait: NULL argument");
    }
    if (ticket >= f->presentNext)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_frame_present_wait: unknown ticket");
    }
    if (ticket < f->presentWaited || ticket + Frame::kPresentRing < f->presentNext)
    {
        return VKRT_SUCCESS; // already waited for (copies finish in order; reused events were synchronised)
    }
    cudaSetDevice(ctx->device);
    VKRT_CUDA_TRY(ctx, cudaEventSynchronize(f->presentDone[ticket % Frame::kPresentRing]));
    f->presentWaited = ticket + 1;
    return VKRT_SUCCESS;
}

int vkrt_host_alloc(VkrtContext* c, size_t bytes, void** out)
{
    Context* ctx = C(c);
    if (!out)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_host_alloc: NULL argument");
    }
    cudaSetDevice(ctx->device);
    VKRT_CUDA_TRY(ctx, cudaHostAlloc(out, std::max<size_t>(bytes, 1), cudaHostAllocDefault));
    return VKRT_SUCCESS;
}

void vkrt_host_free(VkrtContext* c, void* ptr)
{
    if (ptr)
    {
        if (c)
        {
            cudaSetDevice(C(c)->device);
        }
        cudaFreeHost(ptr);
    }
}

void* vkrt_frame_device_bgra8(VkrtFrame* fr)
{
    return fr ? F(fr)->bgra8.ptr : nullptr;
}

int vkrt_rasterize(VkrtContext* c, const VkrtBlas* b, const float* wvp, uint32_t flags, VkrtFrame* fr)
{
    Context* ctx = C(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (!b || !wvp || !fr)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_rasterize: NULL argument");
    }
    cudaSetDevice(ctx->device);
    NvtxRange range("Render"); // DebugMarker::beginLabel(cb, "Render"), src/Rasterizer.cpp:117
    return rasterizeFrame(ctx, static_cast<const Blas*>(b), wvp, flags, F(fr));
}

int vkrt_trace_set_lane_refill(VkrtContext* c, uint32_t lanes)
{
    Context* ctx = C(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (lanes > 32)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_set_lane_refill: lanes must be in 0..32");
    }
    ctx->refillLanes = lanes;
    return VKRT_SUCCESS;
}

int vkrt_trace_rays(VkrtContext* c, const VkrtTlas* tlas, const VkrtRay* rays, uint64_t count, int anyHit, VkrtHit* hits)
{
    Context* ctx = C(c);
    if (!tlas || (count && (!rays || !hits)))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_rays: NULL argument");
    }
    cudaSetDevice(ctx->device);
    return traceRaysHost(ctx, T(tlas), rays, count, anyHit, hits);
}

int vkrt_stats_enable(VkrtContext* c, int countTraversal, int timeStages)
{
    Context* ctx = C(c);
    ctx->countTraversal = countTraversal != 0;
    ctx->timeStages = timeStages != 0;
    return VKRT_SUCCESS;
}

int vkrt_stats_reset(VkrtContext* c)
{
    Context* ctx = C(c);
    cudaSetDevice(ctx->device);
    ctx->stats = VkrtStats{};
    ctx->launches = 0;
    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(ctx->statCounters.ptr, 0, 128, ctx->stream));
    return VKRT_SUCCESS;
}

int vkrt_stats_get(VkrtContext* c, VkrtStats* out)
{
    Context* ctx = C(c);
    if (!out)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_stats_get: NULL argument");
    }
    cudaSetDevice(ctx->device);
    unsigned long long host[16] = {};
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(host, ctx->statCounters.ptr, sizeof(host), cudaMemcpyDeviceToHost, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->stats.raysShadow = host[1];
    ctx->stats.raysReflection = host[2];
    ctx->stats.raysShadowBounce = host[8];
    ctx->stats.nodesVisited = host[3] + host[6];
    ctx->stats.trianglesTested = host[4] + host[7];
    ctx->stats.nodesVisitedAny = host[6];
    ctx->stats.trianglesTestedAny = host[7];
    ctx->stats.kernelLaunches = ctx->launches;
    if (ctx->timeStages && ctx->stageEventCount == 11)
    {
        auto ms = [&](int a, int b) {
            float v = 0.f;
            cudaEventElapsedTime(&v, ctx->stageEvents[a], ctx->stageEvents[b]);
            return v;
        };
        ctx->stats.msRaygen = ms(0, 1);
        ctx->stats.msTraceClosest = ms(1, 2) + ms(5, 6);
        ctx->stats.msShade = ms(2, 3) + ms(4, 5) + ms(6, 7) + ms(8, 9);
        ctx->stats.msTraceAny = ms(3, 4) + ms(7, 8);
        ctx->stats.msResolve = ms(9, 10);
        ctx->stats.msTotal = ms(0, 10);
        for (int k = 0; k < 10; ++k)
        {
            ctx->stats.msStage[k] = ms(k, k + 1);
        }
    }
    *out = ctx->stats;
    return VKRT_SUCCESS;
}

// Test hook (not part of include/vkrt.h): runs the builder's radix sort on host arrays.
VKRT_API int vkrt_debug_sort_pairs(VkrtContext* c, uint32_t* keys, uint32_t* vals, uint32_t count)
{
    Context* ctx = C(c);
    cudaSetDevice(ctx->device);
    return sortPairsForTest(ctx, keys, vals, count);
}

int vkrt_comm_unique_id(uint8_t id[VKRT_UNIQUE_ID_BYTES])
{
    std::string err;
    const int r = commUniqueId(id, err);
    if (r != VKRT_SUCCESS)
    {
        g_createError = err;
    }
    return r;
}

int vkrt_comm_init(VkrtContext* c, const uint8_t id[VKRT_UNIQUE_ID_BYTES], int rank, int size)
{
    Context* ctx = C(c);
    if (size < 1 || rank < 0 || rank >= size || !id)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_comm_init: bad rank/size");
    }
    cudaSetDevice(ctx->device);
    return commInit(ctx, id, rank, size);
}

int vkrt_comm_emulate(VkrtContext* c, int rank, int size)
{
    Context* ctx = C(c);
    if (size < 1 || rank < 0 || rank >= size)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_comm_emulate: bad rank/size");
    }
    ctx->rank = rank;
    ctx->size = size;
    return VKRT_SUCCESS;
}

int vkrt_comm_destroy(VkrtContext* c)
{
    return commDestroy(C(c));
}

uint64_t vkrt_hash_fnv1a64(const void* data, size_t bytes)
{
    const uint8_t* p = static_cast<const uint8_t*>(data);
    uint64_t h = 0xCBF29CE484222325ull;
    for (size_t i = 0; i < bytes; ++i)
    {
        h = (h ^ p[i]) * 0x100000001B3ull;
    }
    return h;
}

int vkrt_tile_owner(uint32_t width, uint32_t x, uint32_t y, int size)
{
    const uint32_t tilesX = (width + kTile - 1) / kTile;
    const uint32_t tile = (y / kTile) * tilesX + (x / kTile);
    return size > 0 ? (int)(tile % (uint32_t)size) : 0;
}

int vkrt_tiles_gather_wait(VkrtContext* c, VkrtFrame* frame)
{
    Context* ctx = C(c);
    if (!frame)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tiles_gather_wait: NULL frame");
    }
    cudaSetDevice(ctx->device);
    return joinGather(ctx, F(frame));
}

int vkrt_textures_broadcast(VkrtContext* c, int root)
{
    Context* ctx = C(c);
    if (!ctx || root < 0 || root >= ctx->size)
    {
        return ctx ? ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_textures_broadcast: bad root") : VKRT_ERROR_INVALID_ARGUMENT;
    }
    cudaSetDevice(ctx->device);
    NvtxRange range("createTextures (broadcast)");
    return broadcastTextures(ctx, root);
}

int vkrt_tiles_gather(VkrtContext* c, VkrtFrame* frame, int root)
{
    Context* ctx = C(c);
    if (!frame || root < 0 || root >= ctx->size)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_tiles_gather: bad argument");
    }
    cudaSetDevice(ctx->device);
    NvtxRange range("Tile gather");
    return gatherTiles(ctx, F(frame), root);
}

} // extern "C"
