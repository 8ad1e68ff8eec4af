// context.hpp — host-side objects behind the opaque handles of include/vkrt.h.
#pragma once

#include "common.cuh"

#include <memory>
#include <utility>

struct VkrtContext
{
};
struct VkrtBlas
{
};
struct VkrtTlas
{
};
struct VkrtFrame
{
};

namespace vkrt
{

struct BuildScratch
{
    DeviceBuffer triBox, packedBuild, keysA, keysB, valsA, valsB, tileHist, binNodes, flags, taskRef, primSource, small, sortedBox;
    DeviceBuffer plocLoA, plocHiA, plocLoB, plocHiB, plocStatus; // PLOC cluster boxes (ping-pong) and look-back words
    DeviceBuffer dp;                                              // cost-optimal collapse: decisions (n), parents (2n)
    DeviceBuffer reinsert;                                        // reinsertion passes: target, gain (4 B each) and lock (8 B) per node
    void release()
    {
        DeviceBuffer* all[] = {&triBox, &packedBuild, &keysA, &keysB, &valsA, &valsB, &tileHist, &binNodes, &flags,
                               &taskRef, &primSource, &small, &sortedBox, &plocLoA, &plocHiA, &plocLoB, &plocHiB, &plocStatus, &dp, &reinsert};
        for (DeviceBuffer* b : all)
        {
            b->release();
        }
    }
};

// Wavefront queues of one frame; grow-only, sized by the number of owned pixel paths.
struct WaveBuffers
{
    DeviceBuffer rayO, rayD;         // float4[P] current closest-hit rays (primary), then reflection queue rays
    DeviceBuffer reflO, reflD;       // float4[P] reflection rays (compacted)
    DeviceBuffer reflPath;           // u32[P] path index of each reflection queue entry
    DeviceBuffer hitA;               // float4[P] (t, b1, b2, tri bits)
    DeviceBuffer hitInst;            // u32[P]
    DeviceBuffer surfA, surfB;       // float4[P] (worldPos, u) / (normal, v)
    DeviceBuffer surfGeom;           // u32[P]
    DeviceBuffer shadowBits;         // u32[P] per queue entry
    DeviceBuffer shadowO, shadowD;   // float4[4*P] per-light shadow ray queues
    DeviceBuffer pathAccum;          // float4[P] finalHitValue of the current sample (xyz) + attenuation (w)
    DeviceBuffer sampleSum;          // float4[P] sum over samples
    DeviceBuffer counters;           // u32[16]: [0..3] shadow queue sizes, [4] reflection queue size, [8..] stats
    DeviceBuffer coneT, surfFoot;    // float[P] each, reserved only in the ray-cone texture LOD mode
    void release()
    {
        DeviceBuffer* all[] = {&rayO, &rayD, &reflO, &reflD, &reflPath, &hitA, &hitInst, &surfA, &surfB, &surfGeom, &shadowBits,
                               &shadowO, &shadowD, &pathAccum, &sampleSum, &counters, &coneT, &surfFoot};
        for (DeviceBuffer* b : all)
        {
            b->release();
        }
    }
};

// Scratch of the software rasteriser (raster.cu); grow-only, reused across frames.
struct RasterBuffers
{
    DeviceBuffer setups;                          // RasterSetup[triangles]
    DeviceBuffer smallList, largeList, largeTiles; // u32[triangles] each
    DeviceBuffer sortKeys, sortKeysB, sortValsB, sortState; // front-to-back order of the large triangles
    DeviceBuffer counters;                        // u32[16]
    DeviceBuffer tileZ;                           // u32 per 16x16 tile: hierarchical-Z bound (raster.cu)
    DeviceBuffer minAlpha;                        // u32 per image: smallest alpha byte
    DeviceBuffer items;                           // uint2 (triangle, tile) work list of the binned large-triangle path
    uint64_t minAlphaGeneration = ~0ull;          // Context::textureGeneration the table was computed for
    void release()
    {
        DeviceBuffer* all[] = {&setups, &smallList, &largeList, &largeTiles, &sortKeys, &sortKeysB, &sortValsB, &sortState, &counters, &tileZ, &minAlpha, &items};
        for (DeviceBuffer* b : all)
        {
            b->release();
        }
    }
};

struct NcclApi; // dlopen'ed NCCL entry points (comm.cpp)
struct Blas;

struct Context : VkrtContext
{
    int device = 0;
    int smCount = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copyStream = nullptr; // vkrt_frame_present copies (created on first use)
    cudaStream_t commStream = nullptr; // vkrt_tiles_gather: NCCL exchange + untile, overlapping the next frame's traversal
    cudaEvent_t commReady = nullptr;   // recorded on `stream` when a gather is queued
    void* buildReadback = nullptr;     // 256 page-locked bytes: the one read-back of a graph-replayed build
    std::string lastError;
    uint64_t launches = 0;

    // shading resources
    DeviceBuffer materials; // VkrtSubmeshInfo[], texture indices made safe (syncMaterials, api.cu)
    uint32_t materialCount = 0;
    std::vector<VkrtSubmeshInfo> materialsHost; // the table as the caller gave it
    bool materialsDirty = false;
    DeviceBuffer textureRecords; // TextureRecord[]
    std::vector<void*> textureAllocs;
    uint32_t textureCount = 0;
    uint64_t textureGeneration = 0; // bumped whenever the texel pool changes (vkrt_textures_set / _broadcast)
    std::vector<TextureRecord> textureHost; // host copy of the records (device pointers), for vkrt_texture_info / _read
    DeviceBuffer textureMips;               // TextureMips[], parallel to textureRecords (vkrt_textures_generate_mips)
    std::vector<TextureMips> mipsHost;
    DeviceBuffer mipPool;                   // levels 1.. of every image
    bool mipsReady = false;
    int textureLod = VKRT_TEXTURE_LOD_ZERO; // vkrt_trace_set_texture_lod
    DeviceBuffer uniforms; // VkrtUniforms

    BuildScratch buildScratch;
    WaveBuffers wave;
    RasterBuffers raster;

    // stats
    uint32_t waveKeyPaths = 0, waveKeySpp = 0, waveSamples = 1; // cached samples-per-wave decision (wavefront.cu)
    uint32_t refillLanes = 0; // vkrt_trace_set_lane_refill: 0 = whole-warp re-arm (default)
    bool countTraversal = false;
    bool timeStages = false;
    VkrtStats stats{};
    DeviceBuffer statCounters; // u64[16] device-side: rays p/s/r, nodes, tris; [8] shadow rays of the bounce
    cudaEvent_t stageEvents[16] = {};
    bool eventsCreated = false;
    int stageEventCount = 0;

    // multi-GPU
    std::vector<const Blas*> liveBlas; // handles handed out by vkrt_blas_build and not yet destroyed
    DeviceBuffer errorFlags;           // u32[4] device-side: [0] traversal stack overflow
    int rank = 0, size = 1;
    void* ncclComm = nullptr;
    DeviceBuffer gatherStaging;

    int fail(int code, const std::string& text)
    {
        lastError = text;
        return code;
    }
};

struct Blas : VkrtBlas
{
    DeviceBuffer vertices, indices;       // device copies of the geometry (kept for shading)
    DeviceBuffer geomFirstTri, geomIndexBase;
    uint32_t vertexStride = 0;
    uint64_t vertexBytes = 0, indexCount = 0;
    uint32_t geometryCount = 0;
    uint32_t triangleCount = 0;
    DeviceBuffer nodes, tris;
    uint32_t nodeCount = 0;
    float boundsMin[3] = {0, 0, 0}, boundsMax[3] = {0, 0, 0};
    float sahCost = 0.0f;
    float buildMs = 0.0f;
    uint32_t flags = 0;
    uint32_t levels = 0;        // wide levels of the hierarchy (bounds the traversal stack)
    float boxScale = 1.0f;      // power of two that normalises box-test distances in this hierarchy (traverse.cuh::setBoxRay)
    uint64_t generation = 0;    // bumped by every (re)build: a TLAS remembers the generation it saw
    bool shadingLayout = false; // vertices are the 64-byte Model::Vertex (stride % 16 == 0, >= 64): vkrt_trace / vkrt_rasterize can shade it
    // Small fast builds (LBVH flavour) replay their whole kernel chain as one CUDA graph on vkrt_blas_rebuild (bvh_build.cu).
    cudaGraphExec_t buildGraph = nullptr;
    uint64_t buildGraphKey[24] = {}; // every pointer and size the captured chain depends on
    uint32_t buildGraphLaunches = 0;
};

struct Tlas : VkrtTlas
{
    DeviceBuffer instances;    // InstanceRecord[]
    DeviceBuffer nodes;        // Node8[]: hierarchy over the instances' world boxes (instanceCount > 1)
    DeviceBuffer leafInstance; // u32[instanceCount]: instance index per top-level leaf
    uint32_t instanceCount = 0;
    uint32_t nodeCount = 0, levels = 0;
    float boxScale = 1.0f;         // as Blas::boxScale, for the top-level hierarchy
    uint32_t maxGeometryCount = 0; // largest geometry count among the referenced BLASes (material table check)
    bool shadingLayout = true;     // every referenced BLAS can be shaded
    float buildMs = 0.0f, sahCost = 0.0f;
    float boundsMin[3] = {0, 0, 0}, boundsMax[3] = {0, 0, 0};
    std::vector<std::pair<const Blas*, uint64_t>> blasRefs; // (BLAS, generation at build) per distinct BLAS
};

struct Frame : VkrtFrame
{
    uint32_t width = 0, height = 0;
    uint32_t tilesX = 0, tilesY = 0;
    int rank = 0, size = 1;
    uint32_t ownedTiles = 0;
    uint32_t paths = 0;       // ownedTiles * 4096
    DeviceBuffer bgra8;       // width*height*4 linear image (root / single GPU)
    DeviceBuffer tileBgra8;   // ownedTiles*4096*4 tile-major (multi-GPU send buffer)
    DeviceBuffer rgb32f;      // width*height*3 floats (AOV)
    DeviceBuffer triangleId;  // width*height u32 (AOV)
    DeviceBuffer shadowMask;  // width*height u8 (AOV)
    DeviceBuffer visibility;  // width*height*8 u64 (depth24 << 32 | triangle) keys of the rasteriser (raster.cu)
    bool gathered = false;
    cudaEvent_t gatherDone = nullptr; // recorded on the comm stream after a gather of this frame
    bool gatherPending = false;       // the context stream has not waited for gatherDone yet
    // presentation ring (vkrt_frame_present): copy n uses presentDone[n % kPresentRing]
    static constexpr int kPresentRing = 8;
    cudaEvent_t presentReady = nullptr;              // recorded on the context stream when a present is queued
    cudaEvent_t presentDone[kPresentRing] = {};      // recorded on the copy stream after each copy
    uint64_t presentNext = 0;                        // tickets handed out so far
    uint64_t presentWaited = 0;                      // every ticket below this has been waited for on the host
};
// Makes the context stream wait for the outstanding presents of `frame` before its image is overwritten.
int waitPresents(Context* ctx, Frame* frame);
// Makes the context stream wait for an outstanding vkrt_tiles_gather of `frame` (before its buffers are read or rewritten).
int joinGather(Context* ctx, Frame* frame);

int buildBlasOnDevice(Context* ctx, Blas* blas);
int buildTlasOnDevice(Context* ctx, Tlas* tlas);
int syncMaterials(Context* ctx);
int checkTlasCurrent(Context* ctx, const Tlas* tlas, const char* who);
int checkSceneForShading(Context* ctx, const Tlas* tlas, const char* who); // stale TLAS, vertex layout, material rows
int checkDeviceErrors(Context* ctx, const char* who); // device-side error flags, read at synchronisation points
int sortPairsForTest(Context* ctx, uint32_t* keys, uint32_t* vals, uint32_t count);
size_t sortPairsStateBytes(uint32_t count);
int sortPairsDevice(Context* ctx, uint32_t* keysA, uint32_t* valsA, uint32_t* keysB, uint32_t* valsB, uint32_t* state, uint32_t count);
int traceFrame(Context* ctx, const Tlas* tlas, const VkrtUniforms* uniforms, uint32_t spp, uint32_t frameIndex, uint32_t flags,
               Frame* frame);
int traceRaysHost(Context* ctx, const Tlas* tlas, const VkrtRay* rays, uint64_t count, int anyHit, VkrtHit* hits);
int generateMips(Context* ctx); // textures.cu
int rasterizeFrame(Context* ctx, const Blas* blas, const float* wvp, uint32_t flags, Frame* frame); // raster.cu
int gatherTiles(Context* ctx, Frame* frame, int root);
int broadcastTextures(Context* ctx, int root);
int commInit(Context* ctx, const uint8_t* id, int rank, int size);
int commDestroy(Context* ctx);
int commUniqueId(uint8_t* id, std::string& err);

} // namespace vkrt
