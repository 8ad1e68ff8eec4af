// wavefront.cu — the wavefront ray engine that replaces vkCmdTraceRaysKHR(W,H,1) (src/Raytracer.cpp:157) and the
// shader pipeline it dispatches (shader.rgen -> traceRayEXT -> shader.rchit / shader.rmiss / shader_shadow.rmiss).
//
// Stages per sample (each a persistent grid-stride kernel; queue sizes stay on the device, no host sync):
//   k_raygen -> k_trace_closest -> k_shade_surface (emits per-light shadow-ray queues, warp-aggregated compaction)
//            -> k_trace_any -> k_shade_resolve (emits the compacted reflection queue)
//            -> k_trace_closest -> k_shade_surface -> k_trace_any -> k_shade_resolve   (the single bounce, rgen:40)
//            -> k_resolve_sample (sample average, B8G8R8A8 store, tile-major copy for the NCCL gather)
#include "context.hpp"
#include "shading.cuh"
#include "traverse.cuh"

#include <algorithm>
#include <cstddef>
#include <cstdlib>
#include <mutex>
#include <string>

namespace vkrt
{

namespace
{
constexpr int kTraceThreads = 128;
#ifndef VKRT_TRACE_MIN_BLOCKS
#define VKRT_TRACE_MIN_BLOCKS 7
#endif
constexpr int kShadeThreads = 128;

struct SceneDev
{
    const InstanceRecord* instances;
    const Node8* topNodes;        // top-level hierarchy over the instances (null when there is a single instance)
    const uint32_t* leafInstance; // instance per top-level leaf
    uint32_t instanceCount;
    uint32_t cullMask;            // traceRayEXT's cullMask: 0xFF in shader.rgen:52 and shader.rchit:133
    uint32_t* errorFlags;         // [0] traversal stack overflow
    float topBoxScale;            // Tlas::boxScale
    const VkrtSubmeshInfo* rows;
    const TextureRecord* textures;
    uint32_t one; // 0x3F800000, see byteToUnitFloat (traverse.cuh)
    const TextureMips* mips; // ray-cone LOD mode only, else nullptr
};

struct WaveDev
{
    float4 *rayO, *rayD, *reflO, *reflD;
    uint32_t* reflPath;
    float4* hitA;
    uint32_t* hitInst;
    float4 *surfA, *surfB;
    uint32_t *surfGeom, *shadowBits;
    float4 *shadowO, *shadowD;
    float4 *pathAccum, *sampleSum;
    uint32_t* counters; // [0..3] shadow queue sizes, [4] reflection queue size
    uint32_t capacity;  // P
    float *coneT, *surfFoot; // ray-cone LOD mode only: primary hit distance per path, squared uv footprint per surface
};

struct FrameDev
{
    TileMap map;
    uint32_t* bgra8;      // linear image or nullptr
    uint32_t* tileBgra8;  // tile-major or nullptr
    float* rgb32f;        // AOVs or nullptr
    uint32_t* triangleId;
    uint8_t* shadowMask;
};

__device__ __forceinline__ uint32_t warpAppend(uint32_t* counter, bool want)
{
    // warp-aggregated queue append: one atomic per warp, slots handed out in lane order
    const uint32_t mask = __ballot_sync(__activemask(), want);
    if (!want)
    {
        return 0;
    }
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader)
    {
        base = atomicAdd(counter, __popc(mask));
    }
    base = __shfl_sync(mask, base, leader);
    return base + __popc(mask & ((1u << lane) - 1u));
}

// Block-aggregated append to the four per-light queues: one atomic per block, light and loop iteration instead of one
// per warp (at 4K x 8 samples the warp form issued 2 M atomics on each of the four counters and k_shade_surface spent
// 73 % of its stall samples waiting for them).  Every thread of the block must call it, the same number of times;
// `parity` alternates so that two iterations never share a staging set.  Returns the slots through `slot[4]`.
constexpr int kShadeWarps = 8; // k_shade_surface block = 256 threads
struct AppendStage
{
    uint32_t count[2][4][kShadeWarps]; // per-warp counts, turned into exclusive prefixes by threads 0..3
    uint32_t base[2][4];
};
__device__ __forceinline__ void blockAppend4(AppendStage& st, uint32_t* counters, uint32_t wantMask, int parity, uint32_t* slot)
{
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t ballots[4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        ballots[i] = __ballot_sync(0xFFFFFFFFu, (wantMask >> i) & 1u);
        if (lane == 0)
        {
            st.count[parity][i][warp] = __popc(ballots[i]);
        }
    }
    __syncthreads();
    if (threadIdx.x < 4)
    {
        uint32_t total = 0;
        for (int k = 0; k < kShadeWarps; ++k)
        {
            const uint32_t c = st.count[parity][threadIdx.x][k];
            st.count[parity][threadIdx.x][k] = total;
            total += c;
        }
        st.base[parity][threadIdx.x] = total ? atomicAdd(&counters[threadIdx.x], total) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 4; ++i)
    {
        slot[i] = st.base[parity][i] + st.count[parity][i][warp] + __popc(ballots[i] & ((1u << lane) - 1u));
    }
}

__global__ void k_reset_counters(uint32_t* counters)
{
    if (threadIdx.x < 16) // [0..3] shadow queue sizes, [4] reflection queue size, [8..11] ray-fetch cursors of the trace kernels
    {
        counters[threadIdx.x] = 0u;
    }
}

// shader.rgen:32-47 — one primary ray per owned pixel path
__global__ void __launch_bounds__(kShadeThreads)
k_raygen(const __grid_constant__ VkrtUniforms u, TileMap map, uint32_t paths, uint32_t sampleBase, uint32_t samples, uint32_t spp,
         float4* __restrict__ rayO, float4* __restrict__ rayD, float4* __restrict__ pathAccum)
{
    // a wave holds `samples` consecutive samples of every owned pixel path, pixel-major: index = path * samples + s, so a
    // warp traces the samples of a few neighbouring pixels together (near-identical rays: they visit the same nodes and
    // triangles in step; sample-major order left the triangle test at 13-16 of 32 lanes)
    const uint32_t total = paths * samples;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        const uint32_t p = i / samples;
        float jx, jy;
        sampleOffset(sampleBase + i % samples, spp, jx, jy);
        uint32_t x, y, tl, pit;
        if (pathToPixel(map, p, x, y, tl, pit))
        {
            F3 o, d;
            primaryRay(u, map.width, map.height, x, y, jx, jy, o, d);
            rayO[i] = make_float4(o.x, o.y, o.z, kRayTMin);
            rayD[i] = make_float4(d.x, d.y, d.z, kRayTMax);
        }
        else
        {
            rayO[i] = make_float4(0.f, 0.f, 0.f, kRayTMin);
            rayD[i] = make_float4(0.f, 0.f, 0.f, -1.0f); // inactive path
        }
        pathAccum[i] = make_float4(0.f, 0.f, 0.f, 1.0f); // finalHitValue = 0, attenuation = 1 (rgen:42-45)
    }
}

// ------------------------------------------------------------------------------------------------------------
// Persistent-thread traversal with dynamic ray fetch (Aila & Laine 2009; Ylitie et al. 2017, section 4.3).
// A warp keeps a private pool of ray indices that it refills from the launch-wide cursor in chunks (one global
// atomic per chunk) and by default re-arms all 32 lanes together once every lane's ray has finished: on coherent
// workloads (Sponza) per-lane replacement below 4..28 busy lanes was measured slower (config 2: 2014 Mrays/s without,
// 1690-1900 with), as was triangle postponing.  The REFILL variant (vkrt_trace_set_lane_refill) re-arms idle lanes
// as soon as fewer than `refill` lanes are traversing; it pays on incoherent scenes such as the config-5 soup, where
// only 13-15 of 32 lanes are busy otherwise.
// Sources map a queue position to a ray and receive the result.
// ------------------------------------------------------------------------------------------------------------
struct ClosestSource
{
    const float4 *rayO, *rayD;
    float4* hitA;
    uint32_t* hitInst;
    __device__ __forceinline__ uint32_t map(uint32_t i) const { return i; }
    __device__ __forceinline__ void load(uint32_t h, F3& o, F3& d, float& tmin, float& tmax) const
    {
        const float4 o4 = rayO[h], d4 = rayD[h];
        o = f3(o4.x, o4.y, o4.z), d = f3(d4.x, d4.y, d4.z);
        tmin = o4.w, tmax = d4.w;
    }
    __device__ __forceinline__ void commit(uint32_t h, const BestHit& best, bool) const
    {
        hitA[h] = make_float4(best.t, best.b1, best.b2, __uint_as_float(best.tri));
        hitInst[h] = best.instance;
    }
};

struct ShadowSource // the four per-light queues, concatenated
{
    const float4 *shadowO, *shadowD;
    uint32_t* shadowBits;
    uint32_t c0, c01, c012, capacity;
    __device__ __forceinline__ uint32_t map(uint32_t i) const
    {
        return i < c0 ? i : (i < c01 ? capacity + (i - c0) : (i < c012 ? 2u * capacity + (i - c01) : 3u * capacity + (i - c012)));
    }
    __device__ __forceinline__ void load(uint32_t h, F3& o, F3& d, float& tmin, float& tmax) const
    {
        const float4 o4 = shadowO[h], d4 = shadowD[h];
        o = f3(o4.x, o4.y, o4.z), d = f3(d4.x, d4.y, d4.z);
        tmin = kRayTMin, tmax = o4.w;
    }
    __device__ __forceinline__ void commit(uint32_t h, const BestHit&, bool occluded) const
    {
        if (occluded)
        {
            const uint32_t light = h / capacity;
            atomicOr(&shadowBits[__float_as_uint(shadowD[h].w)], 1u << light);
        }
    }
};

struct ListAnySource // ray-level API, any-hit
{
    const float4 *rayO, *rayD;
    VkrtHit* out;
    __device__ __forceinline__ uint32_t map(uint32_t i) const { return i; }
    __device__ __forceinline__ void load(uint32_t h, F3& o, F3& d, float& tmin, float& tmax) const
    {
        const float4 o4 = rayO[h], d4 = rayD[h];
        o = f3(o4.x, o4.y, o4.z), d = f3(d4.x, d4.y, d4.z);
        tmin = o4.w, tmax = d4.w;
    }
    __device__ __forceinline__ void commit(uint32_t h, const BestHit& best, bool occluded) const
    {
        VkrtHit o;
        o.t = occluded ? best.t : 0.f, o.b1 = 0.f, o.b2 = 0.f;
        o.instance = occluded ? best.instance : kInvalid;
        o.geometry = o.primitive = occluded ? 0u : kInvalid;
        out[h] = o;
    }
};

template<bool MULTI>
struct BlasBinding
{
    const Node8* nodes;
    const PackedTri* tris;
    uint32_t instance;
    int spFloor; // MULTI: stack height at which the current BLAS was entered, -1 while the ray is at the top level
};

template<bool ANY, bool MULTI, bool STATS, bool REFILL, typename Source>
__device__ __forceinline__ void tracePersistent(const SceneDev& scene, const Source& src, uint32_t total, uint32_t* cursor, uint32_t refill,
                                                TraceCounters& tc)
{
    initSlotTables();
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    uint32_t poolNext = 0, poolEnd = 0; // warp-uniform
    bool exhausted = false;             // warp-uniform: the launch-wide cursor has passed `total`

    RayStack stack;
#if defined(VKRT_SMEM_STACK)
    __shared__ uint2 sStack[kSmemStack * kTraceThreads];
    stack.shared = sStack + threadIdx.x;
#endif
    RayCtx c;
    BestHit best;
    BlasBinding<MULTI> bind;
    bind.nodes = scene.instances[0].nodes, bind.tris = scene.instances[0].tris, bind.instance = 0u, bind.spFloor = 0;
    TlasRefs top;
    top.nodes = scene.topNodes, top.leafInstance = scene.leafInstance, top.instances = scene.instances;
    top.instanceCount = scene.instanceCount, top.cullMask = scene.cullMask, top.boxScale = scene.topBoxScale;
    TraceCounters* const counters = STATS ? &tc : nullptr;
    uint32_t handle = 0;
    bool active = false;

    for (;;)
    {
        const uint32_t idle = __ballot_sync(0xFFFFFFFFu, !active);
        if (idle != 0u && !exhausted)
        {
            const uint32_t need = __popc(idle), avail = poolEnd - poolNext;
            uint32_t newBase = 0, chunk = 0;
            if (avail < need)
            {
                // guided self-scheduling: a chunk is 1/16 of an even share of what is LEFT (32 .. 512 rays), so chunks shrink
                // towards the end of the launch and the last warps finish within about one 32-ray batch of each other.  With
                // the fixed chunk this replaced, a rank of the 8-GPU split lost ~6 % of k_trace_closest to that tail.
                if (lane == 0)
                {
                    const uint32_t seen = *reinterpret_cast<volatile uint32_t*>(cursor);
                    const uint32_t left = total > seen ? total - seen : 0u;
                    chunk = (left / (warpsTotal * 16u) + 31u) & ~31u;
                    chunk = min(max(chunk, 32u), 512u);
                    newBase = atomicAdd(cursor, chunk);
                }
                newBase = __shfl_sync(0xFFFFFFFFu, newBase, 0);
                chunk = __shfl_sync(0xFFFFFFFFu, chunk, 0);
            }
            const uint32_t k = __popc(idle & ((1u << lane) - 1u));
            const uint32_t idx = k < avail ? poolNext + k : newBase + (k - avail);
            if (avail < need)
            {
                poolNext = newBase + (need - avail), poolEnd = newBase + chunk;
                exhausted = newBase >= total;
            }
            else
            {
                poolNext += need;
            }
            if (!active && idx < total)
            {
                handle = src.map(idx);
                F3 o, d;
                float tmin, tmax;
                src.load(handle, o, d, tmin, tmax);
                best.t = tmax, best.b1 = best.b2 = 0.0f;
                best.tri = kInvalid, best.instance = kInvalid, best.buildIndex = kInvalid;
                bool ok = validRay(o, d) && tmax > tmin;
                if (ok)
                {
                    if (MULTI)
                    {
                        beginTop(c, o, d, tmin, 0, scene.topBoxScale);
                        c.ngroup = make_uint2(0u, 0x80000000u); // the top-level root
                        bind.spFloor = -1;
                    }
                    else
                    {
                        const InstanceRecord& inst = scene.instances[0];
                        const F3 oo = affinePoint(inst.w2o, o), dd = affineDir(inst.w2o, d);
                        ok = validRay(oo, dd) && (inst.mask & scene.cullMask) != 0u;
                        if (ok)
                        {
                            beginTraversal(c, oo, dd, tmin, inst.boxScale);
                        }
                    }
                }
                if (ok)
                {
                    active = true;
                }
                else
                {
                    src.commit(handle, best, false);
                }
            }
        }
        if (__ballot_sync(0xFFFFFFFFu, active) == 0u)
        {
            if (exhausted)
            {
                break;
            }
            continue;
        }
        if (active)
        {
            for (;;)
            {
                if (MULTI && bind.spFloor < 0)
                {
                    // top level: on to the next instance whose box the ray enters
                    uint32_t idx = 0;
                    if (topStep(c, stack, top, best.t, counters, scene.one, scene.errorFlags, idx) == kTopDone)
                    {
                        src.commit(handle, best, best.tri != kInvalid);
                        active = false;
                        break;
                    }
                    F3 o, d;
                    float tmin, tmax;
                    src.load(handle, o, d, tmin, tmax);
                    const InstanceRecord& inst = scene.instances[idx];
                    int floor = 0;
                    if (enterInstanceTop(c, stack, inst, o, d, tmin, counters, scene.errorFlags, floor))
                    {
                        bind.instance = idx, bind.nodes = inst.nodes, bind.tris = inst.tris, bind.spFloor = floor;
                    }
                    continue;
                }
                const int r = traversalStep<ANY>(c, stack, bind.nodes, bind.tris, bind.instance, best, counters, scene.one,
                                                 MULTI ? bind.spFloor : 0, scene.errorFlags);
                if (r & kStepFound)
                {
                    src.commit(handle, best, true);
                    active = false;
                    break;
                }
                if (r & kStepDone)
                {
                    if (MULTI)
                    {
                        // back to the top level: the world-space box state is recomputed from the ray, the suspended
                        // groups are on the stack below spFloor
                        F3 o, d;
                        float tmin, tmax;
                        src.load(handle, o, d, tmin, tmax);
                        beginTop(c, o, d, tmin, bind.spFloor, scene.topBoxScale);
                        bind.spFloor = -1;
                        continue;
                    }
                    src.commit(handle, best, best.tri != kInvalid);
                    active = false;
                    break;
                }
                if (REFILL && !exhausted && (uint32_t)__popc(__activemask()) < refill)
                {
                    break; // too few lanes still traversing: leave so that the idle lanes can take new rays
                }
            }
        }
    }
}

template<bool STATS>
__device__ __forceinline__ void flushTraceCounters(const TraceCounters& tc, unsigned long long* statCounters, int nodeSlot)
{
    if (STATS)
    {
        atomicAdd(&statCounters[nodeSlot], (unsigned long long)tc.nodes);
        atomicAdd(&statCounters[nodeSlot + 1], (unsigned long long)tc.tris);
        if (tc.overflow)
        {
            atomicAdd(&statCounters[5], 1ull);
        }
    }
}

// MODE: 0 plain, 1 counting nodes/triangles (stats pass), 2 per-lane ray replacement
template<int MODE, bool MULTI>
__global__ void __launch_bounds__(kTraceThreads, VKRT_TRACE_MIN_BLOCKS)
k_trace_closest(SceneDev scene, const float4* __restrict__ rayO, const float4* __restrict__ rayD, const uint32_t* __restrict__ countPtr,
                uint32_t countImm, float4* __restrict__ hitA, uint32_t* __restrict__ hitInst, unsigned long long* statCounters,
                int statSlot, uint32_t* cursor, uint32_t refill)
{
    constexpr bool STATS = MODE == 1;
    const uint32_t count = countPtr ? *countPtr : countImm;
    TraceCounters tc;
    ClosestSource src;
    src.rayO = rayO, src.rayD = rayD, src.hitA = hitA, src.hitInst = hitInst;
    tracePersistent<false, MULTI, STATS, MODE == 2>(scene, src, count, cursor, refill, tc);
    flushTraceCounters<STATS>(tc, statCounters, 3);
    if (statSlot >= 0 && blockIdx.x == 0 && threadIdx.x == 0 && countPtr)
    {
        atomicAdd(&statCounters[statSlot], (unsigned long long)count);
    }
}

// shader.rchit:88-129 for every hit; shader.rmiss:17 for every miss.  Emits shadow rays.  LOD = true is the ray-cone
// texture mode (vkrt_trace_set_texture_lod); the default instantiation is the reference's LOD-0 sampling.
template<bool LOD>
__global__ void __launch_bounds__(kShadeWarps * 32)
k_shade_surface(const __grid_constant__ VkrtUniforms u, SceneDev scene, WaveDev w, FrameDev f, const float4* __restrict__ rayD,
                const uint32_t* __restrict__ pathOf, const uint32_t* __restrict__ countPtr, uint32_t countImm, int depth, uint32_t aovSamples)
{
    __shared__ AppendStage stage;
    // c / 255.0f for every byte value: the bilinear filter converts 16 texel bytes per texture() call, each an IEEE division
    // (~10 instructions); the table holds the identical quotients (bit-equal results, as in the rasteriser and the mip kernel)
    __shared__ float unorm[256];
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x)
    {
        unorm[i] = (float)i / 255.0f;
    }
    __syncthreads();
    const uint32_t count = countPtr ? *countPtr : countImm;
    const uint32_t rounded = (count + blockDim.x - 1u) / blockDim.x * blockDim.x; // whole blocks stay in the loop (blockAppend4)
    int parity = 0;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < rounded; q += gridDim.x * blockDim.x, parity ^= 1)
    {
        bool active = q < count;
        uint32_t p = 0;
        float4 h = make_float4(0.f, 0.f, 0.f, 0.f);
        if (active)
        {
            p = pathOf ? pathOf[q] : q;
            active = rayD[q].w >= 0.0f; // padded pixels of partial tiles
        }
        bool hit = false;
        SurfacePoint sp;
        LightTerm lt[4];
        uint32_t castMask = 0;
        uint32_t instance = 0, primitive = 0;
        if (active)
        {
            h = w.hitA[q];
            const uint32_t tri = __float_as_uint(h.w);
            hit = tri != kInvalid;
            if (!hit)
            {
                float4 acc = w.pathAccum[p];
                acc.x += 0.8f, acc.y += 0.8f, acc.z += 1.0f; // rmiss:17, rgen:63
                w.pathAccum[p] = acc;
                w.surfGeom[q] = kInvalid;
                if (depth == 0 && aovSamples && p % aovSamples == 0) // sample 0 of the frame only
                {
                    uint32_t x, y, tl, pit;
                    pathToPixel(f.map, p / aovSamples, x, y, tl, pit);
                    f.triangleId[(size_t)y * f.map.width + x] = kInvalid;
                    f.shadowMask[(size_t)y * f.map.width + x] = 0;
                }
            }
            else
            {
                instance = w.hitInst[q];
                const InstanceRecord& inst = scene.instances[instance];
                const float4 tag = __ldg(reinterpret_cast<const float4*>(inst.tris + tri) + 2);
                const uint32_t geometry = __float_as_uint(tag.y);
                primitive = __float_as_uint(tag.z);
                if (LOD)
                {
                    const float4 d4 = rayD[q];
                    float travelled = h.x; // cone width = spread angle x distance along the path (surface spread ignored)
                    if (depth == 0)
                    {
                        w.coneT[q] = h.x;
                    }
                    else
                    {
                        travelled = w.coneT[p] + h.x;
                    }
                    LodInput lod;
                    lod.mips = scene.mips, lod.worldRayDir = f3(d4.x, d4.y, d4.z);
                    lod.coneWidth = pixelSpreadAngle(u, f.map.height) * travelled;
                    sp = shadeSurface<true>(inst, scene.rows, scene.textures, geometry, primitive, h.y, h.z, &lod, unorm);
                    w.surfFoot[q] = sp.footprint2;
                }
                else
                {
                    sp = shadeSurface<false>(inst, scene.rows, scene.textures, geometry, primitive, h.y, h.z, nullptr, unorm);
                }
                for (int i = 0; i < 4; ++i)
                {
                    lt[i] = lightTerm(u, i, sp.worldPos, sp.normal);
                    if (lt[i].castShadow)
                    {
                        castMask |= 1u << i;
                    }
                }
                w.surfA[q] = make_float4(sp.worldPos.x, sp.worldPos.y, sp.worldPos.z, sp.u);
                w.surfB[q] = make_float4(sp.normal.x, sp.normal.y, sp.normal.z, sp.v);
                w.surfGeom[q] = geometry;
                w.shadowBits[q] = castMask << 4;
                if (depth == 0 && aovSamples && p % aovSamples == 0) // sample 0 of the frame only
                {
                    uint32_t x, y, tl, pit;
                    pathToPixel(f.map, p / aovSamples, x, y, tl, pit);
                    f.triangleId[(size_t)y * f.map.width + x] =
                        (instance << 24) + (uint32_t)scene.rows[geometry].indexBufferOffset + primitive;
                }
            }
        }
        // compaction into the 4 per-light shadow queues (light-major keeps rays towards one light together)
        uint32_t slot[4];
        blockAppend4(stage, w.counters, hit ? castMask : 0u, parity, slot);
#pragma unroll
        for (int i = 0; i < 4; ++i)
        {
            if (hit && ((castMask >> i) & 1u))
            {
                const size_t dst = (size_t)i * w.capacity + slot[i];
                w.shadowO[dst] = make_float4(sp.worldPos.x, sp.worldPos.y, sp.worldPos.z, lt[i].distance);
                w.shadowD[dst] = make_float4(lt[i].dir.x, lt[i].dir.y, lt[i].dir.z, __uint_as_float(q));
            }
        }
    }
}

// shader.rchit:130-143 + shader_shadow.rmiss: any-hit rays, tmin 0.001, tmax = light distance
template<int MODE, bool MULTI>
__global__ void __launch_bounds__(kTraceThreads, VKRT_TRACE_MIN_BLOCKS)
k_trace_any(SceneDev scene, WaveDev w, unsigned long long* statCounters, uint32_t* cursor, uint32_t refill, uint32_t bounce)
{
    constexpr bool STATS = MODE == 1;
    const uint32_t c0 = w.counters[0], c1 = w.counters[1], c2 = w.counters[2], c3 = w.counters[3];
    const uint32_t total = c0 + c1 + c2 + c3;
    TraceCounters tc;
    ShadowSource src;
    src.shadowO = w.shadowO, src.shadowD = w.shadowD, src.shadowBits = w.shadowBits;
    src.c0 = c0, src.c01 = c0 + c1, src.c012 = c0 + c1 + c2, src.capacity = w.capacity;
    tracePersistent<true, MULTI, STATS, MODE == 2>(scene, src, total, cursor, refill, tc);
    flushTraceCounters<STATS>(tc, statCounters, 6);
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        atomicAdd(&statCounters[1], (unsigned long long)total);
        if (bounce)
        {
            atomicAdd(&statCounters[8], (unsigned long long)total); // shadow rays cast from reflection hits
        }
    }
}

// shader.rchit:144-171 and the bounce bookkeeping of shader.rgen:63-73
template<bool LOD>
__global__ void __launch_bounds__(kShadeThreads)
k_shade_resolve(const __grid_constant__ VkrtUniforms u, SceneDev scene, WaveDev w, FrameDev f, const float4* __restrict__ rayD,
                const uint32_t* __restrict__ pathOf, const uint32_t* __restrict__ countPtr, uint32_t countImm, int depth, uint32_t aovSamples)
{
    __shared__ float unorm[256]; // see k_shade_surface
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x)
    {
        unorm[i] = (float)i / 255.0f;
    }
    __syncthreads();
    const uint32_t count = countPtr ? *countPtr : countImm;
    const uint32_t rounded = (count + 31u) & ~31u;
    for (uint32_t q = blockIdx.x * blockDim.x + threadIdx.x; q < rounded; q += gridDim.x * blockDim.x)
    {
        bool active = q < count;
        uint32_t geometry = kInvalid;
        if (active)
        {
            geometry = w.surfGeom[q];
            active = geometry != kInvalid && rayD[q].w >= 0.0f;
        }
        bool reflect = false;
        uint32_t p = 0;
        F3 origin = f3(0.f, 0.f, 0.f), dir = f3(0.f, 0.f, 0.f);
        if (active)
        {
            p = pathOf ? pathOf[q] : q;
            const float4 a = w.surfA[q], b = w.surfB[q];
            SurfacePoint sp;
            sp.worldPos = f3(a.x, a.y, a.z), sp.normal = f3(b.x, b.y, b.z);
            sp.u = a.w, sp.v = b.w, sp.geometry = geometry;
            sp.footprint2 = LOD ? w.surfFoot[q] : 0.0f;
            const uint32_t bits = w.shadowBits[q];
            float4 acc = w.pathAccum[p];
            const float4 d4 = rayD[q];
            const ResolveOut r = shadeResolve<LOD>(u, scene.rows, scene.textures, sp, bits & 0xFu, acc.w, f3(d4.x, d4.y, d4.z), scene.mips, unorm);
            acc.x += r.hitValue.x, acc.y += r.hitValue.y, acc.z += r.hitValue.z; // rgen:63
            acc.w = r.attenuation;
            w.pathAccum[p] = acc;
            reflect = r.reflect && depth == 0; // rgen:65-69: maxDepth 2
            origin = sp.worldPos, dir = r.rayDir;
            if (depth == 0 && aovSamples && p % aovSamples == 0) // sample 0 of the frame only
            {
                uint32_t x, y, tl, pit;
                pathToPixel(f.map, p / aovSamples, x, y, tl, pit);
                f.shadowMask[(size_t)y * f.map.width + x] = (uint8_t)bits;
            }
        }
        const uint32_t slot = warpAppend(&w.counters[4], reflect);
        if (reflect)
        {
            w.reflO[slot] = make_float4(origin.x, origin.y, origin.z, kRayTMin);
            w.reflD[slot] = make_float4(dir.x, dir.y, dir.z, kRayTMax);
            w.reflPath[slot] = p;
        }
    }
}

// shader.rgen:76 imageStore + sample averaging; writes the linear image and/or the tile-major send buffer
__global__ void __launch_bounds__(kShadeThreads)
k_resolve_sample(WaveDev w, FrameDev f, uint32_t paths, uint32_t sampleBase, uint32_t samples, uint32_t spp)
{
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < paths; p += gridDim.x * blockDim.x)
    {
        uint32_t x, y, tl, pit;
        const bool valid = pathToPixel(f.map, p, x, y, tl, pit);
        // samples are summed in sample order (the running sum crosses waves through sampleSum), then divided by spp
        F3 sum = f3(0.f, 0.f, 0.f);
        for (uint32_t j = 0; j < samples; ++j)
        {
            const float4 acc = w.pathAccum[(size_t)p * samples + j];
            const F3 a = f3(acc.x, acc.y, acc.z);
            if (j == 0 && sampleBase == 0)
            {
                sum = a;
            }
            else if (j == 0)
            {
                const float4 sprev = w.sampleSum[p];
                sum = f3(sprev.x, sprev.y, sprev.z) + a;
            }
            else
            {
                sum = sum + a;
            }
        }
        if (spp > 1)
        {
            if (sampleBase + samples < spp)
            {
                w.sampleSum[p] = make_float4(sum.x, sum.y, sum.z, 0.f);
                continue;
            }
            sum = sum / (float)spp;
        }
        const uint32_t px = packBgra8(sum);
        if (f.tileBgra8)
        {
            f.tileBgra8[(size_t)tl * kTilePixels + pit] = valid ? px : 0u;
        }
        if (valid)
        {
            if (f.bgra8)
            {
                f.bgra8[(size_t)y * f.map.width + x] = px;
            }
            if (f.rgb32f)
            {
                float* dst = f.rgb32f + ((size_t)y * f.map.width + x) * 3;
                dst[0] = sum.x, dst[1] = sum.y, dst[2] = sum.z;
            }
        }
    }
}

__global__ void k_hits_export(SceneDev scene, const float4* __restrict__ hitA, const uint32_t* __restrict__ hitInst, uint32_t count,
                              VkrtHit* __restrict__ out)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const float4 h = hitA[i];
        const uint32_t tri = __float_as_uint(h.w);
        VkrtHit o;
        if (tri == kInvalid)
        {
            o.t = 0.f, o.b1 = 0.f, o.b2 = 0.f;
            o.instance = o.geometry = o.primitive = kInvalid;
        }
        else
        {
            const uint32_t inst = hitInst[i];
            const float4 tag = reinterpret_cast<const float4*>(scene.instances[inst].tris + tri)[2];
            o.t = h.x, o.b1 = h.y, o.b2 = h.z;
            o.instance = inst;
            o.geometry = __float_as_uint(tag.y);
            o.primitive = __float_as_uint(tag.z);
        }
        out[i] = o;
    }
}

// any-hit over a plain ray list (ray-level API): writes instance = 0 / kInvalid
template<bool STATS, bool MULTI>
__global__ void __launch_bounds__(kTraceThreads)
k_trace_any_list(SceneDev scene, const float4* __restrict__ rayO, const float4* __restrict__ rayD, uint32_t count, VkrtHit* __restrict__ out,
                 unsigned long long* statCounters, uint32_t* cursor)
{
    TraceCounters tc;
    ListAnySource src;
    src.rayO = rayO, src.rayD = rayD, src.out = out;
    tracePersistent<true, MULTI, STATS, false>(scene, src, count, cursor, 0u, tc);
    flushTraceCounters<STATS>(tc, statCounters, 6);
}

int reserveWave(Context* ctx, uint32_t paths)
{
    WaveBuffers& w = ctx->wave;
    const size_t p = std::max<size_t>(paths, 32);
    VKRT_CUDA_TRY(ctx, w.rayO.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.rayD.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.reflO.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.reflD.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.reflPath.reserve(p * 4));
    VKRT_CUDA_TRY(ctx, w.hitA.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.hitInst.reserve(p * 4));
    VKRT_CUDA_TRY(ctx, w.surfA.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.surfB.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.surfGeom.reserve(p * 4));
    VKRT_CUDA_TRY(ctx, w.shadowBits.reserve(p * 4));
    VKRT_CUDA_TRY(ctx, w.shadowO.reserve(p * 16 * 4));
    VKRT_CUDA_TRY(ctx, w.shadowD.reserve(p * 16 * 4));
    VKRT_CUDA_TRY(ctx, w.pathAccum.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.sampleSum.reserve(p * 16));
    VKRT_CUDA_TRY(ctx, w.counters.reserve(64));
    VKRT_CUDA_TRY(ctx, ctx->statCounters.reserve(128));
    if (ctx->textureLod == VKRT_TEXTURE_LOD_RAY_CONES)
    {
        VKRT_CUDA_TRY(ctx, w.coneT.reserve(p * 4));
        VKRT_CUDA_TRY(ctx, w.surfFoot.reserve(p * 4));
    }
    return VKRT_SUCCESS;
}

WaveDev waveDev(Context* ctx, uint32_t paths)
{
    WaveBuffers& b = ctx->wave;
    WaveDev w;
    w.rayO = b.rayO.as<float4>(), w.rayD = b.rayD.as<float4>();
    w.reflO = b.reflO.as<float4>(), w.reflD = b.reflD.as<float4>();
    w.reflPath = b.reflPath.as<uint32_t>();
    w.hitA = b.hitA.as<float4>(), w.hitInst = b.hitInst.as<uint32_t>();
    w.surfA = b.surfA.as<float4>(), w.surfB = b.surfB.as<float4>();
    w.surfGeom = b.surfGeom.as<uint32_t>(), w.shadowBits = b.shadowBits.as<uint32_t>();
    w.shadowO = b.shadowO.as<float4>(), w.shadowD = b.shadowD.as<float4>();
    w.pathAccum = b.pathAccum.as<float4>(), w.sampleSum = b.sampleSum.as<float4>();
    w.counters = b.counters.as<uint32_t>();
    w.coneT = b.coneT.as<float>(), w.surfFoot = b.surfFoot.as<float>();
    w.capacity = std::max<uint32_t>(paths, 32);
    return w;
}

uint32_t persistentGrid(const Context* ctx, int blocksPerSm)
{
    return (uint32_t)(ctx->smCount * blocksPerSm);
}

void launchClosest(Context* ctx, const SceneDev& scene, uint32_t grid, const float4* o, const float4* d, const uint32_t* countPtr,
                   uint32_t countImm, float4* hitA, uint32_t* hitInst, int statSlot, uint32_t* cursor)
{
    unsigned long long* stat = ctx->statCounters.as<unsigned long long>();
    const bool st = ctx->countTraversal, multi = scene.topNodes != nullptr;
    cudaStream_t s = ctx->stream;
    const uint32_t rf = ctx->refillLanes;
#define VKRT_LAUNCH(S, M) k_trace_closest<S, M><<<grid, kTraceThreads, 0, s>>>(scene, o, d, countPtr, countImm, hitA, hitInst, stat, statSlot, cursor, rf)
    if (st)
    {
        multi ? VKRT_LAUNCH(1, true) : VKRT_LAUNCH(1, false);
    }
    else if (rf)
    {
        multi ? VKRT_LAUNCH(2, true) : VKRT_LAUNCH(2, false);
    }
    else
    {
        multi ? VKRT_LAUNCH(0, true) : VKRT_LAUNCH(0, false);
    }
#undef VKRT_LAUNCH
}

void launchAny(Context* ctx, const SceneDev& scene, uint32_t grid, const WaveDev& w, uint32_t* cursor, uint32_t bounce)
{
    unsigned long long* stat = ctx->statCounters.as<unsigned long long>();
    const bool st = ctx->countTraversal, multi = scene.topNodes != nullptr;
    cudaStream_t s = ctx->stream;
    const uint32_t rf = ctx->refillLanes;
#define VKRT_LAUNCH(S, M) k_trace_any<S, M><<<grid, kTraceThreads, 0, s>>>(scene, w, stat, cursor, rf, bounce)
    if (st)
    {
        multi ? VKRT_LAUNCH(1, true) : VKRT_LAUNCH(1, false);
    }
    else if (rf)
    {
        multi ? VKRT_LAUNCH(2, true) : VKRT_LAUNCH(2, false);
    }
    else
    {
        multi ? VKRT_LAUNCH(0, true) : VKRT_LAUNCH(0, false);
    }
#undef VKRT_LAUNCH
}

void launchAnyList(Context* ctx, const SceneDev& scene, uint32_t grid, const float4* o, const float4* d, uint32_t n, VkrtHit* out,
                   uint32_t* cursor)
{
    unsigned long long* stat = ctx->statCounters.as<unsigned long long>();
    const bool st = ctx->countTraversal, multi = scene.topNodes != nullptr;
    cudaStream_t s = ctx->stream;
#define VKRT_LAUNCH(S, M) k_trace_any_list<S, M><<<grid, kTraceThreads, 0, s>>>(scene, o, d, n, out, stat, cursor)
    if (st)
    {
        multi ? VKRT_LAUNCH(true, true) : VKRT_LAUNCH(true, false);
    }
    else
    {
        multi ? VKRT_LAUNCH(false, true) : VKRT_LAUNCH(false, false);
    }
#undef VKRT_LAUNCH
}
} // namespace

int waitPresents(Context* ctx, Frame* f)
{
    if (f->presentNext > f->presentWaited)
    {
        // copies run in order on one stream: the newest event covers all of them
        VKRT_CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, f->presentDone[(f->presentNext - 1) % Frame::kPresentRing], 0));
    }
    return VKRT_SUCCESS;
}

int traceFrame(Context* ctx, const Tlas* tlas, const VkrtUniforms* uniforms, uint32_t spp, uint32_t frameIndex, uint32_t flags,
               Frame* frame)
{
    (void)frameIndex;
    if (spp == 0)
    {
        spp = 1;
    }
    if (int rc = checkSceneForShading(ctx, tlas, "vkrt_trace"))
    {
        return rc;
    }
    const uint32_t paths = frame->paths;
    // Samples per wave: all samples of a pixel that fit the queue budget go through the pipeline together, so a frame
    // costs 12 launches per WAVE instead of per sample (at 8 GPUs the per-launch ramp and tail were 10 % of a sample).
    uint32_t waveSamples = 1;
    if (ctx->waveKeyPaths == paths && ctx->waveKeySpp == spp)
    {
        waveSamples = ctx->waveSamples; // cudaMemGetInfo costs up to milliseconds: ask once per frame shape
    }
    else
    {
        size_t freeBytes = 0, totalBytes = 0;
        cudaMemGetInfo(&freeBytes, &totalBytes);
        const size_t held = ctx->wave.rayO.capacity * 18; // the wave buffers already reserved count as available
        const size_t budget = std::min<size_t>((size_t)24 << 30, (freeBytes + held) / 3);
        const size_t perSample = (size_t)std::max<uint32_t>(paths, 32) * 288;
        size_t fit = std::max<size_t>(1, budget / perSample);
        fit = std::min<size_t>(fit, (size_t)0x0FFFFFFF / std::max<uint32_t>(paths, 1)); // queue indices stay below 2^30
        const char* e = getenv("VKRT_WAVE_SAMPLES");
        if (e && atoi(e) > 0)
        {
            fit = (size_t)atoi(e);
        }
        waveSamples = (uint32_t)std::max<size_t>(1, std::min<size_t>(fit, spp));
        ctx->waveKeyPaths = paths, ctx->waveKeySpp = spp, ctx->waveSamples = waveSamples;
    }
    ctx->stats.samplesPerWave = waveSamples;
    int r = reserveWave(ctx, paths * waveSamples);
    if (r != VKRT_SUCCESS)
    {
        return r;
    }
    cudaStream_t s = ctx->stream;
    const bool aov = !(flags & VKRT_TRACE_NO_AOV);
    if (aov)
    {
        const size_t px = (size_t)frame->width * frame->height;
        VKRT_CUDA_TRY(ctx, frame->rgb32f.reserve(px * 12));
        VKRT_CUDA_TRY(ctx, frame->triangleId.reserve(px * 4));
        VKRT_CUDA_TRY(ctx, frame->shadowMask.reserve(px));
    }

    SceneDev scene;
    scene.instances = tlas->instances.as<InstanceRecord>();
    scene.topNodes = tlas->nodeCount ? tlas->nodes.as<Node8>() : nullptr;
    scene.leafInstance = tlas->leafInstance.as<uint32_t>();
    scene.instanceCount = tlas->instanceCount;
    scene.cullMask = 0xFFu;
    scene.errorFlags = ctx->errorFlags.as<uint32_t>();
    scene.topBoxScale = tlas->boxScale;
    scene.rows = ctx->materials.as<VkrtSubmeshInfo>();
    scene.textures = ctx->textureRecords.as<TextureRecord>();
    scene.one = 0x3F800000u;
    const bool lod = ctx->textureLod == VKRT_TEXTURE_LOD_RAY_CONES;
    if (lod && !ctx->mipsReady)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_trace: VKRT_TEXTURE_LOD_RAY_CONES needs vkrt_textures_generate_mips first");
    }
    scene.mips = lod ? ctx->textureMips.as<TextureMips>() : nullptr;

    WaveDev w = waveDev(ctx, paths * waveSamples);
    FrameDev f;
    f.map.width = frame->width, f.map.height = frame->height;
    f.map.tilesX = frame->tilesX, f.map.tilesY = frame->tilesY;
    f.map.rank = (uint32_t)frame->rank, f.map.size = (uint32_t)frame->size;
    f.map.ownedTiles = frame->ownedTiles;
    f.bgra8 = frame->bgra8.as<uint32_t>();
    f.tileBgra8 = frame->size > 1 ? frame->tileBgra8.as<uint32_t>() : nullptr;
    f.rgb32f = aov ? frame->rgb32f.as<float>() : nullptr;
    f.triangleId = aov ? frame->triangleId.as<uint32_t>() : nullptr;
    f.shadowMask = aov ? frame->shadowMask.as<uint8_t>() : nullptr;

    const VkrtUniforms u = *uniforms;
    unsigned long long* stat = ctx->statCounters.as<unsigned long long>();
    const uint32_t gridTrace = persistentGrid(ctx, VKRT_TRACE_MIN_BLOCKS); // exactly the resident blocks: a persistent kernel gains nothing from more
    const uint32_t gridShade = persistentGrid(ctx, 8);
    const uint32_t gridSurface = persistentGrid(ctx, 4); // 256-thread blocks at 64 registers: 4 per SM
    const bool st = ctx->countTraversal;

    // valid (non-padded) primary rays owned by this rank
    uint64_t validPixels = 0;
    for (uint32_t t = 0; t < frame->ownedTiles; ++t)
    {
        const uint32_t tile = (uint32_t)frame->rank + t * (uint32_t)frame->size;
        const uint32_t tx = tile % frame->tilesX, ty = tile / frame->tilesX;
        const uint32_t wpx = std::min<uint32_t>(kTile, frame->width - tx * kTile), hpx = std::min<uint32_t>(kTile, frame->height - ty * kTile);
        validPixels += (uint64_t)wpx * hpx;
    }

    int ev = 0;
    auto mark = [&]() {
        if (ctx->timeStages && ev < 16)
        {
            cudaEventRecord(ctx->stageEvents[ev++], s);
        }
    };
    if (ctx->timeStages && !ctx->eventsCreated)
    {
        for (auto& e : ctx->stageEvents)
        {
            cudaEventCreate(&e);
        }
        ctx->eventsCreated = true;
    }
    ctx->stats.msRaygen = ctx->stats.msTraceClosest = ctx->stats.msShade = ctx->stats.msTraceAny = ctx->stats.msResolve = 0.f;
    ctx->stats.msTotal = 0.f;

    for (uint32_t sample = 0; sample < spp; sample += waveSamples)
    {
        const uint32_t samples = std::min(waveSamples, spp - sample);
        const uint32_t wavePaths = paths * samples;
        const uint32_t aovSamples = (aov && sample == 0) ? samples : 0u; // AOVs describe sample 0 = every samples-th entry
        const bool timed = ctx->timeStages && sample == 0;
        if (timed)
        {
            mark(); // 0
        }
        k_reset_counters<<<1, 32, 0, s>>>(w.counters);
        k_raygen<<<gridShade, kShadeThreads, 0, s>>>(u, f.map, paths, sample, samples, spp, w.rayO, w.rayD, w.pathAccum);
        if (timed)
        {
            mark(); // 1
        }
        // depth 0: primary rays
        launchClosest(ctx, scene, gridTrace, w.rayO, w.rayD, nullptr, wavePaths, w.hitA, w.hitInst, -1, w.counters + 8);
        if (timed)
        {
            mark(); // 2
        }
        if (lod)
        {
            k_shade_surface<true><<<gridSurface, kShadeWarps * 32, 0, s>>>(u, scene, w, f, w.rayD, nullptr, nullptr, wavePaths, 0, aovSamples);
        }
        else
        {
            k_shade_surface<false><<<gridSurface, kShadeWarps * 32, 0, s>>>(u, scene, w, f, w.rayD, nullptr, nullptr, wavePaths, 0, aovSamples);
        }
        if (timed)
        {
            mark(); // 3
        }
        launchAny(ctx, scene, gridTrace, w, w.counters + 9, 0u);
        if (timed)
        {
            mark(); // 4
        }
        if (lod)
        {
            k_shade_resolve<true><<<gridShade, kShadeThreads, 0, s>>>(u, scene, w, f, w.rayD, nullptr, nullptr, wavePaths, 0, aovSamples);
        }
        else
        {
            k_shade_resolve<false><<<gridShade, kShadeThreads, 0, s>>>(u, scene, w, f, w.rayD, nullptr, nullptr, wavePaths, 0, aovSamples);
        }
        if (timed)
        {
            mark(); // 5
        }
        // depth 1: the compacted reflection queue
        k_reset_counters<<<1, 4, 0, s>>>(w.counters); // only the 4 shadow queues; [4] holds the reflection count
        launchClosest(ctx, scene, gridTrace, w.reflO, w.reflD, w.counters + 4, 0, w.hitA, w.hitInst, 2, w.counters + 10);
        if (timed)
        {
            mark(); // 6
        }
        if (lod)
        {
            k_shade_surface<true><<<gridSurface, kShadeWarps * 32, 0, s>>>(u, scene, w, f, w.reflD, w.reflPath, w.counters + 4, 0, 1, 0);
        }
        else
        {
            k_shade_surface<false><<<gridSurface, kShadeWarps * 32, 0, s>>>(u, scene, w, f, w.reflD, w.reflPath, w.counters + 4, 0, 1, 0);
        }
        if (timed)
        {
            mark(); // 7
        }
        launchAny(ctx, scene, gridTrace, w, w.counters + 11, 1u);
        if (timed)
        {
            mark(); // 8
        }
        if (lod)
        {
            k_shade_resolve<true><<<gridShade, kShadeThreads, 0, s>>>(u, scene, w, f, w.reflD, w.reflPath, w.counters + 4, 0, 1, 0);
        }
        else
        {
            k_shade_resolve<false><<<gridShade, kShadeThreads, 0, s>>>(u, scene, w, f, w.reflD, w.reflPath, w.counters + 4, 0, 1, 0);
        }
        if (timed)
        {
            mark(); // 9
        }
        if (int r = joinGather(ctx, frame)) // ... and an outstanding tile gather still reads the tile buffer / writes the image
        {
            return r;
        }
        if (int r = waitPresents(ctx, frame)) // the image is about to be overwritten: outstanding vkrt_frame_present copies first
        {
            return r;
        }
        k_resolve_sample<<<gridShade, kShadeThreads, 0, s>>>(w, f, paths, sample, samples, spp);
        if (timed)
        {
            mark(); // 10
        }
        ctx->launches += 12;
    }
    ctx->stageEventCount = ev;
    ctx->stats.raysPrimary += validPixels * spp;
    frame->gathered = false;
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    return VKRT_SUCCESS;
}

int traceRaysHost(Context* ctx, const Tlas* tlas, const VkrtRay* rays, uint64_t count, int anyHit, VkrtHit* hits)
{
    if (count == 0)
    {
        return VKRT_SUCCESS;
    }
    if (count > 0x7FFFFFFFull)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_rays: at most 2^31-1 rays per call");
    }
    const uint32_t n = (uint32_t)count;
    if (int rc = checkTlasCurrent(ctx, tlas, "vkrt_trace_rays"))
    {
        return rc;
    }
    int r = reserveWave(ctx, n);
    if (r != VKRT_SUCCESS)
    {
        return r;
    }
    cudaStream_t s = ctx->stream;
    // split the 32-byte host records into the engine's two float4 streams on the host side
    std::vector<float4> o(n), d(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        o[i] = make_float4(rays[i].origin[0], rays[i].origin[1], rays[i].origin[2], rays[i].tmin);
        d[i] = make_float4(rays[i].direction[0], rays[i].direction[1], rays[i].direction[2], rays[i].tmax);
    }
    WaveDev w = waveDev(ctx, n);
    DeviceBuffer out;
    VKRT_CUDA_TRY(ctx, out.reserve((size_t)n * sizeof(VkrtHit)));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(w.rayO, o.data(), (size_t)n * 16, cudaMemcpyHostToDevice, s));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(w.rayD, d.data(), (size_t)n * 16, cudaMemcpyHostToDevice, s));
    SceneDev scene;
    scene.instances = tlas->instances.as<InstanceRecord>();
    scene.topNodes = tlas->nodeCount ? tlas->nodes.as<Node8>() : nullptr;
    scene.leafInstance = tlas->leafInstance.as<uint32_t>();
    scene.instanceCount = tlas->instanceCount;
    scene.cullMask = 0xFFu;
    scene.errorFlags = ctx->errorFlags.as<uint32_t>();
    scene.topBoxScale = tlas->boxScale;
    scene.rows = nullptr;
    scene.textures = nullptr;
    scene.one = 0x3F800000u;
    scene.mips = nullptr;
    unsigned long long* stat = ctx->statCounters.as<unsigned long long>();
    const uint32_t grid = persistentGrid(ctx, 8);
    k_reset_counters<<<1, 32, 0, s>>>(w.counters);
    if (anyHit)
    {
        launchAnyList(ctx, scene, grid, w.rayO, w.rayD, n, out.as<VkrtHit>(), w.counters + 8);
        ctx->launches += 2;
    }
    else
    {
        launchClosest(ctx, scene, grid, w.rayO, w.rayD, nullptr, n, w.hitA, w.hitInst, -1, w.counters + 8);
        k_hits_export<<<grid, 128, 0, s>>>(scene, w.hitA, w.hitInst, n, out.as<VkrtHit>());
        ctx->launches += 3;
    }
    cudaError_t e = cudaMemcpyAsync(hits, out.ptr, (size_t)n * sizeof(VkrtHit), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess)
    {
        e = cudaStreamSynchronize(s);
    }
    out.release();
    if (e != cudaSuccess)
    {
        return ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_trace_rays: ") + cudaGetErrorString(e));
    }
    return checkDeviceErrors(ctx, "vkrt_trace_rays");
}

// ------------------------------------------------------------------------------------------------------------------
// vkrt_trace_rays_static: the ray-level entry without any allocation (include/vkrt.h).  No counterpart in the
// reference; added for callers that cannot call cudaMalloc.  Everything the call
// needs on the device besides the caller's scene lives in one file-scope array; the kernel is the persistent trace
// kernel over a source that reads the 32-byte ray records and writes the 24-byte hit records in place.
// ------------------------------------------------------------------------------------------------------------------
namespace
{
struct StaticScratch
{
    uint32_t header[64];               // [0] ray-pool cursor, [4] traversal stack overflow flag; cleared by every call
    VkrtRay rays[VKRT_STATIC_MAX_RAYS]; // staged input
    VkrtHit hits[VKRT_STATIC_MAX_RAYS]; // staged output
};
static_assert(sizeof(VkrtRay) == 32 && sizeof(VkrtHit) == 24, "record sizes of include/vkrt.h");
static_assert(offsetof(StaticScratch, rays) == 256 && offsetof(StaticScratch, hits) % 16 == 0, "scratch layout");

// Held by every GPU that loads the library until the process exits.  In host code this name is only a handle for
// cudaGetSymbolAddress - never pass it to a kernel.
static __device__ __align__(256) unsigned char g_staticScratch[sizeof(StaticScratch)];

constexpr int kStaticMaxDevices = 64;
std::mutex g_staticMutex;
StaticScratch* g_staticAddress[kStaticMaxDevices]; // device address of g_staticScratch per device ordinal, resolved once

// The device of `ctx` is current.
int staticScratchAddress(Context* ctx, StaticScratch** out)
{
    if (ctx->device < 0 || ctx->device >= kStaticMaxDevices)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_rays_static: device ordinal beyond 63");
    }
    std::lock_guard<std::mutex> lock(g_staticMutex);
    if (!g_staticAddress[ctx->device])
    {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_staticScratch);
        if (e != cudaSuccess || !p)
        {
            return ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_trace_rays_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        }
        g_staticAddress[ctx->device] = static_cast<StaticScratch*>(p);
    }
    *out = g_staticAddress[ctx->device];
    return VKRT_SUCCESS;
}

template<bool ANY>
struct StaticRaySource
{
    const VkrtRay* rays;
    VkrtHit* out;
    const InstanceRecord* instances;
    __device__ __forceinline__ uint32_t map(uint32_t i) const { return i; }
    __device__ __forceinline__ void load(uint32_t h, F3& o, F3& d, float& tmin, float& tmax) const
    {
        const float4* r = reinterpret_cast<const float4*>(rays + h); // 32-byte records at a 256-byte aligned base
        const float4 o4 = r[0], d4 = r[1];
        o = f3(o4.x, o4.y, o4.z), d = f3(d4.x, d4.y, d4.z);
        tmin = o4.w, tmax = d4.w;
    }
    __device__ __forceinline__ void commit(uint32_t h, const BestHit& best, bool found) const
    {
        VkrtHit o;
        if (ANY)
        {
            // what ListAnySource::commit writes
            o.t = found ? best.t : 0.f, o.b1 = 0.f, o.b2 = 0.f;
            o.instance = found ? best.instance : kInvalid;
            o.geometry = o.primitive = found ? 0u : kInvalid;
        }
        else if (best.tri == kInvalid)
        {
            o.t = 0.f, o.b1 = 0.f, o.b2 = 0.f;
            o.instance = o.geometry = o.primitive = kInvalid;
        }
        else
        {
            // what ClosestSource::commit + k_hits_export write
            const float4 tag = reinterpret_cast<const float4*>(instances[best.instance].tris + best.tri)[2];
            o.t = best.t, o.b1 = best.b1, o.b2 = best.b2;
            o.instance = best.instance;
            o.geometry = __float_as_uint(tag.y);
            o.primitive = __float_as_uint(tag.z);
        }
        out[h] = o;
    }
};

template<bool ANY, bool MULTI>
__global__ void __launch_bounds__(kTraceThreads, VKRT_TRACE_MIN_BLOCKS)
k_trace_rays_static(SceneDev scene, StaticScratch* scratch, uint32_t count)
{
    TraceCounters tc;
    StaticRaySource<ANY> src;
    src.rays = scratch->rays, src.out = scratch->hits, src.instances = scene.instances;
    tracePersistent<ANY, MULTI, false, false>(scene, src, count, &scratch->header[0], 0u, tc);
}
} // namespace

int traceRaysStatic(Context* ctx, const Tlas* tlas, const VkrtRay* rays, uint32_t count, int anyHit, VkrtHit* hits)
{
    if (count > VKRT_STATIC_MAX_RAYS)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_rays_static: at most 65536 rays per call (VKRT_STATIC_MAX_RAYS); "
                                                      "split the batch or use vkrt_trace_rays");
    }
    if (count == 0)
    {
        return VKRT_SUCCESS;
    }
    if (int rc = checkTlasCurrent(ctx, tlas, "vkrt_trace_rays_static"))
    {
        return rc;
    }
    StaticScratch* scratch = nullptr;
    if (int rc = staticScratchAddress(ctx, &scratch))
    {
        return rc;
    }
    cudaStream_t s = ctx->stream;
    // the array keeps the last call's contents: the cursor and the overflow flag must start at zero
    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(scratch->header, 0, sizeof(scratch->header), s));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(scratch->rays, rays, (size_t)count * sizeof(VkrtRay), cudaMemcpyDefault, s));
    SceneDev scene;
    scene.instances = tlas->instances.as<InstanceRecord>();
    scene.topNodes = tlas->nodeCount ? tlas->nodes.as<Node8>() : nullptr;
    scene.leafInstance = tlas->leafInstance.as<uint32_t>();
    scene.instanceCount = tlas->instanceCount;
    scene.cullMask = 0xFFu;
    scene.errorFlags = scratch->header + 4;
    scene.topBoxScale = tlas->boxScale;
    scene.rows = nullptr;
    scene.textures = nullptr;
    scene.one = 0x3F800000u;
    scene.mips = nullptr;
    const uint32_t grid = persistentGrid(ctx, VKRT_TRACE_MIN_BLOCKS);
    const bool multi = scene.topNodes != nullptr;
    if (anyHit)
    {
        multi ? k_trace_rays_static<true, true><<<grid, kTraceThreads, 0, s>>>(scene, scratch, count)
              : k_trace_rays_static<true, false><<<grid, kTraceThreads, 0, s>>>(scene, scratch, count);
    }
    else
    {
        multi ? k_trace_rays_static<false, true><<<grid, kTraceThreads, 0, s>>>(scene, scratch, count)
              : k_trace_rays_static<false, false><<<grid, kTraceThreads, 0, s>>>(scene, scratch, count);
    }
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    ctx->launches += 1;
    uint32_t flags[4] = {0, 0, 0, 0};
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(hits, scratch->hits, (size_t)count * sizeof(VkrtHit), cudaMemcpyDefault, s));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(flags, scratch->header + 4, sizeof(flags), cudaMemcpyDeviceToHost, s));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    if (flags[0])
    {
        ctx->stats.traversalOverflows += 1;
        return ctx->fail(VKRT_ERROR_TRAVERSAL_OVERFLOW, "vkrt_trace_rays_static: a ray ran out of traversal stack entries; the hits are incomplete");
    }
    return VKRT_SUCCESS;
}

} // namespace vkrt

extern "C" int vkrt_trace_rays_static(VkrtContext* c, const VkrtTlas* tlas, const VkrtRay* rays, uint32_t count, int anyHit, VkrtHit* hits)
{
    vkrt::Context* ctx = static_cast<vkrt::Context*>(c);
    if (!ctx)
    {
        return VKRT_ERROR_INVALID_ARGUMENT;
    }
    if (!tlas || (count && (!rays || !hits)))
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_trace_rays_static: NULL argument");
    }
    const cudaError_t e = cudaSetDevice(ctx->device);
    if (e != cudaSuccess)
    {
        return ctx->fail(VKRT_ERROR_CUDA, std::string("vkrt_trace_rays_static: cudaSetDevice: ") + cudaGetErrorString(e));
    }
    return vkrt::traceRaysStatic(ctx, static_cast<const vkrt::Tlas*>(tlas), rays, count, anyHit, hits);
}
