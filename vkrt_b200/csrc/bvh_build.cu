// bvh_build.cu — GPU BVH builder kernels and their host driver.
// Replaces Raytracer::createBLAS / vkCmdBuildAccelerationStructuresKHR (src/Raytracer.cpp:1027-1157).
//
// HBM-bound integer/byte work: kernels are sized in multiples of the SM count where they are grid-stride,
// all per-triangle streams are coalesced 4/16-byte accesses, the sort stages tiles through shared memory.
#include "builder.cuh"
#include "context.hpp"
#include "reinsert.cuh"
#include "traverse.cuh"

#include <algorithm>
#include <cfloat>
#include <cstring>

namespace vkrt
{

namespace
{
constexpr int kThreads = 256;

__device__ __forceinline__ uint32_t orderedFromFloat(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float floatFromOrdered(uint32_t o)
{
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o;
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f;
    memcpy(&f, &u, 4);
    return f;
#endif
}

// sceneBounds: 12 ordered uints — [0..2] min of triangle boxes, [3..5] max, [6..8] centroid min, [9..11] centroid max
__global__ void k_init_build(uint32_t* sceneBounds, uint32_t* counters, float* sahAccum, int* root)
{
    const int i = threadIdx.x;
    if (i < 12)
    {
        const bool isMin = (i < 3) || (i >= 6 && i < 9);
        sceneBounds[i] = isMin ? 0xFFFFFFFFu : 0u;
    }
    if (i < 8)
    {
        counters[i] = i == 3 ? 255u : 0u; // [3]: running minimum of the nodes' scale exponents
    }
    if (i < 2)
    {
        sahAccum[i] = 0.0f;
    }
    if (i == 0)
    {
        *root = 0;
    }
}

__device__ __forceinline__ float warpMin(float v)
{
    for (int o = 16; o > 0; o >>= 1)
    {
        v = fminf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    }
    return v;
}
__device__ __forceinline__ float warpMax(float v)
{
    for (int o = 16; o > 0; o >>= 1)
    {
        v = fmaxf(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
    }
    return v;
}

// Per-triangle AABB + reduction of the scene and centroid bounds (warp shuffle -> one atomic per warp).
__global__ void __launch_bounds__(kThreads)
k_tri_bounds(GeometryTable geo, const uint8_t* __restrict__ vertices, uint32_t stride, const uint32_t* __restrict__ indices,
             uint32_t count, float4* __restrict__ triBox, float4* __restrict__ packedBuild, uint32_t* sceneBounds)
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        F3 v0, v1, v2;
        uint32_t g, p;
        fetchTriangle(geo, vertices, stride, indices, i, v0, v1, v2, g, p);
        const float4 l = make_float4(fminf(fminf(v0.x, v1.x), v2.x), fminf(fminf(v0.y, v1.y), v2.y),
                                     fminf(fminf(v0.z, v1.z), v2.z), 0.0f);
        const float4 h = make_float4(fmaxf(fmaxf(v0.x, v1.x), v2.x), fmaxf(fmaxf(v0.y, v1.y), v2.y),
                                     fmaxf(fmaxf(v0.z, v1.z), v2.z), 0.0f);
        triBox[2 * (size_t)i] = l;
        triBox[2 * (size_t)i + 1] = h;
        // the packed triangle in build order: k_pack_tris then only permutes 48-byte records instead of gathering the
        // three (64-byte stride) vertices a second time
        float4* dst = packedBuild + 3 * (size_t)i;
        dst[0] = make_float4(v0.x, v0.y, v0.z, v1.x);
        dst[1] = make_float4(v1.y, v1.z, v2.x, v2.y);
        dst[2] = make_float4(v2.z, __uint_as_float(g), __uint_as_float(p), __uint_as_float(i));
        const float c[3] = {0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)};
        lo[0] = fminf(lo[0], l.x), lo[1] = fminf(lo[1], l.y), lo[2] = fminf(lo[2], l.z);
        hi[0] = fmaxf(hi[0], h.x), hi[1] = fmaxf(hi[1], h.y), hi[2] = fmaxf(hi[2], h.z);
        for (int k = 0; k < 3; ++k)
        {
            clo[k] = fminf(clo[k], c[k]);
            chi[k] = fmaxf(chi[k], c[k]);
        }
    }
    for (int k = 0; k < 3; ++k)
    {
        const float a = warpMin(lo[k]), b = warpMax(hi[k]), c = warpMin(clo[k]), d = warpMax(chi[k]);
        if ((threadIdx.x & 31) == 0)
        {
            atomicMin(&sceneBounds[k], orderedFromFloat(a));
            atomicMax(&sceneBounds[3 + k], orderedFromFloat(b));
            atomicMin(&sceneBounds[6 + k], orderedFromFloat(c));
            atomicMax(&sceneBounds[9 + k], orderedFromFloat(d));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Onesweep LSD radix sort (Adinets & Merrill 2022), key-value (u32,u32), four 8-bit digits.
//   k_morton_hist   : Morton codes + the four global digit histograms in the same pass over the triangles
//   k_onesweep_pass : one launch per digit.  A block takes the next tile (4096 keys) by atomic ticket, ranks its
//                     keys (warp-level match), publishes the tile's digit counts and resolves its global offsets by
//                     decoupled look-back over the preceding tiles (no separate scan kernel), then scatters through
//                     shared memory so that each digit run leaves as consecutive addresses.
// Look-back words: bits 30..31 = state (0 not ready, 1 tile count, 2 inclusive prefix), bits 0..29 = value, hence
// fewer than 2^30 keys.  A block only ever waits for tiles with a smaller ticket, whose blocks are already running.
// ------------------------------------------------------------------------------------------------
constexpr int kSortItems = 16;
constexpr int kSortTile = kThreads * kSortItems; // 4096
constexpr int kSortWarps = kThreads / 32;
constexpr uint32_t kLookAggregate = 0x40000000u, kLookInclusive = 0x80000000u, kLookValue = 0x3FFFFFFFu;

__device__ __forceinline__ void histAccumulate(uint32_t (*hist)[256], uint32_t key)
{
    atomicAdd(&hist[0][key & 0xFFu], 1u);
    atomicAdd(&hist[1][(key >> 8) & 0xFFu], 1u);
    atomicAdd(&hist[2][(key >> 16) & 0xFFu], 1u);
    atomicAdd(&hist[3][key >> 24], 1u);
}
__device__ __forceinline__ void histFlush(uint32_t (*hist)[256], uint32_t* __restrict__ globalHist)
{
    __syncthreads();
    for (int i = threadIdx.x; i < 1024; i += blockDim.x)
    {
        const uint32_t v = hist[i >> 8][i & 255];
        if (v)
        {
            atomicAdd(&globalHist[i], v);
        }
    }
}

__global__ void __launch_bounds__(kThreads)
k_morton_hist(const float4* __restrict__ triBox, const uint32_t* __restrict__ sceneBounds,
              uint32_t count, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ globalHist)
{
    __shared__ uint32_t hist[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x)
    {
        hist[i >> 8][i & 255] = 0u;
    }
    __syncthreads();
    const F3 lo = f3(floatFromOrdered(sceneBounds[6]), floatFromOrdered(sceneBounds[7]), floatFromOrdered(sceneBounds[8]));
    const F3 hi = f3(floatFromOrdered(sceneBounds[9]), floatFromOrdered(sceneBounds[10]), floatFromOrdered(sceneBounds[11]));
    const F3 inv = f3(hi.x > lo.x ? 1.0f / (hi.x - lo.x) : 0.0f, hi.y > lo.y ? 1.0f / (hi.y - lo.y) : 0.0f,
                      hi.z > lo.z ? 1.0f / (hi.z - lo.z) : 0.0f);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const float4 l = triBox[2 * (size_t)i], h = triBox[2 * (size_t)i + 1];
        const F3 c = f3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z));
        const uint32_t key = morton30(c, lo, inv);
        keys[i] = key;
        vals[i] = i;
        histAccumulate(hist, key);
    }
    histFlush(hist, globalHist);
}

// histograms of arbitrary keys (debug sort entry)
__global__ void __launch_bounds__(kThreads) k_onesweep_hist(const uint32_t* __restrict__ keys, uint32_t count, uint32_t* __restrict__ globalHist)
{
    __shared__ uint32_t hist[4][256];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x)
    {
        hist[i >> 8][i & 255] = 0u;
    }
    __syncthreads();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        histAccumulate(hist, keys[i]);
    }
    histFlush(hist, globalHist);
}

// exclusive scan of 256 values, one per thread of a 256-thread block
__device__ __forceinline__ uint32_t blockExclusive256(uint32_t v, uint32_t* warpSums /* >= 8 */)
{
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1)
    {
        const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o)
        {
            incl += t;
        }
    }
    if (lane == 31)
    {
        warpSums[warp] = incl;
    }
    __syncthreads();
    uint32_t base = 0;
    for (uint32_t w = 0; w < warp; ++w)
    {
        base += warpSums[w];
    }
    __syncthreads();
    return base + incl - v;
}

__global__ void __launch_bounds__(kThreads)
k_onesweep_pass(const uint32_t* __restrict__ keysIn, const uint32_t* __restrict__ valsIn, uint32_t count, uint32_t shift,
                const uint32_t* __restrict__ globalHist /* 256 counts of this digit */, uint32_t* __restrict__ ticket,
                uint32_t* __restrict__ status /* numTiles x 256, zeroed */, uint32_t* __restrict__ keysOut, uint32_t* __restrict__ valsOut)
{
    __shared__ uint32_t warpHist[kSortWarps][256];
    __shared__ uint32_t stageKey[kSortTile];
    __shared__ uint32_t stageVal[kSortTile];
    __shared__ uint32_t localStart[256]; // first staging slot of each digit
    __shared__ uint32_t globalStart[256]; // output position of the digit's first key of this tile
    __shared__ uint32_t warpSums[kSortWarps];
    __shared__ uint32_t sTile;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
    {
        sTile = atomicAdd(ticket, 1u);
    }
    for (int w = 0; w < kSortWarps; ++w)
    {
        warpHist[w][threadIdx.x] = 0;
    }
    __syncthreads();
    const uint32_t tile = sTile;

    // warp-striped: warp w owns 512 consecutive keys; item j, lane l -> key w*512 + j*32 + l (stable order)
    const uint32_t base = tile * kSortTile + warp * (32 * kSortItems);
    uint32_t key[kSortItems], rank[kSortItems];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int j = 0; j < kSortItems; ++j)
    {
        const uint32_t idx = base + j * 32 + lane;
        const bool valid = idx < count;
        key[j] = valid ? keysIn[idx] : 0xFFFFFFFFu;
        const uint32_t digit = (key[j] >> shift) & 0xFFu;
        const uint32_t peers = __match_any_sync(0xFFFFFFFFu, valid ? digit : 0x100u);
        const uint32_t leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (valid && lane == leader)
        {
            pre = warpHist[warp][digit];
            warpHist[warp][digit] = pre + __popc(peers);
        }
        pre = __shfl_sync(0xFFFFFFFFu, pre, leader);
        rank[j] = pre + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();

    // thread d owns digit d: exclusive prefix over the warps of this tile, the tile's count, look-back
    const uint32_t d = threadIdx.x;
    uint32_t tileCount = 0;
    for (int w = 0; w < kSortWarps; ++w)
    {
        const uint32_t v = warpHist[w][d];
        warpHist[w][d] = tileCount;
        tileCount += v;
    }
    volatile uint32_t* st = status;
    st[(size_t)tile * 256 + d] = tileCount | (tile == 0 ? kLookInclusive : kLookAggregate);
    const uint32_t digitBase = blockExclusive256(globalHist[d], warpSums);
    localStart[d] = blockExclusive256(tileCount, warpSums);
    uint32_t exclusive = 0;
    if (tile > 0)
    {
        for (uint32_t t = tile - 1;; --t)
        {
            uint32_t w;
            do
            {
                w = st[(size_t)t * 256 + d];
            } while ((w & ~kLookValue) == 0u);
            exclusive += w & kLookValue;
            if (w & kLookInclusive)
            {
                break;
            }
        }
        st[(size_t)tile * 256 + d] = (exclusive + tileCount) | kLookInclusive;
    }
    globalStart[d] = digitBase + exclusive;
    __syncthreads();

    // stage in tile-sorted order
#pragma unroll
    for (int j = 0; j < kSortItems; ++j)
    {
        const uint32_t idx = base + j * 32 + lane;
        if (idx < count)
        {
            const uint32_t digit = (key[j] >> shift) & 0xFFu;
            const uint32_t slot = localStart[digit] + warpHist[warp][digit] + rank[j];
            stageKey[slot] = key[j];
            stageVal[slot] = valsIn[idx];
        }
    }
    __syncthreads();
    const uint32_t tileKeys = min((uint32_t)kSortTile, count - tile * kSortTile);
    for (uint32_t i = threadIdx.x; i < tileKeys; i += kThreads)
    {
        const uint32_t k = stageKey[i];
        const uint32_t digit = (k >> shift) & 0xFFu;
        const uint32_t dst = globalStart[digit] + (i - localStart[digit]);
        keysOut[dst] = k;
        valsOut[dst] = stageVal[i];
    }
}

__global__ void __launch_bounds__(kThreads) k_lbvh(Bvh2 bvh)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < bvh.count)
    {
        if (bvh.nodes32)
        {
            lbvhClimbCompact(bvh, i);
        }
        else
        {
            lbvhClimb(bvh, i);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// PLOC (Meister & Bittner 2018, "Parallel Locally-Ordered Clustering"): the default, higher-quality binary hierarchy.
// Clusters live in Morton order; every round each cluster finds its nearest neighbour (smallest merged surface
// area) within kPlocRadius positions, mutual pairs merge into a new binary node, and the survivors are compacted in
// order.  One kernel per round (PLOC++ style): a persistent grid takes tiles of 1024 clusters by ticket, searches in
// shared memory (tile + halo), and resolves the compaction offset by decoupled look-back.  Equal distances are
// ordered by (i xor j) so that duplicates pair up as buddies instead of chaining.
// ------------------------------------------------------------------------------------------------
constexpr int kPlocRadius = 8;
constexpr int kPlocTile = 1024;
constexpr int kPlocHalo = 2 * kPlocRadius;
constexpr int kPlocThreads = 256;
constexpr int kPlocSpan = kPlocTile + 2 * kPlocHalo;
constexpr int kPlocPerThread = kPlocTile / kPlocThreads;

struct PlocClusters
{
    int* id;      // binary-tree ref: >= 0 internal node, < 0 ~sortedLeaf
    float4* lo;   // w = leaf count (raw bits)
    float4* hi;
};

// state: [0],[1] cluster counts of buffers A/B, [2] ticket, [3] finished blocks, [4] binary nodes allocated
__global__ void __launch_bounds__(kThreads)
k_ploc_init(Bvh2 bvh, PlocClusters c, uint32_t* state, uint32_t* status, uint32_t statusCount)
{
    const uint32_t n = bvh.count;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const uint32_t tri = bvh.vals[i];
        float4 lo = bvh.triBox[2 * (size_t)tri], hi = bvh.triBox[2 * (size_t)tri + 1];
        lo.w = __uint_as_float(1u);
        hi.w = 0.0f;
        c.id[i] = ~(int)i;
        c.lo[i] = lo;
        c.hi[i] = hi;
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < statusCount; i += gridDim.x * blockDim.x)
    {
        status[i] = 0u;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        state[0] = n, state[1] = 0u, state[2] = 0u, state[3] = 0u, state[4] = 0u;
    }
}

__global__ void __launch_bounds__(kPlocThreads)
k_ploc_round(PlocClusters in, PlocClusters out, Bvh2 bvh, uint32_t* state, uint32_t* status, int parity)
{
    __shared__ float4 sLo[kPlocSpan];
    __shared__ float4 sHi[kPlocSpan];
    __shared__ int sId[kPlocSpan];
    __shared__ short sNN[kPlocSpan];
    __shared__ uint32_t sWarp[kPlocThreads / 32];
    __shared__ uint32_t sTile, sExclusive, sNodeBase, sLast;

    const uint32_t n = state[parity];
    if (n <= 1u)
    {
        // nothing to merge: hand the (single) cluster over so that the result always sits in the buffer of the last round
        if (blockIdx.x == 0 && threadIdx.x == 0)
        {
            if (n == 1u)
            {
                out.id[0] = in.id[0], out.lo[0] = in.lo[0], out.hi[0] = in.hi[0];
            }
            state[parity ^ 1] = n;
        }
        return;
    }
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (n <= (uint32_t)kPlocTile)
    {
        // Tail: everything fits one tile, so block 0 runs all remaining rounds out of shared memory in this launch
        // (about 40 of the 74 rounds of a 262 k-triangle build are this small; as launches they cost ~8 us each).
        if (blockIdx.x != 0)
        {
            return;
        }
        for (uint32_t k = threadIdx.x; k < n; k += kPlocThreads)
        {
            sLo[k] = in.lo[k], sHi[k] = in.hi[k], sId[k] = in.id[k];
        }
        uint32_t m = n, nodeNext = state[4];
        __syncthreads();
        while (m > 1u)
        {
            for (uint32_t k = threadIdx.x; k < m; k += kPlocThreads)
            {
                const float4 lo = sLo[k], hi = sHi[k];
                float bestD = INFINITY;
                uint32_t bestX = 0xFFFFFFFFu;
                int bestJ = -1;
                const int j0 = max((int)k - kPlocRadius, 0), j1 = min((int)k + kPlocRadius, (int)m - 1);
                for (int j = j0; j <= j1; ++j)
                {
                    if (j == (int)k)
                    {
                        continue;
                    }
                    const float4 l2 = sLo[j], h2 = sHi[j];
                    const float dx = fmaxf(hi.x, h2.x) - fminf(lo.x, l2.x), dy = fmaxf(hi.y, h2.y) - fminf(lo.y, l2.y),
                                dz = fmaxf(hi.z, h2.z) - fminf(lo.z, l2.z);
                    const float d = dx * dy + dy * dz + dz * dx;
                    const uint32_t x = k ^ (uint32_t)j;
                    if (d < bestD || (d == bestD && x < bestX) || bestJ < 0)
                    {
                        bestD = d, bestX = x, bestJ = j;
                    }
                }
                sNN[k] = (short)bestJ;
            }
            __syncthreads();
            // kPlocPerThread consecutive clusters per thread, results held in registers until everyone has read
            uint32_t keepMask = 0, mergeMask = 0;
            float4 rLo[kPlocPerThread], rHi[kPlocPerThread];
            int rId[kPlocPerThread], rP[kPlocPerThread];
#pragma unroll
            for (int e = 0; e < kPlocPerThread; ++e)
            {
                const uint32_t k = threadIdx.x * kPlocPerThread + e;
                rP[e] = -1;
                if (k < m)
                {
                    const int p = sNN[k];
                    const bool mutual = sNN[p] == (short)k;
                    rLo[e] = sLo[k], rHi[e] = sHi[k], rId[e] = sId[k];
                    if (!mutual || (int)k < p)
                    {
                        keepMask |= 1u << e;
                    }
                    if (mutual && (int)k < p)
                    {
                        mergeMask |= 1u << e;
                        rP[e] = p;
                        const float4 l2 = sLo[p], h2 = sHi[p];
                        const uint32_t leaves = __float_as_uint(rLo[e].w) + __float_as_uint(l2.w);
                        rLo[e] = make_float4(fminf(rLo[e].x, l2.x), fminf(rLo[e].y, l2.y), fminf(rLo[e].z, l2.z), __uint_as_float(leaves));
                        rHi[e] = make_float4(fmaxf(rHi[e].x, h2.x), fmaxf(rHi[e].y, h2.y), fmaxf(rHi[e].z, h2.z), 0.0f);
                    }
                }
            }
            const uint32_t mine = (uint32_t)__popc(keepMask) | ((uint32_t)__popc(mergeMask) << 16);
            uint32_t incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                if (lane >= (uint32_t)o)
                {
                    incl += t;
                }
            }
            if (lane == 31)
            {
                sWarp[warp] = incl;
            }
            __syncthreads(); // also: every thread has finished reading the shared cluster arrays
            uint32_t warpBase = 0, total = 0;
            for (uint32_t w2 = 0; w2 < kPlocThreads / 32; ++w2)
            {
                const uint32_t v = sWarp[w2];
                if (w2 < warp)
                {
                    warpBase += v;
                }
                total += v;
            }
            const uint32_t excl = warpBase + incl - mine;
            uint32_t dst = excl & 0xFFFFu, node = nodeNext + (excl >> 16);
#pragma unroll
            for (int e = 0; e < kPlocPerThread; ++e)
            {
                if (!((keepMask >> e) & 1u))
                {
                    continue;
                }
                int id = rId[e];
                if ((mergeMask >> e) & 1u)
                {
                    bvh.nodes[node].lo = rLo[e];
                    bvh.nodes[node].hi = rHi[e];
                    bvh.nodes[node].children = make_int2(id, sId[rP[e]]);
                    id = (int)node;
                    ++node;
                }
                rId[e] = id;
                rP[e] = (int)dst; // destination slot
                ++dst;
            }
            __syncthreads(); // partner ids (sId) have been read
#pragma unroll
            for (int e = 0; e < kPlocPerThread; ++e)
            {
                if ((keepMask >> e) & 1u)
                {
                    sLo[rP[e]] = rLo[e], sHi[rP[e]] = rHi[e], sId[rP[e]] = rId[e];
                }
            }
            m = total & 0xFFFFu;
            nodeNext += total >> 16;
            __syncthreads();
        }
        if (threadIdx.x == 0)
        {
            out.id[0] = sId[0], out.lo[0] = sLo[0], out.hi[0] = sHi[0];
            state[4] = nodeNext;
            state[parity ^ 1] = 1u;
        }
        return;
    }
    const uint32_t numTiles = (n + kPlocTile - 1) / kPlocTile;
    volatile uint32_t* st = status;

    for (;;)
    {
        if (threadIdx.x == 0)
        {
            sTile = atomicAdd(&state[2], 1u);
        }
        __syncthreads();
        const uint32_t tile = sTile;
        if (tile >= numTiles)
        {
            break;
        }
        const int64_t start = (int64_t)tile * kPlocTile - kPlocHalo; // global index of span slot 0
        for (int k = threadIdx.x; k < kPlocSpan; k += kPlocThreads)
        {
            const int64_t g = start + k;
            if (g >= 0 && g < (int64_t)n)
            {
                sLo[k] = in.lo[g];
                sHi[k] = in.hi[g];
                sId[k] = in.id[g];
            }
        }
        __syncthreads();
        // nearest neighbour of every cluster whose choice can matter to the tile: slots [radius, span - radius)
        for (int k = kPlocRadius + threadIdx.x; k < kPlocSpan - kPlocRadius; k += kPlocThreads)
        {
            const int64_t g = start + k;
            if (g < 0 || g >= (int64_t)n)
            {
                continue;
            }
            const float4 lo = sLo[k], hi = sHi[k];
            float bestD = INFINITY;
            uint32_t bestX = 0xFFFFFFFFu;
            int bestJ = -1;
#pragma unroll 4
            for (int j = k - kPlocRadius; j <= k + kPlocRadius; ++j)
            {
                const int64_t gj = start + j;
                if (j == k || gj < 0 || gj >= (int64_t)n)
                {
                    continue;
                }
                const float4 l2 = sLo[j], h2 = sHi[j];
                const float dx = fmaxf(hi.x, h2.x) - fminf(lo.x, l2.x), dy = fmaxf(hi.y, h2.y) - fminf(lo.y, l2.y),
                            dz = fmaxf(hi.z, h2.z) - fminf(lo.z, l2.z);
                const float d = dx * dy + dy * dz + dz * dx;
                const uint32_t x = (uint32_t)g ^ (uint32_t)gj;
                if (d < bestD || (d == bestD && x < bestX) || bestJ < 0)
                {
                    bestD = d, bestX = x, bestJ = j;
                }
            }
            sNN[k] = (short)bestJ;
        }
        __syncthreads();

        // merge decisions for the tile's own clusters, kPlocPerThread consecutive ones per thread
        uint32_t keepMask = 0, mergeMask = 0;
#pragma unroll
        for (int e = 0; e < kPlocPerThread; ++e)
        {
            const int k = kPlocHalo + threadIdx.x * kPlocPerThread + e;
            const int64_t g = start + k;
            if (g < (int64_t)n)
            {
                const int p = sNN[k];
                const bool mutual = p >= 0 && sNN[p] == k;
                if (!mutual || k < p)
                {
                    keepMask |= 1u << e;
                }
                if (mutual && k < p)
                {
                    mergeMask |= 1u << e;
                }
            }
        }
        // block scan of (kept, merged) packed as 16 + 16 bits
        const uint32_t mine = (uint32_t)__popc(keepMask) | ((uint32_t)__popc(mergeMask) << 16);
        uint32_t incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1)
        {
            const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= (uint32_t)o)
            {
                incl += t;
            }
        }
        if (lane == 31)
        {
            sWarp[warp] = incl;
        }
        __syncthreads();
        uint32_t warpBase = 0, total = 0;
        for (uint32_t w2 = 0; w2 < kPlocThreads / 32; ++w2)
        {
            const uint32_t v = sWarp[w2];
            if (w2 < warp)
            {
                warpBase += v;
            }
            total += v;
        }
        const uint32_t excl = warpBase + incl - mine;
        if (threadIdx.x == 0)
        {
            const uint32_t tileKeep = total & 0xFFFFu, tileMerge = total >> 16;
            st[tile] = tileKeep | (tile == 0 ? kLookInclusive : kLookAggregate);
            uint32_t exclusive = 0;
            if (tile > 0)
            {
                for (uint32_t t = tile - 1;; --t)
                {
                    uint32_t w2;
                    do
                    {
                        w2 = st[t];
                    } while ((w2 & ~kLookValue) == 0u);
                    exclusive += w2 & kLookValue;
                    if (w2 & kLookInclusive)
                    {
                        break;
                    }
                }
                st[tile] = (exclusive + tileKeep) | kLookInclusive;
            }
            sExclusive = exclusive;
            sNodeBase = tileMerge ? atomicAdd(&state[4], tileMerge) : 0u;
            if (tile == numTiles - 1)
            {
                state[parity ^ 1] = exclusive + tileKeep;
            }
        }
        __syncthreads();
        uint32_t dst = sExclusive + (excl & 0xFFFFu);
        uint32_t node = sNodeBase + (excl >> 16);
#pragma unroll
        for (int e = 0; e < kPlocPerThread; ++e)
        {
            if (!((keepMask >> e) & 1u))
            {
                continue;
            }
            const int k = kPlocHalo + threadIdx.x * kPlocPerThread + e;
            float4 lo = sLo[k], hi = sHi[k];
            int id = sId[k];
            if ((mergeMask >> e) & 1u)
            {
                const int p = sNN[k];
                const float4 l2 = sLo[p], h2 = sHi[p];
                const uint32_t leaves = __float_as_uint(lo.w) + __float_as_uint(l2.w);
                lo = make_float4(fminf(lo.x, l2.x), fminf(lo.y, l2.y), fminf(lo.z, l2.z), __uint_as_float(leaves));
                hi = make_float4(fmaxf(hi.x, h2.x), fmaxf(hi.y, h2.y), fmaxf(hi.z, h2.z), 0.0f);
                bvh.nodes[node].lo = lo;
                bvh.nodes[node].hi = hi;
                bvh.nodes[node].children = make_int2(id, sId[p]);
                id = (int)node;
                ++node;
            }
            out.id[dst] = id;
            out.lo[dst] = lo;
            out.hi[dst] = hi;
            ++dst;
        }
        __syncthreads(); // shared arrays are reused by the next tile
    }

    // the last block to finish re-arms ticket and look-back words for the next round
    __threadfence();
    if (threadIdx.x == 0)
    {
        sLast = atomicAdd(&state[3], 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (sLast)
    {
        for (uint32_t i = threadIdx.x; i < numTiles; i += kPlocThreads)
        {
            status[i] = 0u;
        }
        if (threadIdx.x == 0)
        {
            state[2] = 0u, state[3] = 0u;
        }
    }
}

// root of the binary hierarchy = the single cluster left in buffer `parity`
__global__ void k_ploc_finish(PlocClusters c, Bvh2 bvh, const uint32_t* state, int parity, uint32_t* remaining)
{
    *remaining = state[parity];
    if (state[parity] == 1u && bvh.count > 1)
    {
        *bvh.root = c.id[0];
    }
}

// Level-synchronous collapse without host round trips.  `range` = {first, last, ticket, levels}: wide nodes
// [first,last) were allocated by the previous level.  Narrow levels (<= kNarrowMax nodes) are swept by ONE block that
// loops over levels with __syncthreads(); wide levels take one multi-block launch each, whose last block publishes
// the next range.  Each kernel is a no-op when the current level belongs to the other kind, so the host just
// alternates them (see buildBlasOnDevice).
constexpr int kGraphWidePairs = 12;     // (wide, narrow) launch pairs in a graph-replayed build (no read-back in between)
constexpr uint32_t kNarrowLevel = 1024; // threads of the one-block sweep
constexpr uint32_t kNarrowMax = 256;    // largest level the one-block sweep takes: above, a launch over all SMs is faster than one SM
                                        // (Sponza 1.78 -> 1.54 ms default, 0.70 -> 0.61 ms fast build; profiles/r02_y_narrow_max.txt)

__device__ __forceinline__ void flushSah(const CollapseArgs& a, const float* sah)
{
    // one atomic per warp instead of two per wide node (same-address float atomics serialise in L2)
    float s0 = sah[0], s1 = sah[1];
    for (int o = 16; o > 0; o >>= 1)
    {
        s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
        s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
    }
    if ((threadIdx.x & 31) == 0 && a.sahAccum)
    {
        atomicAdd(&a.sahAccum[0], s0);
        atomicAdd(&a.sahAccum[1], s1);
    }
}

// Cooperative form of collapseNode (builder.cuh): 8 lanes build one wide node, lane k holding child k, so the
// dependent chain of child fetches, the greedy slot assignment and the quantisation run across lanes instead of as
// ~500-step loops over local arrays in one thread.  Same decisions as the scalar form (lowest index wins ties).
// All 32 lanes of a warp must call it; `valid` marks groups that have a node.  `stage` = 80 bytes per group.
__device__ __forceinline__ void collapseNodeCoop(const CollapseArgs& a, uint32_t wideIndex, bool valid, uint8_t* stage, float* sahLocal)
{
    const Bvh2& b = a.bvh2;
    const uint32_t lane = threadIdx.x & 31u, l8 = lane & 7u, base = lane & ~7u;
    const unsigned full = 0xFFFFFFFFu;

    int rootRef = ~0;
    CollapseChild self;
    self.ref = 0, self.lo = self.hi = f3(0.f, 0.f, 0.f), self.count = 0;
    if (valid)
    {
        rootRef = a.taskRef[wideIndex];
        self = collapseFetch(b, rootRef);
    }
    // children: lane k holds child k while k < n
    CollapseChild c;
    c.ref = 0, c.lo = c.hi = f3(0.f, 0.f, 0.f), c.count = 0;
    uint32_t n = 0;
    if (valid)
    {
        if (rootRef < 0 || self.count <= b.maxLeaf)
        {
            n = 1;
            if (l8 == 0)
            {
                c = self;
            }
        }
        else if (!b.dpChoice)
        {
            n = 2;
            const int2 ch = binChildren(b, rootRef);
            if (l8 < 2)
            {
                c = collapseFetch(b, l8 == 0 ? ch.x : ch.y);
            }
        }
    }
    if (b.dpChoice)
    {
        // cost-optimal collapse: lane 0 of the group walks the decisions, then every lane fetches its entry
        int refs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t m = 0;
        if (valid && n == 0 && l8 == 0)
        {
            m = (uint32_t)dpExpand(b, rootRef, refs);
        }
        m = __shfl_sync(full, m, base);
        int myRef = 0;
#pragma unroll
        for (int j = 0; j < 8; ++j)
        {
            const int r = __shfl_sync(full, refs[j], base);
            if (l8 == (uint32_t)j)
            {
                myRef = r;
            }
        }
        if (m > 0)
        {
            n = m;
            if (l8 < n)
            {
                c = collapseFetch(b, myRef);
            }
        }
    }
    // open children: first those that must become inner nodes otherwise (> 3 triangles), then multi-triangle leaves
    // (with the cost-optimal collapse only the latter: which subtrees become nodes is already decided)
    const uint32_t innerCat = b.dpChoice ? 0u : 2u;
    bool opening = valid && n >= 2;
#pragma unroll 1
    for (int iter = 0; iter < 6; ++iter)
    {
        const uint32_t cat = (l8 < n && opening) ? (c.count > b.maxLeaf ? innerCat : (c.count > 1u ? 1u : 0u)) : 0u;
        float area = cat ? halfArea(c.lo, c.hi) : -1.0f;
        uint32_t bestCat = cat, bestIdx = l8;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1)
        {
            const uint32_t oc = __shfl_xor_sync(full, bestCat, o);
            const float oa = __shfl_xor_sync(full, area, o);
            const uint32_t oi = __shfl_xor_sync(full, bestIdx, o);
            const bool take = oc > bestCat || (oc == bestCat && (oa > area || (oa == area && oi < bestIdx)));
            if (take)
            {
                bestCat = oc, area = oa, bestIdx = oi;
            }
        }
        if (bestCat == 0u)
        {
            opening = false;
        }
        const int pickRef = __shfl_sync(full, c.ref, base + bestIdx);
        if (opening && n < 8)
        {
            const int2 ch = binChildren(b, pickRef);
            if (l8 == bestIdx)
            {
                c = collapseFetch(b, ch.x);
            }
            else if (l8 == n)
            {
                c = collapseFetch(b, ch.y);
            }
            ++n;
            if (n == 8)
            {
                opening = false;
            }
        }
        if (!__any_sync(full, opening))
        {
            break;
        }
    }

    // octant-ordered slot assignment: greedy maximum of +-cx +-cy +-cz over (child, free slot)
    const bool has = valid && l8 < n;
    const float ox = 0.5f * (c.lo.x + c.hi.x) - 0.5f * (self.lo.x + self.hi.x);
    const float oy = 0.5f * (c.lo.y + c.hi.y) - 0.5f * (self.lo.y + self.hi.y);
    const float oz = 0.5f * (c.lo.z + c.hi.z) - 0.5f * (self.lo.z + self.hi.z);
    uint32_t freeMask = 0xFFu;
    int mySlot = -1;
#pragma unroll 1
    for (int iter = 0; iter < 8; ++iter)
    {
        float cost = -INFINITY;
        int slot = -1;
        if (has && mySlot < 0)
        {
#pragma unroll
            for (int sidx = 0; sidx < 8; ++sidx)
            {
                if ((freeMask >> sidx) & 1u)
                {
                    const float v = ((sidx & 1) ? ox : -ox) + ((sidx & 2) ? oy : -oy) + ((sidx & 4) ? oz : -oz);
                    if (v > cost || slot < 0)
                    {
                        cost = v, slot = sidx;
                    }
                }
            }
        }
        float bestCost = cost;
        int bestSlot = slot;
        uint32_t bestK = slot >= 0 ? l8 : 8u;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1)
        {
            const float oc = __shfl_xor_sync(full, bestCost, o);
            const int os = __shfl_xor_sync(full, bestSlot, o);
            const uint32_t ok = __shfl_xor_sync(full, bestK, o);
            const bool take = ok < 8u && (bestK >= 8u || oc > bestCost || (oc == bestCost && ok < bestK));
            if (take)
            {
                bestCost = oc, bestSlot = os, bestK = ok;
            }
        }
        if (bestK < 8u)
        {
            if (l8 == bestK)
            {
                mySlot = bestSlot;
            }
            freeMask &= ~(1u << bestSlot);
        }
        if (!__any_sync(full, bestK < 8u))
        {
            break;
        }
    }

    // bring the children into slot order: lane s takes the child whose slot is s
    int src = -1;
#pragma unroll
    for (int j = 0; j < 8; ++j)
    {
        const int sj = __shfl_sync(full, mySlot, base + j);
        if (sj == (int)l8)
        {
            src = j;
        }
    }
    const uint32_t from = base + (src < 0 ? 0 : src);
    CollapseChild d;
    d.ref = __shfl_sync(full, c.ref, from);
    d.lo.x = __shfl_sync(full, c.lo.x, from), d.lo.y = __shfl_sync(full, c.lo.y, from), d.lo.z = __shfl_sync(full, c.lo.z, from);
    d.hi.x = __shfl_sync(full, c.hi.x, from), d.hi.y = __shfl_sync(full, c.hi.y, from), d.hi.z = __shfl_sync(full, c.hi.z, from);
    d.count = __shfl_sync(full, c.count, from);
    const bool present = src >= 0;
    const bool inner = present && d.count > b.maxLeaf;
    const bool leaf = present && !inner;

    const uint32_t innerMask = (__ballot_sync(full, inner) >> base) & 0xFFu;
    const uint32_t innerCount = __popc(innerMask), innerRank = __popc(innerMask & ((1u << l8) - 1u));
    uint32_t leafIncl = leaf ? d.count : 0u;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1)
    {
        const uint32_t t = __shfl_up_sync(full, leafIncl, o, 8);
        if (l8 >= (uint32_t)o)
        {
            leafIncl += t;
        }
    }
    const uint32_t primOffset = leafIncl - (leaf ? d.count : 0u);
    const uint32_t primCount = __shfl_sync(full, leafIncl, base + 7);

    // allocation: one pair of atomics per warp (4 nodes)
    uint32_t needNodes = (l8 == 0 && valid) ? innerCount : 0u, needPrims = (l8 == 0 && valid) ? primCount : 0u;
    uint32_t nodesBefore = 0, primsBefore = 0, nodesTotal = 0, primsTotal = 0;
#pragma unroll
    for (int g = 0; g < 4; ++g)
    {
        const uint32_t gn = __shfl_sync(full, needNodes, g * 8), gp = __shfl_sync(full, needPrims, g * 8);
        if ((uint32_t)g * 8u < base)
        {
            nodesBefore += gn, primsBefore += gp;
        }
        nodesTotal += gn, primsTotal += gp;
    }
    uint32_t warpNodeBase = 0, warpPrimBase = 0;
    if (lane == 0)
    {
        warpNodeBase = nodesTotal ? atomicAdd(a.nodeCounter, nodesTotal) : 0u;
        warpPrimBase = primsTotal ? atomicAdd(a.primCounter, primsTotal) : 0u;
    }
    warpNodeBase = __shfl_sync(full, warpNodeBase, 0), warpPrimBase = __shfl_sync(full, warpPrimBase, 0);
    const uint32_t childBase = innerCount ? warpNodeBase + nodesBefore : 0u;
    const uint32_t primBase = primCount ? warpPrimBase + primsBefore : 0u;
    const bool overflow = valid && childBase + innerCount > a.nodeCapacity;
    if (overflow && l8 == 0)
    {
        *a.overflow = 1u;
    }

    // Node8::leafBits: slot l8 contributes its unary triangle count at bit 3*l8; OR over the 8 lanes of the group
    uint32_t leafBits = leaf ? ((1u << d.count) - 1u) << (3u * l8) : 0u;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1)
    {
        leafBits |= __shfl_xor_sync(full, leafBits, o, 8);
    }

    if (valid && !overflow)
    {
        uint32_t ex, ey, ez;
        const uint32_t em = quantExponents(self.lo, self.hi, ex, ey, ez);
        const float sx = __uint_as_float(ex << 23), sy = __uint_as_float(ey << 23), sz = __uint_as_float(ez << 23);
        uint8_t q[6] = {0, 0, 0, 0, 0, 0};
        if (present)
        {
            q[0] = (uint8_t)quantLo(d.lo.x, self.lo.x, sx), q[1] = (uint8_t)quantLo(d.lo.y, self.lo.y, sy);
            q[2] = (uint8_t)quantLo(d.lo.z, self.lo.z, sz), q[3] = (uint8_t)quantHi(d.hi.x, self.lo.x, sx);
            q[4] = (uint8_t)quantHi(d.hi.y, self.lo.y, sy), q[5] = (uint8_t)quantHi(d.hi.z, self.lo.z, sz);
        }
        if (inner)
        {
            a.taskRef[childBase + innerRank] = d.ref;
        }
        else if (leaf)
        {
            uint32_t leaves[4];
            collectLeaves(b, d.ref, leaves);
            for (uint32_t j = 0; j < d.count; ++j)
            {
                a.primSource[primBase + primOffset + j] = b.vals[leaves[j]];
            }
            sahLocal[1] += halfArea(d.lo, d.hi) * (float)d.count;
        }
#pragma unroll
        for (int f = 0; f < 6; ++f)
        {
            stage[32 + 8 * f + l8] = q[f];
        }
        if (l8 == 0)
        {
            float* hf = reinterpret_cast<float*>(stage);
            uint32_t* hu = reinterpret_cast<uint32_t*>(stage);
            hf[0] = self.lo.x, hf[1] = self.lo.y, hf[2] = self.lo.z;
            hu[3] = (uint32_t)packScaleExponent(ex) | ((uint32_t)packScaleExponent(ey) << 16);
            hu[4] = childBase, hu[5] = primBase, hu[6] = leafBits;
            (void)em;
            hu[7] = (uint32_t)packScaleExponent(ez) | (innerMask << 16);
            sahLocal[0] += halfArea(self.lo, self.hi);
        }
    }
    __syncwarp();
    if (valid && !overflow && l8 < 5)
    {
        reinterpret_cast<uint4*>(a.nodes + wideIndex)[l8] = reinterpret_cast<const uint4*>(stage)[l8];
    }
    __syncwarp();
}

__global__ void __launch_bounds__(kThreads) k_dp_parents(Bvh2 bvh)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 < bvh.count)
    {
        dpLinkParents(bvh, i);
    }
}

__global__ void __launch_bounds__(kThreads) k_dp_climb(Bvh2 bvh)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < bvh.count)
    {
        dpClimb(bvh, i);
    }
}

// ---- reinsertion passes (reinsert.cuh), VKRT_BUILD_PREFER_FAST_TRACE ----
// state[0] = surface area of the root box (float bits), state[1] = moves applied so far
__global__ void k_reinsert_begin(Bvh2 bvh, uint32_t* state)
{
    const int root = *bvh.root;
    bvh.dpParent[bvh.count + (uint32_t)root] = -1;
    const float4 lo = bvh.nodes[root].lo, hi = bvh.nodes[root].hi;
    state[0] = floatAsUint(halfArea(f3(lo.x, lo.y, lo.z), f3(hi.x, hi.y, hi.z)));
    state[1] = 0u;
}

// Only nodes whose surface area lies in [loMul, hiMul) x the root's area are looked at in this pass: large nodes first.
__global__ void __launch_bounds__(128) k_reinsert_search(ReinsertArgs a, const uint32_t* state, float hiMul, float loMul)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot < 2u * a.bvh.count - 1u)
    {
        const float rootArea = uintAsFloat(state[0]);
        a.areaHi = hiMul * rootArea, a.areaLo = loMul * rootArea;
        reinsertSearch(a, slot);
    }
}

__global__ void __launch_bounds__(kThreads) k_reinsert_lock(ReinsertArgs a)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot < 2u * a.bvh.count - 1u)
    {
        reinsertLock(a, slot);
    }
}

__global__ void __launch_bounds__(kThreads) k_reinsert_verify(ReinsertArgs a)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot < 2u * a.bvh.count - 1u)
    {
        reinsertVerify(a, slot);
    }
}

__global__ void __launch_bounds__(kThreads) k_reinsert_apply(ReinsertArgs a)
{
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot < 2u * a.bvh.count - 1u)
    {
        reinsertApply(a, slot);
    }
}

__global__ void __launch_bounds__(kThreads) k_reinsert_refit(Bvh2 bvh)
{
    const uint32_t leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf < bvh.count)
    {
        reinsertRefit(bvh, leaf);
    }
}

__global__ void k_collapse_seed(CollapseArgs a, uint32_t* range)
{
    a.taskRef[0] = a.bvh2.count > 1 ? *a.bvh2.root : ~0;
    *a.nodeCounter = 1u;
    *a.primCounter = 0u;
    range[0] = 0u, range[1] = 1u, range[2] = 0u, range[3] = 0u;
}

__global__ void __launch_bounds__(kNarrowLevel) k_collapse_narrow(CollapseArgs a, uint32_t* range, uint32_t narrowMax)
{
    __shared__ uint32_t sFirst, sLast;
    __shared__ __align__(16) uint8_t sStage[kNarrowLevel / 8][80];
    if (threadIdx.x == 0)
    {
        sFirst = range[0], sLast = range[1];
    }
    __syncthreads();
    uint32_t levels = 0;
    float sah[2] = {0.0f, 0.0f};
    const uint32_t group = threadIdx.x >> 3;
    for (;;)
    {
        const uint32_t first = sFirst, last = sLast;
        if (first >= last || last - first > narrowMax)
        {
            break;
        }
        for (uint32_t chunk = first; chunk < last; chunk += kNarrowLevel / 8)
        {
            const uint32_t i = chunk + group;
            collapseNodeCoop(a, i, i < last, sStage[group], sah);
        }
        __threadfence_block();
        __syncthreads();
        if (threadIdx.x == 0)
        {
            sFirst = last;
            sLast = *(volatile uint32_t*)a.nodeCounter;
        }
        ++levels;
        __syncthreads();
    }
    flushSah(a, sah);
    if (threadIdx.x == 0)
    {
        range[0] = sFirst, range[1] = sLast, range[3] += levels;
    }
}

__global__ void __launch_bounds__(128) k_collapse_wide(CollapseArgs a, uint32_t* range, uint32_t narrowMax)
{
    const uint32_t first = range[0], last = range[1];
    if (last - first <= narrowMax) // also covers the finished state first == last
    {
        return;
    }
    __shared__ __align__(16) uint8_t sStage[128 / 8][80];
    float sah[2] = {0.0f, 0.0f};
    const uint32_t groupsPerBlock = blockDim.x >> 3, group = threadIdx.x >> 3;
    if (last - first <= gridDim.x * blockDim.x)
    {
        // latency-bound level: 8 lanes per node
        for (uint32_t chunk = first + blockIdx.x * groupsPerBlock; chunk < last; chunk += gridDim.x * groupsPerBlock)
        {
            const uint32_t i = chunk + group;
            collapseNodeCoop(a, i, i < last, sStage[group], sah);
        }
    }
    else
    {
        // throughput-bound level (more nodes than threads): one thread per node costs fewer instructions in total
        for (uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x; i < last; i += gridDim.x * blockDim.x)
        {
            collapseNode(a, i, sah);
        }
    }
    flushSah(a, sah);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const uint32_t ticket = atomicAdd(&range[2], 1u);
        if (ticket == gridDim.x - 1) // every block has read the range and finished its nodes
        {
            range[0] = last;
            range[1] = atomicAdd(a.nodeCounter, 0u);
            range[2] = 0u;
            range[3] += 1u;
        }
    }
}

__global__ void k_empty_root(Node8* nodes)
{
    Node8 n;
    memset(&n, 0, sizeof(n));
    n.sx = n.sy = n.sz = packScaleExponent(1);
    nodes[0] = n;
}

__global__ void __launch_bounds__(kThreads)
k_pack_tris(const float4* __restrict__ packedBuild, const uint32_t* __restrict__ primSource, uint32_t count, PackedTri* __restrict__ out)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const float4* src = packedBuild + 3 * (size_t)primSource[i];
        const float4 a = src[0], b = src[1], c = src[2];
        float4* dst = reinterpret_cast<float4*>(out + i);
        dst[0] = a, dst[1] = b, dst[2] = c;
    }
}

inline uint32_t gridFor(uint32_t count, int threads, int smCount)
{
    const uint32_t blocks = (count + threads - 1) / threads;
    const uint32_t cap = (uint32_t)smCount * 16u; // grid-stride kernels: a whole number of waves
    return std::max(1u, std::min(blocks, cap));
}
} // namespace

// Sorts (keys, vals) by key; result ends in keysA/valsA (4 passes ping-pong A->B->A->B->A).  `sortState` holds
// {globalHist[4][256], ticket[4]} (the histograms already filled by k_morton_hist / k_onesweep_hist, tickets zero)
// followed by the 4 * numTiles * 256 zeroed look-back words.
static int radixSort(Context* ctx, uint32_t* keysA, uint32_t* valsA, uint32_t* keysB, uint32_t* valsB, uint32_t* sortState,
                     uint32_t count)
{
    const uint32_t numTiles = (count + kSortTile - 1) / kSortTile;
    uint32_t *kin = keysA, *vin = valsA, *kout = keysB, *vout = valsB;
    uint32_t* globalHist = sortState;
    uint32_t* ticket = sortState + 1024;
    uint32_t* status = sortState + 1040;
    for (uint32_t pass = 0; pass < 4; ++pass)
    {
        k_onesweep_pass<<<numTiles, kThreads, 0, ctx->stream>>>(kin, vin, count, pass * 8, globalHist + 256 * pass, ticket + pass,
                                                                   status + (size_t)pass * numTiles * 256, kout, vout);
        ctx->launches += 1;
        std::swap(kin, kout);
        std::swap(vin, vout);
    }
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    return VKRT_SUCCESS;
}
static size_t sortStateBytes(uint32_t count)
{
    const size_t numTiles = ((size_t)count + kSortTile - 1) / kSortTile;
    return (1040 + 4 * numTiles * 256) * sizeof(uint32_t);
}

// Device-resident (key, value) sort for other parts of the library (the rasteriser orders its large triangles front to
// back with it).  Everything is stream-ordered; the sorted pairs end up in keysA / valsA.
size_t sortPairsStateBytes(uint32_t count)
{
    return sortStateBytes(count);
}
int sortPairsDevice(Context* ctx, uint32_t* keysA, uint32_t* valsA, uint32_t* keysB, uint32_t* valsB, uint32_t* state, uint32_t count)
{
    if (count == 0)
    {
        return VKRT_SUCCESS;
    }
    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(state, 0, sortStateBytes(count), ctx->stream));
    k_onesweep_hist<<<gridFor(count, kThreads, ctx->smCount), kThreads, 0, ctx->stream>>>(keysA, count, state);
    ctx->launches += 1;
    return radixSort(ctx, keysA, valsA, keysB, valsB, state, count);
}

// Debug entry used by the tests: sorts host arrays on the device with the builder's sort.
int sortPairsForTest(Context* ctx, uint32_t* keys, uint32_t* vals, uint32_t count)
{
    if (count == 0)
    {
        return VKRT_SUCCESS;
    }
    DeviceBuffer a, b, c, d, h;
    if (count >= 0x40000000u)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "sort: at most 2^30-1 keys");
    }
    VKRT_CUDA_TRY(ctx, a.reserve(count * 4ull));
    VKRT_CUDA_TRY(ctx, b.reserve(count * 4ull));
    VKRT_CUDA_TRY(ctx, c.reserve(count * 4ull));
    VKRT_CUDA_TRY(ctx, d.reserve(count * 4ull));
    VKRT_CUDA_TRY(ctx, h.reserve(sortStateBytes(count)));
    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(h.ptr, 0, sortStateBytes(count), ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(a.ptr, keys, count * 4ull, cudaMemcpyHostToDevice, ctx->stream));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(b.ptr, vals, count * 4ull, cudaMemcpyHostToDevice, ctx->stream));
    k_onesweep_hist<<<gridFor(count, kThreads, ctx->smCount), kThreads, 0, ctx->stream>>>(a.as<uint32_t>(), count, h.as<uint32_t>());
    int r = radixSort(ctx, a.as<uint32_t>(), b.as<uint32_t>(), c.as<uint32_t>(), d.as<uint32_t>(), h.as<uint32_t>(), count);
    if (r == VKRT_SUCCESS)
    {
        cudaMemcpyAsync(keys, a.ptr, count * 4ull, cudaMemcpyDeviceToHost, ctx->stream);
        cudaMemcpyAsync(vals, b.ptr, count * 4ull, cudaMemcpyDeviceToHost, ctx->stream);
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess)
        {
            r = ctx->fail(VKRT_ERROR_CUDA, cudaGetErrorString(e));
        }
    }
    a.release(), b.release(), c.release(), d.release(), h.release();
    return r;
}

// ------------------------------------------------------------------------------------------------
// Top-level input: world boxes of the instances (replaces the instance-buffer read of the TLAS build,
// src/Raytracer.cpp:1179-1215).  Same outputs as k_tri_bounds: one box per primitive + scene / centroid bounds.
// ------------------------------------------------------------------------------------------------
namespace
{
__global__ void __launch_bounds__(kThreads)
k_instance_bounds(const InstanceRecord* __restrict__ instances, uint32_t count, float4* __restrict__ triBox, uint32_t* sceneBounds)
{
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        float l[3], h[3];
        instanceWorldBounds(instances[i], l, h);
        triBox[2 * (size_t)i] = make_float4(l[0], l[1], l[2], 0.0f);
        triBox[2 * (size_t)i + 1] = make_float4(h[0], h[1], h[2], 0.0f);
        for (int k = 0; k < 3; ++k)
        {
            const float c = 0.5f * (l[k] + h[k]);
            lo[k] = fminf(lo[k], l[k]), hi[k] = fmaxf(hi[k], h[k]);
            clo[k] = fminf(clo[k], c), chi[k] = fmaxf(chi[k], c);
        }
    }
    for (int k = 0; k < 3; ++k)
    {
        const float a = warpMin(lo[k]), b = warpMax(hi[k]), c = warpMin(clo[k]), d = warpMax(chi[k]);
        if ((threadIdx.x & 31) == 0)
        {
            atomicMin(&sceneBounds[k], orderedFromFloat(a));
            atomicMax(&sceneBounds[3 + k], orderedFromFloat(b));
            atomicMin(&sceneBounds[6 + k], orderedFromFloat(c));
            atomicMax(&sceneBounds[9 + k], orderedFromFloat(d));
        }
    }
}

// Top-level leaves: Node8::reserved = mask of leaf slots (one instance each), see traverse.cuh.
__global__ void __launch_bounds__(kThreads) k_tlas_finalize(Node8* nodes, uint32_t count)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count)
    {
        const uint32_t bits = nodes[i].leafBits;
        uint32_t m = 0;
        for (int sl = 0; sl < 8; ++sl)
        {
            m |= ((bits >> (3 * sl)) & 1u) << sl;
        }
        nodes[i].reserved = (uint8_t)m;
    }
}

// smallest scale exponent over all nodes of a hierarchy (-> boxScaleFromExponent); counters[3], preset to 255 by k_init_build
__global__ void __launch_bounds__(kThreads) k_min_scale_exponent(const Node8* __restrict__ nodes, uint32_t count, uint32_t* out)
{
    uint32_t e = 255u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const Node8& n = nodes[i];
        if (n.imask != 0 || (n.leafBits & 0xFFFFFFu) != 0u)
        {
            e = min(e, min(min((uint32_t)n.sx >> 7, (uint32_t)n.sy >> 7), (uint32_t)n.sz >> 7));
        }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        e = min(e, __shfl_xor_sync(0xFFFFFFFFu, e, o));
    }
    if ((threadIdx.x & 31) == 0 && e < 255u)
    {
        atomicMin(out, e);
    }
}

// the same with the node count read on the device (graph-replayed builds never learn it on the host in between)
__global__ void __launch_bounds__(kThreads) k_min_scale_exponent_dev(const Node8* __restrict__ nodes, const uint32_t* __restrict__ countPtr,
                                                                      uint32_t* out)
{
    const uint32_t count = *countPtr;
    uint32_t e = 255u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    {
        const Node8& n = nodes[i];
        if (n.imask != 0 || (n.leafBits & 0xFFFFFFu) != 0u)
        {
            e = min(e, min(min((uint32_t)n.sx >> 7, (uint32_t)n.sy >> 7), (uint32_t)n.sz >> 7));
        }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        e = min(e, __shfl_xor_sync(0xFFFFFFFFu, e, o));
    }
    if ((threadIdx.x & 31) == 0 && e < 255u)
    {
        atomicMin(out, e);
    }
}

struct HierarchyResult
{
    uint32_t nodeCount = 0;
    uint32_t minExponent = 127; // smallest scale exponent of any node
    uint32_t levels = 0; // wide levels = longest root-to-leaf node path
    float boundsMin[3] = {0, 0, 0}, boundsMax[3] = {0, 0, 0};
    float sahCost = 0.0f;
    uint32_t reinsertMoves = 0; // VKRT_BUILD_PREFER_FAST_TRACE: nodes moved by the reinsertion passes
};

int reserveBuildScratch(Context* ctx, uint32_t n, uint32_t nodeCapacity, bool packed)
{
    BuildScratch& sc = ctx->buildScratch;
    const size_t n1 = std::max<size_t>(n, 1);
    VKRT_CUDA_TRY(ctx, sc.triBox.reserve(n1 * 2 * sizeof(float4)));
    if (packed)
    {
        VKRT_CUDA_TRY(ctx, sc.packedBuild.reserve(n1 * sizeof(PackedTri)));
    }
    VKRT_CUDA_TRY(ctx, sc.keysA.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.keysB.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.valsA.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.valsB.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.tileHist.reserve(sortStateBytes(n)));
    VKRT_CUDA_TRY(ctx, sc.binNodes.reserve(n1 * sizeof(BinNode)));
    VKRT_CUDA_TRY(ctx, sc.flags.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.taskRef.reserve((size_t)nodeCapacity * 4));
    VKRT_CUDA_TRY(ctx, sc.primSource.reserve(n1 * 4));
    VKRT_CUDA_TRY(ctx, sc.small.reserve(256));
    return VKRT_SUCCESS;
}

// The part of the builder that only sees boxes: Morton codes -> onesweep sort -> PLOC / LBVH -> (cost-optimal)
// collapse into `nodes`.  On entry buildScratch.triBox holds the n boxes and `small` the reduced scene bounds
// (k_tri_bounds / k_instance_bounds, stream-ordered).  On return buildScratch.primSource[0..n) is the primitive order
// the leaves refer to, and everything is complete on the host side (the function synchronises).
int buildHierarchy(Context* ctx, uint32_t n, uint32_t maxLeaf, bool fast, bool fastTrace, Node8* nodes, uint32_t nodeCapacity,
                   HierarchyResult& out, bool noSync = false)
{
    cudaStream_t s = ctx->stream;
    const int sm = ctx->smCount;
    BuildScratch& sc = ctx->buildScratch;
    const size_t n1 = std::max<size_t>(n, 1);
    uint32_t* sceneBounds = sc.small.as<uint32_t>();           // 12 uints
    uint32_t* counters = sc.small.as<uint32_t>() + 16;         // [0] nodes, [1] prims, [2] overflow
    float* sahAccum = reinterpret_cast<float*>(sc.small.as<uint32_t>() + 32);
    int* root = reinterpret_cast<int*>(sc.small.as<uint32_t>() + 40);

    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(sc.tileHist.ptr, 0, sortStateBytes(n), s));
    k_morton_hist<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(sc.triBox.as<float4>(), sceneBounds, n,
                                                                   sc.keysA.as<uint32_t>(), sc.valsA.as<uint32_t>(),
                                                                   sc.tileHist.as<uint32_t>());
    ctx->launches += 1;
    int r = radixSort(ctx, sc.keysA.as<uint32_t>(), sc.valsA.as<uint32_t>(), sc.keysB.as<uint32_t>(),
                      sc.valsB.as<uint32_t>(), sc.tileHist.as<uint32_t>(), n);
    if (r != VKRT_SUCCESS)
    {
        return r;
    }

    // cost-optimal collapse (default; the fast-build path keeps the greedy one)
    float dpPrimCost = fast ? 0.0f : kDpPrimCost;
    if (const char* e = getenv("VKRT_BUILD_DP_PRIM_COST"))
    {
        dpPrimCost = (float)atof(e);
    }
    const bool useDp = n > maxLeaf && n > 1 && dpPrimCost > 0.0f;
    // compact 32-byte binary nodes: LBVH + greedy collapse only (the cost-optimal collapse keeps its costs in the padding of
    // the 64-byte record, PLOC its leaf counts in lo.w)
    const bool compact = fast && !useDp && n > 1 && n <= kCompactMaxPrims && !getenv("VKRT_BUILD_NO_COMPACT");
    if (compact)
    {
        VKRT_CUDA_TRY(ctx, sc.sortedBox.reserve(n1 * 2 * sizeof(float4)));
    }
    Bvh2 bvh;
    bvh.dpChoice = nullptr, bvh.dpParent = nullptr, bvh.dpPrimCost = 0.0f;
    bvh.maxLeaf = maxLeaf;
    bvh.nodes32 = compact ? sc.binNodes.as<BinNode32>() : nullptr;
    bvh.sortedBox = compact ? sc.sortedBox.as<float4>() : nullptr;
    bvh.nodes = sc.binNodes.as<BinNode>();
    bvh.flags = sc.flags.as<int>();
    bvh.keys = sc.keysA.as<uint32_t>();
    bvh.vals = sc.valsA.as<uint32_t>();
    bvh.triBox = sc.triBox.as<float4>();
    bvh.count = n;
    bvh.root = root;
    if (n > 1 && fast)
    {
        VKRT_CUDA_TRY(ctx, cudaMemsetAsync(bvh.flags, 0xFF, (size_t)(n - 1) * 4, s)); // -1
        k_lbvh<<<(n + kThreads - 1) / kThreads, kThreads, 0, s>>>(bvh);
        ctx->launches += 1;
    }
    else if (n > 1)
    {
        const uint32_t statusCount = (n + kPlocTile - 1) / kPlocTile;
        VKRT_CUDA_TRY(ctx, sc.plocLoA.reserve(n1 * sizeof(float4)));
        VKRT_CUDA_TRY(ctx, sc.plocHiA.reserve(n1 * sizeof(float4)));
        VKRT_CUDA_TRY(ctx, sc.plocLoB.reserve(n1 * sizeof(float4)));
        VKRT_CUDA_TRY(ctx, sc.plocHiB.reserve(n1 * sizeof(float4)));
        VKRT_CUDA_TRY(ctx, sc.plocStatus.reserve((size_t)statusCount * 4 + 64));
        PlocClusters cl[2];
        cl[0].id = sc.flags.as<int>(), cl[0].lo = sc.plocLoA.as<float4>(), cl[0].hi = sc.plocHiA.as<float4>();
        cl[1].id = reinterpret_cast<int*>(sc.keysB.as<uint32_t>()), cl[1].lo = sc.plocLoB.as<float4>(), cl[1].hi = sc.plocHiB.as<float4>();
        uint32_t* state = sc.plocStatus.as<uint32_t>(); // 16 words of state, then the look-back words
        uint32_t* status = state + 16;
        k_ploc_init<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(bvh, cl[0], state, status, statusCount);
        ctx->launches += 1;
        int round = 0;
        int batch = 8;
        for (uint32_t m = n; m > (uint32_t)kPlocTile; m >>= 1)
        {
            batch += 4; // measured: a round leaves ~84 % of the clusters; the last <= 1024 take one launch
        }
        for (;;)
        {
            for (int k = 0; k < batch; ++k, ++round)
            {
                k_ploc_round<<<(uint32_t)sm * 4u, kPlocThreads, 0, s>>>(cl[round & 1], cl[(round + 1) & 1], bvh, state, status, round & 1);
            }
            k_ploc_finish<<<1, 1, 0, s>>>(cl[round & 1], bvh, state, round & 1, state + 8);
            ctx->launches += batch + 1;
            uint32_t remaining = 0;
            VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(&remaining, state + 8, 4, cudaMemcpyDeviceToHost, s));
            VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
            if (remaining == 1)
            {
                break;
            }
            if (round > 4096)
            {
                return ctx->fail(VKRT_ERROR_CUDA, "PLOC: did not converge");
            }
            batch = 16;
        }
    }

    // quality passes on the binary hierarchy (VKRT_BUILD_PREFER_FAST_TRACE): parallel reinsertion, large nodes first
    out.reinsertMoves = 0;
    if (fastTrace && !compact && n > 3)
    {
        const uint32_t slots = 2u * n - 1u;
        VKRT_CUDA_TRY(ctx, sc.dp.reserve((size_t)n * 12));
        VKRT_CUDA_TRY(ctx, sc.reinsert.reserve((size_t)slots * 16 + 64));
        bvh.dpParent = reinterpret_cast<int*>(sc.dp.as<uint32_t>() + n);
        ReinsertArgs ra;
        ra.bvh = bvh;
        ra.lock = sc.reinsert.as<unsigned long long>();
        ra.target = reinterpret_cast<int*>(ra.lock + slots);
        ra.gain = reinterpret_cast<float*>(ra.target + slots);
        uint32_t* state = reinterpret_cast<uint32_t*>(ra.gain + slots);
        ra.moved = state + 1;
        ra.areaLo = 0.0f, ra.areaHi = FLT_MAX;
        k_dp_parents<<<(n + kThreads - 1) / kThreads, kThreads, 0, s>>>(bvh);
        k_reinsert_begin<<<1, 1, 0, s>>>(bvh, state);
        ctx->launches += 2;
        int sweeps = kReinsertSweeps, bands = kReinsertBands, iters = kReinsertIterations;
        float ratio = kReinsertBandRatio;
        if (const char* e = getenv("VKRT_BUILD_REINSERT")) // experiments: "sweeps,bands,iterations,ratio"
        {
            sscanf(e, "%d,%d,%d,%f", &sweeps, &bands, &iters, &ratio);
        }
        const uint32_t gridSlots = (slots + kThreads - 1) / kThreads, gridLeaves = (n + kThreads - 1) / kThreads;
        for (int sweep = 0; sweep < sweeps; ++sweep)
        {
            for (int band = 0; band < bands; ++band)
            {
                const float hiMul = band == 0 ? FLT_MAX : powf(ratio, -(float)band);
                const float loMul = band == bands - 1 ? 0.0f : powf(ratio, -(float)(band + 1));
                for (int it = 0; it < iters; ++it)
                {
                    k_reinsert_search<<<(slots + 127) / 128, 128, 0, s>>>(ra, state, hiMul, loMul);
                    k_reinsert_lock<<<gridSlots, kThreads, 0, s>>>(ra);
                    k_reinsert_verify<<<gridSlots, kThreads, 0, s>>>(ra);
                    k_reinsert_apply<<<gridSlots, kThreads, 0, s>>>(ra);
                    VKRT_CUDA_TRY(ctx, cudaMemsetAsync(bvh.flags, 0, (size_t)(n - 1) * 4, s));
                    k_reinsert_refit<<<gridLeaves, kThreads, 0, s>>>(bvh);
                    ctx->launches += 5;
                }
            }
        }
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(&out.reinsertMoves, ra.moved, 4, cudaMemcpyDeviceToHost, s));
        bvh.dpParent = nullptr;
    }

    // cost-optimal collapse: parents, then a bottom-up climb
    if (useDp)
    {
        VKRT_CUDA_TRY(ctx, sc.dp.reserve((size_t)n * 12));
        bvh.dpChoice = sc.dp.as<uint32_t>();
        bvh.dpParent = reinterpret_cast<int*>(sc.dp.as<uint32_t>() + n);
        bvh.dpPrimCost = dpPrimCost;
        VKRT_CUDA_TRY(ctx, cudaMemsetAsync(bvh.flags, 0, (size_t)(n - 1) * 4, s));
        k_dp_parents<<<(n + kThreads - 1) / kThreads, kThreads, 0, s>>>(bvh);
        k_dp_climb<<<(n + kThreads - 1) / kThreads, kThreads, 0, s>>>(bvh);
        ctx->launches += 2;
    }

    CollapseArgs ca;
    ca.bvh2 = bvh;
    ca.nodes = nodes;
    ca.taskRef = sc.taskRef.as<int>();
    ca.primSource = sc.primSource.as<uint32_t>();
    ca.nodeCounter = counters + 0;
    ca.primCounter = counters + 1;
    ca.overflow = counters + 2;
    ca.sahAccum = sahAccum;
    ca.nodeCapacity = nodeCapacity;
    uint32_t* range = counters + 4; // {first, last, ticket, levels}
    k_collapse_seed<<<1, 1, 0, s>>>(ca, range);
    ctx->launches += 1;
    uint32_t last = 0;
    // levels of at most narrowMax nodes are swept by one block without leaving the kernel; larger ones take a launch each
    uint32_t narrowMax = kNarrowMax;
    int widePairs = 8;
    if (const char* e = getenv("VKRT_BUILD_NARROW_MAX")) // experiments
    {
        narrowMax = std::min<uint32_t>(std::max(atoi(e), 8), kNarrowLevel);
    }
    if (const char* e = getenv("VKRT_BUILD_WIDE_PAIRS")) // experiments
    {
        widePairs = std::min(std::max(atoi(e), 1), 32);
    }
    if (noSync)
    {
        // for a CUDA graph: enough (wide, narrow) pairs for any hierarchy of this size class, no read-back in between; the
        // caller reads the counters once at the end and takes the normal path if the collapse had not finished
        k_collapse_narrow<<<1, kNarrowLevel, 0, s>>>(ca, range, narrowMax);
        const int pairs = getenv("VKRT_BUILD_WIDE_PAIRS") ? widePairs : kGraphWidePairs;
        for (int k = 0; k < pairs; ++k)
        {
            k_collapse_wide<<<(uint32_t)sm * 8u, 128, 0, s>>>(ca, range, narrowMax);
            k_collapse_narrow<<<1, kNarrowLevel, 0, s>>>(ca, range, narrowMax);
        }
        k_min_scale_exponent_dev<<<gridFor(nodeCapacity, kThreads, sm), kThreads, 0, s>>>(nodes, counters + 0, counters + 3);
        ctx->launches += 2 + 2 * pairs;
        VKRT_CUDA_TRY(ctx, cudaGetLastError());
        return VKRT_SUCCESS;
    }
    for (int batch = 0;; ++batch)
    {
        // one narrow sweep, then (wide, narrow) pairs: enough for 8 wide levels per batch
        k_collapse_narrow<<<1, kNarrowLevel, 0, s>>>(ca, range, narrowMax);
        ctx->launches += 1;
        if (n > narrowMax) // with fewer primitives than one narrow level no level can be wide
        {
            for (int k = 0; k < widePairs; ++k)
            {
                k_collapse_wide<<<(uint32_t)sm * 8u, 128, 0, s>>>(ca, range, narrowMax);
                k_collapse_narrow<<<1, kNarrowLevel, 0, s>>>(ca, range, narrowMax);
            }
            ctx->launches += 2 * widePairs;
        }
        uint32_t hostCounters[8];
        VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(hostCounters, counters, sizeof(hostCounters), cudaMemcpyDeviceToHost, s));
        VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
        if (hostCounters[2])
        {
            return ctx->fail(VKRT_ERROR_OUT_OF_MEMORY, "BVH collapse: wide node capacity exceeded");
        }
        if (hostCounters[4] == hostCounters[5]) // first == last: no wide node left to fill
        {
            last = hostCounters[0];
            out.levels = hostCounters[7];
            if (hostCounters[1] != n)
            {
                return ctx->fail(VKRT_ERROR_CUDA, "BVH collapse: packed " + std::to_string(hostCounters[1]) + " of " +
                                                      std::to_string(n) + " primitives");
            }
            break;
        }
        if (batch > 64)
        {
            return ctx->fail(VKRT_ERROR_CUDA, "BVH collapse: did not terminate");
        }
    }
    out.nodeCount = last;
    k_min_scale_exponent<<<gridFor(last, kThreads, sm), kThreads, 0, s>>>(nodes, last, counters + 3);
    ctx->launches += 1;
    uint32_t minExp = 255u;
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(&minExp, counters + 3, 4, cudaMemcpyDeviceToHost, s));

    uint32_t hb[12];
    float hs[2];
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(hb, sceneBounds, sizeof(hb), cudaMemcpyDeviceToHost, s));
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(hs, sahAccum, sizeof(hs), cudaMemcpyDeviceToHost, s));
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    for (int k = 0; k < 3; ++k)
    {
        out.boundsMin[k] = floatFromOrdered(hb[k]);
        out.boundsMax[k] = floatFromOrdered(hb[3 + k]);
    }
    const float dx = out.boundsMax[0] - out.boundsMin[0], dy = out.boundsMax[1] - out.boundsMin[1],
                dz = out.boundsMax[2] - out.boundsMin[2];
    const float rootArea = dx * dy + dy * dz + dz * dx;
    out.sahCost = rootArea > 0.0f ? (hs[0] + hs[1]) / rootArea : 0.0f;
    out.minExponent = minExp == 255u ? 127u : minExp;
    return VKRT_SUCCESS;
}
} // namespace

// Small fast builds as ONE CUDA graph.  A Sponza-sized LBVH build is ~40 launches of 2-50 us each and three host round trips:
// launch gaps and synchronisation are a third of its 0.6 ms.  The chain is captured without any read-back in between (fixed
// number of collapse launches, node count read on the device) and replayed by later rebuilds of the same BLAS; the counters
// come back in one copy at the end.  `done` stays false if the collapse had not finished within the captured launches (the
// caller then builds the normal way).
constexpr uint32_t kGraphMaxPrims = 4u << 20;

static int graphBuild(Context* ctx, Blas* blas, const GeometryTable& geo, uint32_t n, uint32_t nodeCapacity, HierarchyResult& h, bool& done)
{
    done = false;
    cudaStream_t s = ctx->stream;
    const int sm = ctx->smCount;
    BuildScratch& sc = ctx->buildScratch;
    if (!ctx->buildReadback)
    {
        VKRT_CUDA_TRY(ctx, cudaMallocHost(&ctx->buildReadback, 256));
    }
    VKRT_CUDA_TRY(ctx, sc.sortedBox.reserve(std::max<size_t>(n, 1) * 2 * sizeof(float4))); // before the key: capture must not allocate
    uint32_t* sceneBounds = sc.small.as<uint32_t>();
    uint32_t* counters = sc.small.as<uint32_t>() + 16;
    float* sahAccum = reinterpret_cast<float*>(sc.small.as<uint32_t>() + 32);
    int* root = reinterpret_cast<int*>(sc.small.as<uint32_t>() + 40);
    auto addr = [](const DeviceBuffer& b) { return (uint64_t)(uintptr_t)b.ptr; };
    const uint64_t key[24] = {n, blas->flags, nodeCapacity, blas->vertexStride, blas->geometryCount, addr(blas->vertices), addr(blas->indices),
                              addr(blas->geomFirstTri), addr(blas->geomIndexBase), addr(blas->nodes), addr(blas->tris), addr(sc.triBox),
                              addr(sc.packedBuild), addr(sc.keysA), addr(sc.keysB), addr(sc.valsA), addr(sc.valsB), addr(sc.tileHist),
                              addr(sc.binNodes), addr(sc.flags), addr(sc.taskRef), addr(sc.primSource), addr(sc.small), addr(sc.sortedBox)};
    if (!blas->buildGraph || memcmp(key, blas->buildGraphKey, sizeof(key)) != 0)
    {
        if (blas->buildGraph)
        {
            cudaGraphExecDestroy(blas->buildGraph);
            blas->buildGraph = nullptr;
        }
        const uint64_t launchesBefore = ctx->launches;
        VKRT_CUDA_TRY(ctx, cudaStreamBeginCapture(s, cudaStreamCaptureModeRelaxed));
        k_init_build<<<1, 32, 0, s>>>(sceneBounds, counters, sahAccum, root);
        k_tri_bounds<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(geo, blas->vertices.as<uint8_t>(), blas->vertexStride,
                                                                      blas->indices.as<uint32_t>(), n, sc.triBox.as<float4>(),
                                                                      sc.packedBuild.as<float4>(), sceneBounds);
        ctx->launches += 2;
        HierarchyResult unused;
        int r = buildHierarchy(ctx, n, kMaxLeafTris, true, false, blas->nodes.as<Node8>(), nodeCapacity, unused, true);
        k_pack_tris<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(sc.packedBuild.as<float4>(), sc.primSource.as<uint32_t>(), n,
                                                                     blas->tris.as<PackedTri>());
        ctx->launches += 1;
        cudaMemcpyAsync(ctx->buildReadback, sc.small.ptr, 192, cudaMemcpyDeviceToHost, s);
        cudaGraph_t graph = nullptr;
        const cudaError_t e = cudaStreamEndCapture(s, &graph);
        blas->buildGraphLaunches = (uint32_t)(ctx->launches - launchesBefore);
        ctx->launches = launchesBefore;
        if (r != VKRT_SUCCESS || e != cudaSuccess || !graph)
        {
            if (graph)
            {
                cudaGraphDestroy(graph);
            }
            cudaGetLastError();
            return VKRT_SUCCESS; // not an error: the normal path builds
        }
        const cudaError_t ei = cudaGraphInstantiate(&blas->buildGraph, graph, 0);
        cudaGraphDestroy(graph);
        if (ei != cudaSuccess)
        {
            blas->buildGraph = nullptr;
            cudaGetLastError();
            return VKRT_SUCCESS;
        }
        memcpy(blas->buildGraphKey, key, sizeof(key));
    }
    VKRT_CUDA_TRY(ctx, cudaGraphLaunch(blas->buildGraph, s));
    ctx->launches += blas->buildGraphLaunches;
    VKRT_CUDA_TRY(ctx, cudaStreamSynchronize(s));
    const uint32_t* w = static_cast<const uint32_t*>(ctx->buildReadback);
    const uint32_t* c = w + 16;
    if (c[2] != 0u || c[4] != c[5] || c[1] != n)
    {
        return VKRT_SUCCESS; // capacity exceeded or more wide levels than the captured launches cover: the normal path decides
    }
    h = HierarchyResult{};
    h.nodeCount = c[0];
    h.levels = c[7];
    h.minExponent = c[3] == 255u ? 127u : c[3];
    for (int k = 0; k < 3; ++k)
    {
        h.boundsMin[k] = floatFromOrdered(w[k]);
        h.boundsMax[k] = floatFromOrdered(w[3 + k]);
    }
    const float dx = h.boundsMax[0] - h.boundsMin[0], dy = h.boundsMax[1] - h.boundsMin[1], dz = h.boundsMax[2] - h.boundsMin[2];
    const float rootArea = dx * dy + dy * dz + dz * dx;
    float sah[2];
    memcpy(sah, w + 32, sizeof(sah));
    h.sahCost = rootArea > 0.0f ? (sah[0] + sah[1]) / rootArea : 0.0f;
    done = true;
    return VKRT_SUCCESS;
}

// Runs the whole builder over the BLAS's device geometry copy.
int buildBlasOnDevice(Context* ctx, Blas* blas)
{
    const uint32_t n = blas->triangleCount;
    cudaStream_t s = ctx->stream;
    const int sm = ctx->smCount;

    cudaEvent_t ev0, ev1;
    VKRT_CUDA_TRY(ctx, cudaEventCreate(&ev0));
    VKRT_CUDA_TRY(ctx, cudaEventCreate(&ev1));
    VKRT_CUDA_TRY(ctx, cudaEventRecord(ev0, s));

    // node capacity: every wide node with internal children covers > 3 triangles per child
    const uint32_t nodeCapacity = kMaxLeafTris >= 3 ? std::max(1u, n / 2u + 8u) : n + 8u;
    VKRT_CUDA_TRY(ctx, blas->nodes.reserve((size_t)nodeCapacity * sizeof(Node8)));
    VKRT_CUDA_TRY(ctx, blas->tris.reserve(std::max<size_t>(1, n) * sizeof(PackedTri)));
    if (int r = reserveBuildScratch(ctx, n, nodeCapacity, true))
    {
        return r;
    }
    BuildScratch& sc = ctx->buildScratch;
    uint32_t* sceneBounds = sc.small.as<uint32_t>();
    uint32_t* counters = sc.small.as<uint32_t>() + 16;
    float* sahAccum = reinterpret_cast<float*>(sc.small.as<uint32_t>() + 32);
    int* root = reinterpret_cast<int*>(sc.small.as<uint32_t>() + 40);
    k_init_build<<<1, 32, 0, s>>>(sceneBounds, counters, sahAccum, root);
    ctx->launches += 1;

    if (n == 0)
    {
        k_empty_root<<<1, 1, 0, s>>>(blas->nodes.as<Node8>());
        ctx->launches += 1;
        blas->nodeCount = 1;
        blas->levels = 1;
        blas->boxScale = 1.0f;
        for (int k = 0; k < 3; ++k)
        {
            blas->boundsMin[k] = 0.0f;
            blas->boundsMax[k] = 0.0f;
        }
        blas->sahCost = 0.0f;
    }
    else
    {
        GeometryTable geo;
        geo.firstTri = blas->geomFirstTri.as<uint32_t>();
        geo.indexBase = blas->geomIndexBase.as<uint64_t>();
        geo.geometryCount = blas->geometryCount;
        HierarchyResult h;
        const bool fast = (blas->flags & VKRT_BUILD_PREFER_FAST_BUILD) != 0;
        bool fastTrace = !fast && ((blas->flags & VKRT_BUILD_PREFER_FAST_TRACE) != 0 || getenv("VKRT_BUILD_FAST_TRACE") != nullptr);
        bool graphDone = false;
        // (rebuilds only: capturing and instantiating costs about as much as one small build)
        if (fast && blas->generation > 0 && n >= 2 && n <= kGraphMaxPrims && !getenv("VKRT_BUILD_NO_GRAPH") && !getenv("VKRT_BUILD_DP_PRIM_COST") &&
            !getenv("VKRT_BUILD_NO_COMPACT"))
        {
            if (int r = graphBuild(ctx, blas, geo, n, nodeCapacity, h, graphDone))
            {
                return r;
            }
            if (!graphDone)
            {
                k_init_build<<<1, 32, 0, s>>>(sceneBounds, counters, sahAccum, root);
                ctx->launches += 1;
            }
        }
        for (; !graphDone;)
        {
            k_tri_bounds<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(geo, blas->vertices.as<uint8_t>(), blas->vertexStride,
                                                                          blas->indices.as<uint32_t>(), n, sc.triBox.as<float4>(),
                                                                          sc.packedBuild.as<float4>(), sceneBounds);
            ctx->launches += 1;
            h = HierarchyResult{};
            if (int r = buildHierarchy(ctx, n, kMaxLeafTris, fast, fastTrace, blas->nodes.as<Node8>(), nodeCapacity, h))
            {
                return r;
            }
            const char* forced = getenv("VKRT_BUILD_TEST_DEPTH_FALLBACK"); // test hook: pretend the reinsertion made the tree too deep
            if (fastTrace && ((int)h.levels > kMaxBlasLevels || forced))
            {
                // the reinsertion passes do not bound the depth: build again without them rather than refuse the scene
                fastTrace = false;
                k_init_build<<<1, 32, 0, s>>>(sceneBounds, counters, sahAccum, root);
                ctx->launches += 1;
                continue;
            }
            break;
        }
        if ((int)h.levels > kMaxBlasLevels)
        {
            return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_blas_build: hierarchy has " + std::to_string(h.levels) +
                                                              " levels, more than the traversal stack holds (" +
                                                              std::to_string(kMaxBlasLevels) + ")");
        }
        blas->nodeCount = h.nodeCount;
        blas->levels = h.levels;
        blas->boxScale = boxScaleFromExponent(h.minExponent);
        if (!graphDone) // (part of the graph otherwise)
        {
            k_pack_tris<<<gridFor(n, kThreads, sm), kThreads, 0, s>>>(sc.packedBuild.as<float4>(), sc.primSource.as<uint32_t>(), n,
                                                                         blas->tris.as<PackedTri>());
            ctx->launches += 1;
        }
        for (int k = 0; k < 3; ++k)
        {
            blas->boundsMin[k] = h.boundsMin[k];
            blas->boundsMax[k] = h.boundsMax[k];
        }
        blas->sahCost = h.sahCost;
    }

    VKRT_CUDA_TRY(ctx, cudaEventRecord(ev1, s));
    VKRT_CUDA_TRY(ctx, cudaEventSynchronize(ev1));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    blas->buildMs = ms;
    blas->generation += 1;
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    return VKRT_SUCCESS;
}

// Top level: instance world boxes (k_instance_bounds) -> the same hierarchy builder with one instance per leaf slot.
// tlas->instances must hold the uploaded records.  A single instance needs no top level (tlas->nodeCount = 0).
int buildTlasOnDevice(Context* ctx, Tlas* tlas)
{
    const uint32_t n = tlas->instanceCount;
    cudaStream_t s = ctx->stream;
    tlas->nodeCount = 0;
    tlas->levels = 0;
    if (n <= 1)
    {
        return VKRT_SUCCESS;
    }
    cudaEvent_t ev0, ev1;
    VKRT_CUDA_TRY(ctx, cudaEventCreate(&ev0));
    VKRT_CUDA_TRY(ctx, cudaEventCreate(&ev1));
    VKRT_CUDA_TRY(ctx, cudaEventRecord(ev0, s));
    const uint32_t nodeCapacity = n + 8u;
    VKRT_CUDA_TRY(ctx, tlas->nodes.reserve((size_t)nodeCapacity * sizeof(Node8)));
    VKRT_CUDA_TRY(ctx, tlas->leafInstance.reserve((size_t)n * 4));
    if (int r = reserveBuildScratch(ctx, n, nodeCapacity, false))
    {
        return r;
    }
    BuildScratch& sc = ctx->buildScratch;
    uint32_t* sceneBounds = sc.small.as<uint32_t>();
    k_init_build<<<1, 32, 0, s>>>(sceneBounds, sc.small.as<uint32_t>() + 16, reinterpret_cast<float*>(sc.small.as<uint32_t>() + 32),
                                  reinterpret_cast<int*>(sc.small.as<uint32_t>() + 40));
    k_instance_bounds<<<gridFor(n, kThreads, ctx->smCount), kThreads, 0, s>>>(tlas->instances.as<InstanceRecord>(), n,
                                                                              sc.triBox.as<float4>(), sceneBounds);
    ctx->launches += 2;
    HierarchyResult h;
    if (int r = buildHierarchy(ctx, n, 1u, false, false, tlas->nodes.as<Node8>(), nodeCapacity, h))
    {
        return r;
    }
    k_tlas_finalize<<<(h.nodeCount + kThreads - 1) / kThreads, kThreads, 0, s>>>(tlas->nodes.as<Node8>(), h.nodeCount);
    ctx->launches += 1;
    VKRT_CUDA_TRY(ctx, cudaMemcpyAsync(tlas->leafInstance.ptr, sc.primSource.ptr, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    VKRT_CUDA_TRY(ctx, cudaEventRecord(ev1, s));
    VKRT_CUDA_TRY(ctx, cudaEventSynchronize(ev1));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, ev0, ev1);
    cudaEventDestroy(ev0);
    cudaEventDestroy(ev1);
    tlas->nodeCount = h.nodeCount;
    tlas->levels = h.levels;
    tlas->boxScale = boxScaleFromExponent(h.minExponent);
    tlas->buildMs = ms;
    tlas->sahCost = h.sahCost;
    for (int k = 0; k < 3; ++k)
    {
        tlas->boundsMin[k] = h.boundsMin[k];
        tlas->boundsMax[k] = h.boundsMax[k];
    }
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    return VKRT_SUCCESS;
}

} // namespace vkrt
