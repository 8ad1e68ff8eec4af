// raster.cu — CUDA kernels of the software rasteriser (logic in raster.cuh) that replaces the reference's `Rasterizer`
// draw loop: vkCmdDrawIndexed per submesh inside one render pass with 8x MSAA and a resolve (src/Rasterizer.cpp:119-147).
//
//   k_raster_clear     visibility keys <- empty                                     (vkCmdBeginRenderPass clears, :119-131)
//   k_raster_setup     one thread per triangle: vertex shader, cull, planes, bounding box; small triangles (bounding box
//                      of at most 32x32 pixels) and large ones go to two lists
//   k_raster_scan      exclusive scan of the large triangles' 16x16-pixel tile counts (one block)
//   k_raster_small     one warp per small triangle, lanes over the pixels of its bounding box
//   k_raster_large     one warp per (large triangle, tile): binary search in the scanned counts, conservative tile
//                      rejection against the three edges, lanes over the 256 pixels of the tile
//   k_raster_resolve   one thread per pixel: fragment colour of each sample's winner, 8-sample average, B8G8R8A8 store
//
// Triangles write 64-bit (depth, triangle) keys with atomicMin, so no ordering between warps is needed.
// Bytes per frame: 96 B set-up write + ~76 B vertex/index read per triangle; 64 B of keys per pixel cleared, ~64 B read
// by the resolve, plus the atomics of the covered samples - the key buffer (59 MB at 720p, 531 MB at 4K) is the HBM term.
#include "context.hpp"
#include "builder.cuh"
#include "raster.cuh"

#include <algorithm>
#include <cstdlib>

namespace vkrt
{

namespace
{
constexpr uint32_t kSmallExtent = 32; // bounding boxes up to 32x32 pixels are walked by one warp
constexpr uint32_t kRasterItemCapacity = 4u << 20; // explicit (triangle, tile) work list of the binned large-triangle path, 8 bytes each

struct RasterDev
{
    const RasterSetup* setups;
    const TextureRecord* textures;
    const TextureMips* mips;
    unsigned long long* visibility;
    uint32_t width, height;
    // multi-GPU: this rank draws the pixels of the 64x64 image tiles it owns (tile % size == rank), like vkrt_trace
    uint32_t tilesX, rank, size, ownedTiles;
    // hierarchical Z: per 16x16 tile the smallest "every sample of this tile is at most this deep" bound published so far by
    // an opaque triangle that covers the whole tile; a triangle whose nearest depth in the tile lies behind it is skipped
    // without touching the keys.  Keys only ever decrease and the final key is a minimum, so skipping a certain loser is
    // independent of what has been written when.
    uint32_t* tileZ;
    uint32_t tilesX16, tilesY16;
    const uint32_t* minAlpha; // per image: smallest alpha byte (an occluder must not discard, shader.frag:19)
    uint32_t itemCapacity;    // entries of the explicit work list (kRasterItemCapacity; VKRT_RASTER_ITEMS overrides it for tests)
};

__device__ __forceinline__ bool ownsPixel(const RasterDev& r, uint32_t x, uint32_t y)
{
    return r.size == 1u || ((y / (uint32_t)kTile) * r.tilesX + x / (uint32_t)kTile) % r.size == r.rank;
}

struct Mat4
{
    float m[16];
};

struct AtomicMinKey
{
    __device__ __forceinline__ void operator()(uint64_t* key, uint64_t candidate) const
    {
        atomicMin(reinterpret_cast<unsigned long long*>(key), (unsigned long long)candidate);
    }
};

__global__ void __launch_bounds__(256) k_raster_clear(ulonglong2* __restrict__ keys, size_t pairs, uint32_t* counters, RasterDev r)
{
    if (r.size == 1u)
    {
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < pairs; i += (size_t)gridDim.x * blockDim.x)
        {
            keys[i] = make_ulonglong2(~0ull, ~0ull);
        }
    }
    else
    {
        // only the owned tiles: 4 key pairs (64 bytes) per pixel, rows of 64 pixels are contiguous
        const size_t total = (size_t)r.ownedTiles * kTilePixels * (kRasterSamples / 2);
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
        {
            const uint32_t pair = (uint32_t)(i % (kRasterSamples / 2));
            const size_t p = i / (kRasterSamples / 2);
            const uint32_t tile = r.rank + (uint32_t)(p / kTilePixels) * r.size, pit = (uint32_t)(p % kTilePixels);
            const uint32_t x = (tile % r.tilesX) * (uint32_t)kTile + pit % (uint32_t)kTile, y = (tile / r.tilesX) * (uint32_t)kTile + pit / (uint32_t)kTile;
            if (x < r.width && y < r.height)
            {
                keys[((size_t)y * r.width + x) * (kRasterSamples / 2) + pair] = make_ulonglong2(~0ull, ~0ull);
            }
        }
    }
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < r.tilesX16 * r.tilesY16; i += gridDim.x * blockDim.x)
    {
        r.tileZ[i] = 0xFFFFFFFFu;
    }
    if (blockIdx.x == 0 && threadIdx.x < 4)
    {
        counters[threadIdx.x] = 0u; // [0] small triangles, [1] large triangles, [2] large work items
    }
}

// smallest alpha byte of every image (level 0; the mip levels are rounded averages of it, so they stay within 1 of it)
__global__ void __launch_bounds__(256) k_texture_min_alpha(const TextureRecord* __restrict__ textures, uint32_t* __restrict__ minAlpha)
{
    const TextureRecord t = textures[blockIdx.y];
    const uint32_t n = t.width * t.height;
    uint32_t m = 255u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        m = min(m, (uint32_t)t.texels[4 * (size_t)i + 3]);
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        m = min(m, __shfl_xor_sync(0xFFFFFFFFu, m, o));
    }
    if ((threadIdx.x & 31) == 0)
    {
        atomicMin(&minAlpha[blockIdx.y], m);
    }
}

__global__ void __launch_bounds__(256)
k_raster_setup(const __grid_constant__ Mat4 wvpHolder, GeometryTable geo, const uint8_t* __restrict__ vertices, uint32_t stride,
               const uint32_t* __restrict__ indices, const VkrtSubmeshInfo* __restrict__ rows, uint32_t textureCount, uint32_t triangles, uint32_t width, uint32_t height,
               RasterSetup* __restrict__ setups, uint32_t* __restrict__ smallList, uint32_t* __restrict__ largeList, uint32_t* __restrict__ largeTiles,
               uint32_t* __restrict__ sortKeys, uint32_t* __restrict__ counters)
{
    const float* wvp = wvpHolder.m;
    for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < triangles; t += gridDim.x * blockDim.x)
    {
        const uint32_t geometry = findGeometry(geo, t);
        const uint32_t primitive = t - geo.firstTri[geometry];
        const uint64_t base = geo.indexBase[geometry] + 3ull * primitive;
        const uint32_t i0 = indices[base], i1 = indices[base + 1], i2 = indices[base + 2];
        const float4 p0 = loadF4(vertices, (size_t)i0 * stride), p1 = loadF4(vertices, (size_t)i1 * stride), p2 = loadF4(vertices, (size_t)i2 * stride);
        const float4 t0 = loadF4(vertices, (size_t)i0 * stride + 32), t1 = loadF4(vertices, (size_t)i1 * stride + 32),
                     t2 = loadF4(vertices, (size_t)i2 * stride + 32);
        RasterSetup s;
        // the reference binds materials[material].baseColor unchecked (src/Rasterizer.cpp:143); indices outside the table go to image 0
        const int32_t baseColor = rows[geometry].baseColorTextureIndex;
        const uint32_t texture = (baseColor < 0 || (uint32_t)baseColor >= textureCount) ? 0u : (uint32_t)baseColor;
        const bool drawn = rasterSetup(wvp, f3(p0.x, p0.y, p0.z), f3(p1.x, p1.y, p1.z), f3(p2.x, p2.y, p2.z), t0.x, t0.y, t1.x, t1.y, t2.x, t2.y,
                                       texture, width, height, s);
        float4* dst = reinterpret_cast<float4*>(setups + t);
        const float4* src = reinterpret_cast<const float4*>(&s);
#pragma unroll
        for (int k = 0; k < 6; ++k)
        {
            dst[k] = src[k];
        }
        if (sortKeys)
        {
            sortKeys[t] = 0xFFFFFFFFu, largeList[t] = t; // the sort moves the large triangles to the front, nearest first
        }
        if (!drawn)
        {
            continue;
        }
        const uint32_t x0 = s.bboxX & 0xFFFFu, x1 = s.bboxX >> 16, y0 = s.bboxY & 0xFFFFu, y1 = s.bboxY >> 16;
        if (x1 - x0 < kSmallExtent && y1 - y0 < kSmallExtent)
        {
            smallList[atomicAdd(&counters[0], 1u)] = t;
        }
        else if (sortKeys)
        {
            // nearest depth the triangle reaches on screen, as ordered bits: min over the bounding-box corners of the depth
            // plane (it is linear in x, y), clamped to [0, 1]
            const float fx0 = (float)x0, fx1 = (float)(x1 + 1u), fy0 = (float)y0, fy1 = (float)(y1 + 1u);
            const float zmin = fminf(fminf(plane3(s.za, s.zb, s.zc, fx0, fy0), plane3(s.za, s.zb, s.zc, fx1, fy0)),
                                     fminf(plane3(s.za, s.zb, s.zc, fx0, fy1), plane3(s.za, s.zb, s.zc, fx1, fy1)));
            sortKeys[t] = __float_as_uint(fminf(fmaxf(zmin, 0.0f), 1.0f));
            atomicAdd(&counters[1], 1u);
        }
        else
        {
            const uint32_t tiles = (x1 / kRasterTile - x0 / kRasterTile + 1u) * (y1 / kRasterTile - y0 / kRasterTile + 1u);
            const uint32_t slot = atomicAdd(&counters[1], 1u);
            largeList[slot] = t;
            largeTiles[slot] = tiles;
        }
    }
}

// Tile counts of the large triangles in their sorted order.
__global__ void __launch_bounds__(256)
k_raster_tiles(const RasterSetup* __restrict__ setups, const uint32_t* __restrict__ largeList, uint32_t* __restrict__ largeTiles,
               const uint32_t* __restrict__ counters)
{
    const uint32_t n = counters[1];
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        const RasterSetup* s = setups + largeList[i];
        const uint32_t bx = s->bboxX, by = s->bboxY;
        const uint32_t x0 = bx & 0xFFFFu, x1 = bx >> 16, y0 = by & 0xFFFFu, y1 = by >> 16;
        largeTiles[i] = (x1 / kRasterTile - x0 / kRasterTile + 1u) * (y1 / kRasterTile - y0 / kRasterTile + 1u);
    }
}

// Exclusive scan of largeTiles[0 .. counters[1]) in place; total -> counters[2].  One block of 1024 threads.
__global__ void __launch_bounds__(1024) k_raster_scan(uint32_t* __restrict__ largeTiles, uint32_t* __restrict__ counters)
{
    __shared__ uint32_t warpSums[32];
    __shared__ uint32_t carry;
    const uint32_t n = counters[1];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0)
    {
        carry = 0u;
    }
    __syncthreads();
    for (uint32_t base = 0; base < n; base += 1024u)
    {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < n ? largeTiles[i] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            incl += lane >= (uint32_t)d ? o : 0u;
        }
        if (lane == 31u)
        {
            warpSums[warp] = incl;
        }
        __syncthreads();
        if (warp == 0)
        {
            const uint32_t w = warpSums[lane];
            uint32_t wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1)
            {
                const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, wi, d);
                wi += lane >= (uint32_t)d ? o : 0u;
            }
            warpSums[lane] = wi - w; // exclusive
        }
        __syncthreads();
        const uint32_t excl = carry + warpSums[warp] + incl - v;
        if (i < n)
        {
            largeTiles[i] = excl;
        }
        __syncthreads();
        if (threadIdx.x == 1023u)
        {
            carry = excl + v;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0)
    {
        counters[2] = carry;
    }
}

// c / 255.0f for every byte value in shared memory (blocks of 256 threads): the texture filter of a fragment converts 32
// texels' channels; ncu showed the resolve kernel with the xu pipe (conversions, reciprocals of the IEEE divisions) at 58 %.
__device__ __forceinline__ void fillUnormTable(float* table)
{
    table[threadIdx.x] = (float)threadIdx.x / 255.0f;
    __syncthreads();
}

__device__ __forceinline__ RasterSetup loadSetup(const RasterSetup* setups, uint32_t t)
{
    RasterSetup s;
    const float4* src = reinterpret_cast<const float4*>(setups + t);
    float4* dst = reinterpret_cast<float4*>(&s);
#pragma unroll
    for (int k = 0; k < 6; ++k)
    {
        dst[k] = __ldg(src + k);
    }
    return s;
}

template<int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS) k_raster_small(RasterDev r, const uint32_t* __restrict__ smallList, const uint32_t* __restrict__ counters)
{
    __shared__ float unorm[256];
    fillUnormTable(unorm);
    const uint32_t count = counters[0];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    for (uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < count; w += warpsTotal)
    {
        const uint32_t tri = smallList[w];
        const RasterSetup s = loadSetup(r.setups, tri);
        const uint32_t x0 = s.bboxX & 0xFFFFu, x1 = s.bboxX >> 16, y0 = s.bboxY & 0xFFFFu, y1 = s.bboxY >> 16;
        {
            // hierarchical Z: hidden when, in every 16x16 tile the bounding box touches (at most 3 x 3), the triangle's nearest
            // depth lies behind the tile's published bound (the large triangles have been drawn before this kernel)
            uint32_t zmin24, zmax24;
            if (rasterDepthRange(s, x0, y0, x1, y1, zmin24, zmax24))
            {
                const uint32_t tx0 = x0 / kRasterTile, tx1 = x1 / kRasterTile, ty0 = y0 / kRasterTile, ty1 = y1 / kRasterTile;
                const uint32_t tw = tx1 - tx0 + 1u, tn = tw * (ty1 - ty0 + 1u);
                bool visible = false;
                for (uint32_t k = lane; k < tn; k += 32u)
                {
                    visible = visible || !(zmin24 > r.tileZ[(ty0 + k / tw) * r.tilesX16 + tx0 + k % tw]);
                }
                if (!__any_sync(0xFFFFFFFFu, visible))
                {
                    continue;
                }
            }
        }
        const uint32_t bw = x1 - x0 + 1u, n = bw * (y1 - y0 + 1u);
        for (uint32_t i = lane; i < n; i += 32u)
        {
            const uint32_t px = x0 + i % bw, py = y0 + i / bw;
            if (ownsPixel(r, px, py))
            {
                rasterPixel(s, tri, r.textures, r.mips, r.width, px, py, reinterpret_cast<uint64_t*>(r.visibility), AtomicMinKey(), unorm);
            }
        }
    }
}

// Implicit form: warp per bounding-box tile, found by binary search in the scanned tile counts.  Draws the frame only when
// the explicit work list of the binned form overflowed.
template<int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS)
k_raster_large(RasterDev r, const uint32_t* __restrict__ largeList, const uint32_t* __restrict__ largeBase, const uint32_t* __restrict__ counters)
{
    if (counters[3] <= r.itemCapacity)
    {
        return;
    }
    __shared__ float unorm[256];
    fillUnormTable(unorm);
    const uint32_t triCount = counters[1], items = counters[2];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    // a warp takes runs of kRun consecutive work items: one binary search per run, then a walk along the scanned counts
    // (consecutive items are neighbouring tiles of the same triangle; the search was 13 % of this kernel's stall samples)
    constexpr uint32_t kRun = 8;
    for (uint32_t base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * kRun; base < items; base += warpsTotal * kRun)
    {
      uint32_t lo = 0, hi = triCount; // largeBase[lo] <= base < largeBase[hi]
      while (hi - lo > 1u)
      {
          const uint32_t mid = (lo + hi) >> 1;
          if (largeBase[mid] <= base)
          {
              lo = mid;
          }
          else
          {
              hi = mid;
          }
      }
      for (uint32_t w = base; w < min(base + kRun, items); ++w)
      {
        while (lo + 1u < triCount && largeBase[lo + 1u] <= w)
        {
            ++lo;
        }
        const uint32_t tri = largeList[lo];
        const RasterSetup s = loadSetup(r.setups, tri);
        const uint32_t x0 = s.bboxX & 0xFFFFu, x1 = s.bboxX >> 16, y0 = s.bboxY & 0xFFFFu, y1 = s.bboxY >> 16;
        const uint32_t tilesX = x1 / kRasterTile - x0 / kRasterTile + 1u;
        const uint32_t local = w - largeBase[lo];
        const uint32_t tx = x0 / kRasterTile + local % tilesX, ty = y0 / kRasterTile + local / tilesX;
        const uint32_t px0 = max(tx * kRasterTile, x0), px1 = min(tx * kRasterTile + kRasterTile - 1u, x1);
        const uint32_t py0 = max(ty * kRasterTile, y0), py1 = min(ty * kRasterTile + kRasterTile - 1u, y1);
        if (!ownsPixel(r, px0, py0) || rasterTileOutside(s, px0, py0, px1, py1)) // a 16x16 tile lies inside one 64x64 ownership tile
        {
            continue;
        }
        {
            // hierarchical Z (see RasterDev::tileZ): skip when certainly hidden; publish a bound when this triangle is opaque
            // and covers every sample of the whole tile
            uint32_t* tz = r.tileZ + ty * r.tilesX16 + tx;
            uint32_t zmin24, zmax24;
            if (rasterDepthRange(s, px0, py0, px1, py1, zmin24, zmax24))
            {
                if (zmin24 > *reinterpret_cast<volatile uint32_t*>(tz))
                {
                    continue;
                }
            }
        }
        const uint32_t bw = px1 - px0 + 1u, n = bw * (py1 - py0 + 1u);
        for (uint32_t i = lane; i < n; i += 32u)
        {
            rasterPixel(s, tri, r.textures, r.mips, r.width, px0 + i % bw, py0 + i / bw, reinterpret_cast<uint64_t*>(r.visibility), AtomicMinKey(), unorm);
        }
      }
    }
}

// ------------------------------------------------------------------------------------------------------------
// Large triangles, binned form (default).  Three launches over the large-triangle list:
//   k_raster_tiles<true>   lanes over the (triangle, 16x16 bounding-box tile) pairs: opaque triangles that cover a whole
//                          tile publish the hierarchical-Z bound (nothing else is touched);
//   k_raster_tiles<false>  same walk: every tile that is owned, not outside an edge and not behind its tile's bound is
//                          appended to an explicit work list (warp-aggregated append);
//   k_raster_items         warp per listed (triangle, tile): lanes over the tile's pixels.
// Against the implicit form (k_raster_large: warp per bounding-box tile, found by binary search) no warp is spent on a
// tile that is rejected, and the bounds are complete before anything is drawn.  The list has a fixed capacity; a frame
// that would overflow it (counters[3] > capacity) is drawn by k_raster_large instead - both draw kernels are launched,
// one of them returns at once.
// ------------------------------------------------------------------------------------------------------------

template<bool PUBLISH>
__global__ void __launch_bounds__(256)
k_raster_tiles(RasterDev r, const uint32_t* __restrict__ largeList, const uint32_t* __restrict__ largeBase, uint32_t* __restrict__ counters,
               uint2* __restrict__ items)
{
    // LANES over the implicit (triangle, bounding-box tile) pairs, 32 consecutive ones per warp iteration: balanced whatever
    // the sizes of the triangles (a warp per triangle left the few screen-filling ones to single warps: 230 us per pass).
    // One binary search in the scanned tile counts per warp iteration; each lane then walks forward to its own triangle.
    const uint32_t triCount = counters[1], pairs = counters[2];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    for (uint32_t base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32u; base < pairs; base += warpsTotal * 32u)
    {
        uint32_t lo = 0, hi = triCount; // largeBase[lo] <= base < largeBase[hi]
        while (hi - lo > 1u)
        {
            const uint32_t mid = (lo + hi) >> 1;
            if (largeBase[mid] <= base)
            {
                lo = mid;
            }
            else
            {
                hi = mid;
            }
        }
        const uint32_t w = base + lane;
        bool keep = w < pairs;
        uint32_t tri = 0, tx = 0, ty = 0;
        if (keep)
        {
            while (lo + 1u < triCount && largeBase[lo + 1u] <= w)
            {
                ++lo;
            }
            tri = largeList[lo];
            const RasterSetup s = loadSetup(r.setups, tri);
            const uint32_t x0 = s.bboxX & 0xFFFFu, x1 = s.bboxX >> 16, y0 = s.bboxY & 0xFFFFu, y1 = s.bboxY >> 16;
            const uint32_t tw = x1 / kRasterTile - x0 / kRasterTile + 1u;
            const uint32_t local = w - largeBase[lo];
            tx = x0 / kRasterTile + local % tw, ty = y0 / kRasterTile + local / tw;
            const uint32_t px0 = max(tx * kRasterTile, x0), px1 = min(tx * kRasterTile + kRasterTile - 1u, x1);
            const uint32_t py0 = max(ty * kRasterTile, y0), py1 = min(ty * kRasterTile + kRasterTile - 1u, y1);
            keep = ownsPixel(r, px0, py0) && !rasterTileOutside(s, px0, py0, px1, py1);
            uint32_t zmin24, zmax24;
            if (keep && rasterDepthRange(s, px0, py0, px1, py1, zmin24, zmax24))
            {
                uint32_t* tz = r.tileZ + ty * r.tilesX16 + tx;
                if (PUBLISH)
                {
                    const uint32_t fx1 = min(tx * kRasterTile + kRasterTile - 1u, r.width - 1u), fy1 = min(ty * kRasterTile + kRasterTile - 1u, r.height - 1u);
                    if (px0 == tx * kRasterTile && py0 == ty * kRasterTile && px1 == fx1 && py1 == fy1 && zmax24 < 0xFFFFFFu &&
                        r.minAlpha[s.texture] >= 28u && rasterTileInside(s, px0, py0, px1, py1))
                    {
                        atomicMin(tz, zmax24);
                    }
                }
                else
                {
                    keep = !(zmin24 > *tz);
                }
            }
        }
        if (!PUBLISH)
        {
            const uint32_t ballot = __ballot_sync(0xFFFFFFFFu, keep);
            if (ballot)
            {
                uint32_t slot = 0;
                if (lane == 0)
                {
                    slot = atomicAdd(&counters[3], (uint32_t)__popc(ballot));
                }
                slot = __shfl_sync(0xFFFFFFFFu, slot, 0) + __popc(ballot & ((1u << lane) - 1u));
                if (keep && slot < r.itemCapacity)
                {
                    items[slot] = make_uint2(tri, tx | (ty << 16));
                }
            }
        }
    }
}

template<int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS)
k_raster_items(RasterDev r, const uint2* __restrict__ items, const uint32_t* __restrict__ counters)
{
    __shared__ float unorm[256];
    fillUnormTable(unorm);
    const uint32_t count = counters[3];
    if (count > r.itemCapacity)
    {
        return; // the work list overflowed: k_raster_large draws this frame
    }
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t warpsTotal = gridDim.x * (blockDim.x >> 5);
    for (uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < count; w += warpsTotal)
    {
        const uint2 item = __ldg(items + w);
        const uint32_t tri = item.x, tx = item.y & 0xFFFFu, ty = item.y >> 16;
        const RasterSetup s = loadSetup(r.setups, tri);
        const uint32_t x0 = s.bboxX & 0xFFFFu, x1 = s.bboxX >> 16, y0 = s.bboxY & 0xFFFFu, y1 = s.bboxY >> 16;
        const uint32_t px0 = max(tx * kRasterTile, x0), px1 = min(tx * kRasterTile + kRasterTile - 1u, x1);
        const uint32_t py0 = max(ty * kRasterTile, y0), py1 = min(ty * kRasterTile + kRasterTile - 1u, y1);
        const uint32_t bw = px1 - px0 + 1u, n = bw * (py1 - py0 + 1u);
        for (uint32_t i = lane; i < n; i += 32u)
        {
            rasterPixel(s, tri, r.textures, r.mips, r.width, px0 + i % bw, py0 + i / bw, reinterpret_cast<uint64_t*>(r.visibility), AtomicMinKey(), unorm);
        }
    }
}

__global__ void __launch_bounds__(256)
k_raster_resolve(RasterDev r, uint32_t* __restrict__ bgra8, uint32_t* __restrict__ tileBgra8, uint32_t* __restrict__ triangleId)
{
    __shared__ float unorm[256];
    fillUnormTable(unorm);
    if (r.size == 1u)
    {
        const uint32_t total = r.width * r.height;
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
        {
            const uint32_t px = i % r.width, py = i / r.width;
            uint32_t first;
            bgra8[i] = rasterResolve(r.setups, r.textures, r.mips, r.width, px, py, reinterpret_cast<const uint64_t*>(r.visibility), first, unorm);
            if (triangleId)
            {
                triangleId[i] = first;
            }
        }
        return;
    }
    // owned tiles only; the tile-major copy is the NCCL send buffer of vkrt_tiles_gather (row-major inside a tile)
    const uint32_t total = r.ownedTiles * (uint32_t)kTilePixels;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x)
    {
        const uint32_t tile = r.rank + (i / (uint32_t)kTilePixels) * r.size, pit = i % (uint32_t)kTilePixels;
        const uint32_t px = (tile % r.tilesX) * (uint32_t)kTile + pit % (uint32_t)kTile, py = (tile / r.tilesX) * (uint32_t)kTile + pit / (uint32_t)kTile;
        uint32_t value = 0u;
        if (px < r.width && py < r.height)
        {
            uint32_t first;
            value = rasterResolve(r.setups, r.textures, r.mips, r.width, px, py, reinterpret_cast<const uint64_t*>(r.visibility), first, unorm);
            bgra8[(size_t)py * r.width + px] = value;
            if (triangleId)
            {
                triangleId[(size_t)py * r.width + px] = first;
            }
        }
        tileBgra8[i] = value;
    }
}
} // namespace

int rasterizeFrame(Context* ctx, const Blas* blas, const float* wvp, uint32_t flags, Frame* frame)
{
    if (ctx->materialCount < blas->geometryCount || ctx->textureCount == 0)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_rasterize: materials (one row per geometry) and textures must be set first");
    }
    if (!blas->shadingLayout)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_rasterize: the vertex stage reads the 64-byte Model::Vertex (position and uv as vec4 "
                                                      "at offsets 0 and 32): the geometry needs vertexStride >= 64 and a multiple of 16");
    }
    if (int r = syncMaterials(ctx))
    {
        return r;
    }
    if (!ctx->mipsReady)
    {
        return ctx->fail(VKRT_ERROR_NOT_READY, "vkrt_rasterize: the fragment shader samples the mip chain; call vkrt_textures_generate_mips first");
    }
    if (frame->width > 32768u || frame->height > 32768u)
    {
        return ctx->fail(VKRT_ERROR_INVALID_ARGUMENT, "vkrt_rasterize: frame larger than 32768 pixels on a side");
    }
    const uint32_t n = blas->triangleCount, W = frame->width, H = frame->height;
    const size_t samples = (size_t)W * H * kRasterSamples;
    RasterBuffers& b = ctx->raster;
    VKRT_CUDA_TRY(ctx, frame->visibility.reserve(samples * 8));
    VKRT_CUDA_TRY(ctx, b.setups.reserve(std::max<size_t>(n, 1) * sizeof(RasterSetup)));
    VKRT_CUDA_TRY(ctx, b.smallList.reserve(std::max<size_t>(n, 1) * 4));
    VKRT_CUDA_TRY(ctx, b.largeList.reserve(std::max<size_t>(n, 1) * 4));
    VKRT_CUDA_TRY(ctx, b.largeTiles.reserve(std::max<size_t>(n, 1) * 4));
    VKRT_CUDA_TRY(ctx, b.counters.reserve(64));
    const uint32_t tilesX16 = (W + kRasterTile - 1u) / kRasterTile, tilesY16 = (H + kRasterTile - 1u) / kRasterTile;
    VKRT_CUDA_TRY(ctx, b.tileZ.reserve((size_t)tilesX16 * tilesY16 * 4));
    VKRT_CUDA_TRY(ctx, b.minAlpha.reserve((size_t)ctx->textureCount * 4));
    VKRT_CUDA_TRY(ctx, b.items.reserve((size_t)kRasterItemCapacity * sizeof(uint2)));
    // Optional (VKRT_RASTER_SORT=1): large triangles rasterised nearest first, so that later ones are mostly rejected by
    // the keys already in place.  The image does not depend on the order.  Measured on B200, Sponza stand-in at 4K:
    // 4.90 ms sorted against 4.68 ms unsorted - the kernel is bound by latency at low occupancy, not by its atomics -
    // so the default is off.
    static const bool sortLarge = getenv("VKRT_RASTER_SORT") && atoi(getenv("VKRT_RASTER_SORT")) != 0;
    if (sortLarge && n > 0)
    {
        VKRT_CUDA_TRY(ctx, b.sortKeys.reserve((size_t)n * 4));
        VKRT_CUDA_TRY(ctx, b.sortKeysB.reserve((size_t)n * 4));
        VKRT_CUDA_TRY(ctx, b.sortValsB.reserve((size_t)n * 4));
        VKRT_CUDA_TRY(ctx, b.sortState.reserve(sortPairsStateBytes(n)));
    }
    const bool aov = !(flags & VKRT_TRACE_NO_AOV);
    if (aov)
    {
        VKRT_CUDA_TRY(ctx, frame->triangleId.reserve((size_t)W * H * 4));
    }
    cudaStream_t s = ctx->stream;
    const uint32_t grid = (uint32_t)ctx->smCount * 8u;
    uint32_t* counters = b.counters.as<uint32_t>();

    Mat4 holder;
    std::copy(wvp, wvp + 16, holder.m);
    GeometryTable geo;
    geo.firstTri = blas->geomFirstTri.as<uint32_t>();
    geo.indexBase = blas->geomIndexBase.as<uint64_t>();
    geo.geometryCount = blas->geometryCount;
    RasterDev r;
    r.setups = b.setups.as<RasterSetup>();
    r.textures = ctx->textureRecords.as<TextureRecord>();
    r.mips = ctx->textureMips.as<TextureMips>();
    r.visibility = frame->visibility.as<unsigned long long>();
    r.width = W, r.height = H;
    r.tilesX = frame->tilesX, r.rank = (uint32_t)frame->rank, r.size = (uint32_t)frame->size, r.ownedTiles = frame->ownedTiles;
    r.tileZ = b.tileZ.as<uint32_t>(), r.tilesX16 = tilesX16, r.tilesY16 = tilesY16;
    r.minAlpha = b.minAlpha.as<uint32_t>();
    static const uint32_t itemCapacity = getenv("VKRT_RASTER_ITEMS") ? (uint32_t)std::min<long>(std::max<long>(atol(getenv("VKRT_RASTER_ITEMS")), 1), (long)kRasterItemCapacity) : kRasterItemCapacity;
    r.itemCapacity = itemCapacity;
    if (b.minAlphaGeneration != ctx->textureGeneration)
    {
        VKRT_CUDA_TRY(ctx, cudaMemsetAsync(b.minAlpha.ptr, 0xFF, (size_t)ctx->textureCount * 4, s));
        k_texture_min_alpha<<<dim3(32, ctx->textureCount), 256, 0, s>>>(r.textures, b.minAlpha.as<uint32_t>());
        ctx->launches += 1;
        b.minAlphaGeneration = ctx->textureGeneration;
    }

    k_raster_clear<<<grid, 256, 0, s>>>(frame->visibility.as<ulonglong2>(), samples / 2, counters, r);
    if (n > 0)
    {
        k_raster_setup<<<grid, 256, 0, s>>>(holder, geo, blas->vertices.as<uint8_t>(), blas->vertexStride, blas->indices.as<uint32_t>(),
                                            ctx->materials.as<VkrtSubmeshInfo>(), ctx->textureCount, n, W, H, b.setups.as<RasterSetup>(), b.smallList.as<uint32_t>(),
                                            b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), sortLarge ? b.sortKeys.as<uint32_t>() : nullptr, counters);
        if (sortLarge)
        {
            if (int e = sortPairsDevice(ctx, b.sortKeys.as<uint32_t>(), b.largeList.as<uint32_t>(), b.sortKeysB.as<uint32_t>(), b.sortValsB.as<uint32_t>(),
                                        b.sortState.as<uint32_t>(), n))
            {
                return e;
            }
            k_raster_tiles<<<grid, 256, 0, s>>>(b.setups.as<RasterSetup>(), b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters);
            ctx->launches += 1;
        }
        k_raster_scan<<<1, 1024, 0, s>>>(b.largeTiles.as<uint32_t>(), counters);
        // resident blocks per SM of the two triangle kernels: 4 (64 registers, 17 spilled words; default) or 3 (80 registers).
        // They are latency-bound (ncu on the first, 89-register build: long-scoreboard and wait stalls with 22 % of the warp
        // slots filled); Sponza stand-in at 4K on B200: 4.68 ms per frame at 2 blocks, 4.12 ms at 3, 3.87 ms at 4.
        static const int minBlocks = getenv("VKRT_RASTER_BLOCKS") ? atoi(getenv("VKRT_RASTER_BLOCKS")) : 4;
        if (minBlocks >= 4)
        {
            const uint32_t g4 = (uint32_t)ctx->smCount * 4u;
            // large triangles first: they publish the hierarchical-Z bounds that the small ones are tested against
            k_raster_tiles<true><<<g4, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters, b.items.as<uint2>());
            k_raster_tiles<false><<<g4, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters, b.items.as<uint2>());
            k_raster_items<4><<<g4, 256, 0, s>>>(r, b.items.as<uint2>(), counters);
            k_raster_large<4><<<g4, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters);
            k_raster_small<4><<<g4, 256, 0, s>>>(r, b.smallList.as<uint32_t>(), counters);
        }
        else
        {
            const uint32_t g3 = (uint32_t)ctx->smCount * 3u;
            k_raster_tiles<true><<<g3, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters, b.items.as<uint2>());
            k_raster_tiles<false><<<g3, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters, b.items.as<uint2>());
            k_raster_items<3><<<g3, 256, 0, s>>>(r, b.items.as<uint2>(), counters);
            k_raster_large<3><<<g3, 256, 0, s>>>(r, b.largeList.as<uint32_t>(), b.largeTiles.as<uint32_t>(), counters);
            k_raster_small<3><<<g3, 256, 0, s>>>(r, b.smallList.as<uint32_t>(), counters);
        }
        ctx->launches += 7;
    }
    if (int e = joinGather(ctx, frame))
    {
        return e;
    }
    if (int e = waitPresents(ctx, frame)) // the image is about to be overwritten
    {
        return e;
    }
    k_raster_resolve<<<grid, 256, 0, s>>>(r, frame->bgra8.as<uint32_t>(), frame->size > 1 ? frame->tileBgra8.as<uint32_t>() : nullptr,
                                          aov ? frame->triangleId.as<uint32_t>() : nullptr);
    ctx->launches += 2;
    frame->gathered = false;
    VKRT_CUDA_TRY(ctx, cudaGetLastError());
    return VKRT_SUCCESS;
}

} // namespace vkrt
