// shading.cuh — ray generation, closest-hit shading and miss logic, restated from the reference's GLSL:
//   shaders/shader.rgen:30-77, shaders/shader.rchit:78-172, shaders/shader.rmiss:15-18, shaders/shader_shadow.rmiss:6-9
// FP32, fixed left-to-right evaluation order, no FMA contraction (see device_math.cuh).
//
// The closest-hit shader is split around its shadow rays for the wavefront engine:
//   shadeSurface()  = rchit:88-108 (vertex fetch, interpolation, TBN, normal map)  -> SurfacePoint
//   lightTerm()     = rchit:121-128 per light (also decides whether a shadow ray is cast, rchit:129)
//   shadeResolve()  = rchit:130-171 given the 4 shadow results (light sum, base colour, reflection set-up)
#pragma once

#include "common.cuh"
#include "device_math.cuh"

namespace vkrt
{

struct SurfacePoint
{
    F3 worldPos;
    F3 normal; // perturbedNormal (may be NaN when the tangent is zero, see SURVEY §7 "NaN path")
    float u, v;
    uint32_t geometry;
};

// Halton radical inverse, digits accumulated in FP32 exactly as the oracle does.
VKRT_HD float radicalInverse(uint32_t s, uint32_t base)
{
    const float inv = 1.0f / (float)base;
    float f = inv, r = 0.0f;
    while (s)
    {
        r += (float)(s % base) * f;
        s /= base;
        f *= inv;
    }
    return r;
}

// In-pixel offset of sample s: spp==1 -> pixel centre (rgen:32); spp>1 -> Halton(2,3), 1-based index.
VKRT_HD void sampleOffset(uint32_t s, uint32_t spp, float& jx, float& jy)
{
    if (spp <= 1)
    {
        jx = 0.5f, jy = 0.5f;
    }
    else
    {
        jx = radicalInverse(s + 1, 2);
        jy = radicalInverse(s + 1, 3);
    }
}

// shader.rgen:32-38
VKRT_HD void primaryRay(const VkrtUniforms& u, uint32_t width, uint32_t height, uint32_t x, uint32_t y, float jx, float jy,
                        F3& origin, F3& direction)
{
    const float pcx = (float)x + jx, pcy = (float)y + jy;
    const float inx = pcx / (float)width, iny = pcy / (float)height;
    const float ux = inx * 2.0f - 1.0f, uy = iny * 2.0f - 1.0f;
    const F3 target = mat4MulVec(u.projInverse, ux, uy, 1.0f, 1.0f);
    const F3 tn = normalize3(target);
    direction = mat4MulVec(u.viewInverse, tn.x, tn.y, tn.z, 0.0f);
    origin = mat4MulVec(u.viewInverse, 0.0f, 0.0f, 0.0f, 1.0f);
}

VKRT_HD uint32_t wrapRepeat(float f, uint32_t n) // f integral; REPEAT addressing
{
    float m = fmodf(f, (float)n);
    if (m < 0.0f)
    {
        m += (float)n;
    }
    const uint32_t i = (uint32_t)m;
    return i >= n ? 0u : i;
}

VKRT_HD uint32_t loadTexel(const uint8_t* texels, size_t index)
{
#if defined(__CUDA_ARCH__)
    return __ldg(reinterpret_cast<const uint32_t*>(texels) + index);
#else
    uint32_t v;
    memcpy(&v, texels + index * 4, 4);
    return v;
#endif
}

// texture(sampler2D, uv) at LOD 0: LINEAR, REPEAT, RGBA8 UNORM (src/Raytracer.cpp:436-458,467).
VKRT_HD void sampleTexture(const TextureRecord& tex, float u, float v, float rgba[4])
{
    float x = u * (float)tex.width - 0.5f;
    float y = v * (float)tex.height - 0.5f;
    if (!finite1(x))
    {
        x = -0.5f;
    }
    if (!finite1(y))
    {
        y = -0.5f;
    }
    const float fx = floorf(x), fy = floorf(y);
    const float a = x - fx, b = y - fy;
    const uint32_t x0 = wrapRepeat(fx, tex.width), y0 = wrapRepeat(fy, tex.height);
    const uint32_t x1 = (x0 + 1 == tex.width) ? 0u : x0 + 1, y1 = (y0 + 1 == tex.height) ? 0u : y0 + 1;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    const uint32_t p00 = loadTexel(tex.texels, (size_t)y0 * tex.width + x0);
    const uint32_t p10 = loadTexel(tex.texels, (size_t)y0 * tex.width + x1);
    const uint32_t p01 = loadTexel(tex.texels, (size_t)y1 * tex.width + x0);
    const uint32_t p11 = loadTexel(tex.texels, (size_t)y1 * tex.width + x1);
    for (int ch = 0; ch < 4; ++ch)
    {
        const float t00 = (float)((p00 >> (8 * ch)) & 0xFFu) / 255.0f, t10 = (float)((p10 >> (8 * ch)) & 0xFFu) / 255.0f;
        const float t01 = (float)((p01 >> (8 * ch)) & 0xFFu) / 255.0f, t11 = (float)((p11 >> (8 * ch)) & 0xFFu) / 255.0f;
        rgba[ch] = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
    }
}

VKRT_HD float4 loadF4(const uint8_t* base, size_t byteOffset)
{
#if defined(__CUDA_ARCH__)
    return __ldg(reinterpret_cast<const float4*>(base + byteOffset));
#else
    float4 v;
    memcpy(&v, base + byteOffset, 16);
    return v;
#endif
}

// shader.rchit:88-108
VKRT_HD SurfacePoint shadeSurface(const InstanceRecord& inst, const VkrtSubmeshInfo* rows, const TextureRecord* textures,
                                  uint32_t geometry, uint32_t primitive, float b1, float b2)
{
    SurfacePoint sp;
    sp.geometry = geometry;
    const VkrtSubmeshInfo row = rows[geometry];
    const uint64_t tri = (uint64_t)(uint32_t)row.indexBufferOffset + primitive; // rchit:88-89
    const uint32_t i0 = inst.indices[3 * tri + 0], i1 = inst.indices[3 * tri + 1], i2 = inst.indices[3 * tri + 2];
    const uint8_t* vb = inst.vertices;
    const size_t s = inst.vertexStride;
    const float bx = 1.0f - b1 - b2, by = b1, bz = b2; // rchit:94

    const float4 uv0 = loadF4(vb, i0 * s + 32), uv1 = loadF4(vb, i1 * s + 32), uv2 = loadF4(vb, i2 * s + 32);
    sp.u = (uv0.x * bx + uv1.x * by) + uv2.x * bz; // rchit:95
    sp.v = (uv0.y * bx + uv1.y * by) + uv2.y * bz;

    const float4 p0 = loadF4(vb, i0 * s), p1 = loadF4(vb, i1 * s), p2 = loadF4(vb, i2 * s);
    const F3 position = (f3(p0.x, p0.y, p0.z) * bx + f3(p1.x, p1.y, p1.z) * by) + f3(p2.x, p2.y, p2.z) * bz; // rchit:97
    sp.worldPos = affinePoint(inst.o2w, position);                                                         // rchit:98

    const float4 n0 = loadF4(vb, i0 * s + 16), n1 = loadF4(vb, i1 * s + 16), n2 = loadF4(vb, i2 * s + 16);
    const F3 normal = (f3(n0.x, n0.y, n0.z) * bx + f3(n1.x, n1.y, n1.z) * by) + f3(n2.x, n2.y, n2.z) * bz; // rchit:100
    const float* w = inst.w2o; // normal * gl_WorldToObjectEXT (row vector times the 3x3 part), rchit:101
    const F3 nW = f3((normal.x * w[0] + normal.y * w[4]) + normal.z * w[8], (normal.x * w[1] + normal.y * w[5]) + normal.z * w[9],
                     (normal.x * w[2] + normal.y * w[6]) + normal.z * w[10]);
    const F3 worldNormal = normalize3(nW);

    const float4 t0 = loadF4(vb, i0 * s + 48), t1 = loadF4(vb, i1 * s + 48), t2 = loadF4(vb, i2 * s + 48);
    const F3 tangent = (f3(t0.x, t0.y, t0.z) * bx + f3(t1.x, t1.y, t1.z) * by) + f3(t2.x, t2.y, t2.z) * bz; // rchit:103

    // getTBN, rchit:78-84 (M = identity)
    const F3 N = worldNormal;
    const F3 T = normalize3(tangent);
    const F3 B = cross3(T, N);
    float texel[4];
    sampleTexture(textures[(uint32_t)row.normalTextureIndex], sp.u, sp.v, texel); // rchit:106-107
    const F3 mapN = f3(texel[0] * 2.0f - 1.0f, texel[1] * 2.0f - 1.0f, texel[2] * 2.0f - 1.0f);
    const F3 mn = normalize3(mapN);
    sp.normal = normalize3((T * mn.x + B * mn.y) + N * mn.z); // rchit:108
    return sp;
}

struct LightTerm
{
    F3 dir;         // lightDir
    float distance; // lightDistance (shadow ray tmax)
    float weight;   // diffuse * lightPower
    bool castShadow;
};

// shader.rchit:121-129 for light i
VKRT_HD LightTerm lightTerm(const VkrtUniforms& u, int i, F3 worldPos, F3 normal)
{
    LightTerm lt;
    const F3 lightVec = f3(u.lightPositions[i][0], u.lightPositions[i][1], u.lightPositions[i][2]) - worldPos;
    lt.distance = length3(lightVec);
    lt.dir = lightVec / lt.distance;
    const float ndl = dot3(normal, lt.dir);
    const float diffuse = clamp01(ndl);
    const float lightPower = 10.0f / (lt.distance * lt.distance);
    lt.weight = diffuse * lightPower;
    lt.castShadow = ndl > 0.0f;
    return lt;
}

struct ResolveOut
{
    F3 hitValue;
    float attenuation; // updated payload.attenuation
    bool reflect;      // payload.done == 0
    F3 rayDir;         // reflection direction (origin = worldPos)
};

// shader.rchit:130-171. shadowBits: bit i set = light i occluded.
VKRT_HD ResolveOut shadeResolve(const VkrtUniforms& u, const VkrtSubmeshInfo* rows, const TextureRecord* textures,
                                const SurfacePoint& sp, uint32_t shadowBits, float attenuation, F3 worldRayDir)
{
    ResolveOut out;
    float totalLightAmount = 0.0f;
    for (int i = 0; i < 4; ++i)
    {
        const LightTerm lt = lightTerm(u, i, sp.worldPos, sp.normal);
        const float shadowMultiplier = (lt.castShadow && ((shadowBits >> i) & 1u)) ? 0.3f : 1.0f;
        totalLightAmount += lt.weight * shadowMultiplier;
    }
    const VkrtSubmeshInfo row = rows[sp.geometry];
    const float ambient = 0.1f;
    float texel[4];
    sampleTexture(textures[(uint32_t)row.baseColorTextureIndex], sp.u, sp.v, texel); // rchit:156-157
    const F3 baseColor = f3(texel[0], texel[1], texel[2]);
    out.hitValue = (baseColor * totalLightAmount) * attenuation + baseColor * ambient; // rchit:158
    sampleTexture(textures[(uint32_t)row.metallicRoughnessTextureIndex], sp.u, sp.v, texel); // rchit:161-162
    const float metallic = texel[2];
    out.attenuation = attenuation;
    out.reflect = false;
    out.rayDir = f3(0.0f, 0.0f, 0.0f);
    if (metallic > 0.1f) // rchit:163-171
    {
        const float reflectAmount = 0.5f * metallic;
        out.attenuation = attenuation * reflectAmount;
        out.hitValue = out.hitValue * (1.0f - out.attenuation);
        out.reflect = true;
        const float k = 2.0f * dot3(sp.normal, worldRayDir); // reflect(I,N) = I - 2*dot(N,I)*N
        out.rayDir = worldRayDir - sp.normal * k;
    }
    return out;
}

// UNORM8 store of the colour attachment (VK_FORMAT_B8G8R8A8_UNORM, src/VulkanUtils.hpp:24)
VKRT_HD uint32_t packBgra8(F3 c)
{
    const uint32_t r = (uint32_t)(clamp01(c.x) * 255.0f + 0.5f);
    const uint32_t g = (uint32_t)(clamp01(c.y) * 255.0f + 0.5f);
    const uint32_t b = (uint32_t)(clamp01(c.z) * 255.0f + 0.5f);
    return b | (g << 8) | (r << 16) | (255u << 24);
}

// Interleaved tile ownership: pixel path index -> (x, y).  Paths are laid out tile-major; inside a 64x64 tile
// the 4096 pixels are ordered as 8x4-pixel blocks so one warp covers a compact screen patch.
struct TileMap
{
    uint32_t width, height;
    uint32_t tilesX, tilesY;
    uint32_t rank, size; // this context owns tiles with (tile % size) == rank
    uint32_t ownedTiles;
};

VKRT_HD bool pathToPixel(const TileMap& m, uint32_t path, uint32_t& x, uint32_t& y, uint32_t& tileLocal, uint32_t& pxInTile)
{
    tileLocal = path / (uint32_t)kTilePixels;
    const uint32_t within = path % (uint32_t)kTilePixels;
    const uint32_t block = within >> 5, lane = within & 31u;
    const uint32_t bx = block & 7u, by = block >> 3; // 8 blocks across, 16 down
    // pixels of a block along a Morton curve: 4 consecutive paths are a 2x2 quad, 8 a 4x2 patch (with several samples
    // per pixel in adjacent lanes a warp then covers a compact patch rather than a row)
    const uint32_t lx = (lane & 1u) | ((lane >> 1) & 2u) | ((lane >> 2) & 4u), ly = ((lane >> 1) & 1u) | ((lane >> 2) & 2u);
    const uint32_t px = bx * 8u + lx, py = by * 4u + ly;
    const uint32_t tile = m.rank + tileLocal * m.size;
    const uint32_t tx = tile % m.tilesX, ty = tile / m.tilesX;
    x = tx * (uint32_t)kTile + px;
    y = ty * (uint32_t)kTile + py;
    pxInTile = py * (uint32_t)kTile + px;
    return tileLocal < m.ownedTiles && x < m.width && y < m.height;
}

} // namespace vkrt
