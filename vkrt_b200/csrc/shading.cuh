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
    float footprint2; // ray-cone mode only: squared uv-space footprint of the ray's cone (coneFootprint2)
};

// Ray-cone level-of-detail inputs of one hit (VKRT_TEXTURE_LOD_RAY_CONES); unused in the default LOD-0 mode.
struct LodInput
{
    const TextureMips* mips; // parallel to the texture array
    F3 worldRayDir;
    float coneWidth; // spread angle times the distance travelled along the path
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

// Pixel spread angle of the primary rays for the ray cones: 2 tan(fovy/2) / height (small-angle form of RTG ch. 20
// eq. 30); tan(fovy/2) is |projInverse[1][1]|.
VKRT_HD float pixelSpreadAngle(const VkrtUniforms& u, uint32_t height)
{
    return (2.0f * fabsf(u.projInverse[5])) / (float)height;
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

// Bilinear REPEAT sample of one RGBA8 image level (LINEAR mag/min, src/Raytracer.cpp:436-458,467); Vulkan texel
// addressing u*W - 0.5.  `unorm` (optional) is a 256-entry table of c / 255.0f: the identical IEEE quotients, looked up
// (shared memory) instead of divided - the rasteriser kernels, which filter a texture per fragment, pass one.
VKRT_HD void sampleLevel(const uint8_t* texels, uint32_t width, uint32_t height, float u, float v, float rgba[4], const float* unorm = nullptr)
{
    float x = u * (float)width - 0.5f;
    float y = v * (float)height - 0.5f;
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
    const uint32_t x0 = wrapRepeat(fx, width), y0 = wrapRepeat(fy, height);
    const uint32_t x1 = (x0 + 1 == width) ? 0u : x0 + 1, y1 = (y0 + 1 == height) ? 0u : y0 + 1;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    const uint32_t p00 = loadTexel(texels, (size_t)y0 * width + x0);
    const uint32_t p10 = loadTexel(texels, (size_t)y0 * width + x1);
    const uint32_t p01 = loadTexel(texels, (size_t)y1 * width + x0);
    const uint32_t p11 = loadTexel(texels, (size_t)y1 * width + x1);
    for (int ch = 0; ch < 4; ++ch)
    {
        if (unorm)
        {
            const float t00 = unorm[(p00 >> (8 * ch)) & 0xFFu], t10 = unorm[(p10 >> (8 * ch)) & 0xFFu];
            const float t01 = unorm[(p01 >> (8 * ch)) & 0xFFu], t11 = unorm[(p11 >> (8 * ch)) & 0xFFu];
            rgba[ch] = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
        }
        else
        {
            const float t00 = (float)((p00 >> (8 * ch)) & 0xFFu) / 255.0f, t10 = (float)((p10 >> (8 * ch)) & 0xFFu) / 255.0f;
            const float t01 = (float)((p01 >> (8 * ch)) & 0xFFu) / 255.0f, t11 = (float)((p11 >> (8 * ch)) & 0xFFu) / 255.0f;
            rgba[ch] = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
        }
    }
}

// texture(sampler2D, uv) at LOD 0 (what the reference's ray-tracing shaders get: no derivatives).
VKRT_HD void sampleTexture(const TextureRecord& tex, float u, float v, float rgba[4], const float* unorm = nullptr)
{
    sampleLevel(tex.texels, tex.width, tex.height, u, v, rgba, unorm);
}

// ---- Mip chain and ray-cone level of detail (off by default, the parity frames sample LOD 0) ----------
// Level k of a WxH image is max(W>>k,1) x max(H>>k,1) (src/Raytracer.cpp:467-468,612,629-630); levels 1.. follow each
// other tightly packed in TextureMips::texels.
VKRT_HD uint32_t mipExtent(uint32_t extent, uint32_t level)
{
    const uint32_t e = extent >> level;
    return e ? e : 1u;
}

VKRT_HD uint32_t mipLevelCount(uint32_t width, uint32_t height) // floor(log2(max(w,h))) + 1, src/Raytracer.cpp:467
{
    uint32_t m = width > height ? width : height, levels = 0;
    while (m)
    {
        ++levels;
        m >>= 1;
    }
    return levels ? levels : 1u;
}

// One destination texel of vkCmdBlitImage(level-1 -> level, VK_FILTER_LINEAR) (src/Raytracer.cpp:610-625): the source
// coordinate of destination texel centre i is (i + 0.5) * srcExtent / dstExtent, filtered bilinearly with
// clamp-to-edge; for an even source extent that is the mean of a 2x2 block.  UNORM8 store as packBgra8 rounds.
// `unorm` (optional) is a 256-entry table of c / 255.0f: the same IEEE quotients, looked up instead of divided.
VKRT_HD uint32_t blitTexel(const uint8_t* src, uint32_t srcW, uint32_t srcH, uint32_t dstW, uint32_t dstH, uint32_t i, uint32_t j,
                           const float* unorm = nullptr)
{
    const float sx = ((float)i + 0.5f) * ((float)srcW / (float)dstW) - 0.5f;
    const float sy = ((float)j + 0.5f) * ((float)srcH / (float)dstH) - 0.5f;
    const float fx = floorf(sx), fy = floorf(sy);
    const float a = sx - fx, b = sy - fy;
    const int ix = (int)fx, iy = (int)fy;
    const uint32_t x0 = (uint32_t)(ix < 0 ? 0 : (ix >= (int)srcW ? (int)srcW - 1 : ix));
    const uint32_t y0 = (uint32_t)(iy < 0 ? 0 : (iy >= (int)srcH ? (int)srcH - 1 : iy));
    const uint32_t x1 = (ix + 1 >= (int)srcW) ? srcW - 1 : (uint32_t)(ix + 1 < 0 ? 0 : ix + 1);
    const uint32_t y1 = (iy + 1 >= (int)srcH) ? srcH - 1 : (uint32_t)(iy + 1 < 0 ? 0 : iy + 1);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    const uint32_t p00 = loadTexel(src, (size_t)y0 * srcW + x0), p10 = loadTexel(src, (size_t)y0 * srcW + x1);
    const uint32_t p01 = loadTexel(src, (size_t)y1 * srcW + x0), p11 = loadTexel(src, (size_t)y1 * srcW + x1);
    uint32_t out = 0;
    for (int ch = 0; ch < 4; ++ch)
    {
        const uint32_t c00 = (p00 >> (8 * ch)) & 0xFFu, c10 = (p10 >> (8 * ch)) & 0xFFu, c01 = (p01 >> (8 * ch)) & 0xFFu, c11 = (p11 >> (8 * ch)) & 0xFFu;
        const float t00 = unorm ? unorm[c00] : (float)c00 / 255.0f, t10 = unorm ? unorm[c10] : (float)c10 / 255.0f;
        const float t01 = unorm ? unorm[c01] : (float)c01 / 255.0f, t11 = unorm ? unorm[c11] : (float)c11 / 255.0f;
        const float c = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
        out |= (uint32_t)(clamp01(c) * 255.0f + 0.5f) << (8 * ch);
    }
    return out;
}

// log2 of a finite x > 0 from its exponent and a cubic-corrected mantissa term (|error| < 1.5e-4): only + and *, fixed
// order, so the CPU oracle and the device agree bit for bit (libm's log2f and CUDA's do not).
VKRT_HD float lodLog2(float x)
{
    uint32_t bits;
    memcpy(&bits, &x, 4);
    const float e = (float)((int)(bits >> 23) - 127);
    const uint32_t mb = (bits & 0x007FFFFFu) | 0x3F800000u;
    float m;
    memcpy(&m, &mb, 4);
    const float f = m - 1.0f;
    const float poly = 0.4380733f + f * (-0.2366934f + f * 0.0803073f);
    return e + (f + (f * (1.0f - f)) * poly);
}

// Ray-cone footprint (Akenine-Moller et al., "Texture Level of Detail Strategies for Real-Time Ray Tracing", Ray Tracing
// Gems 2019, ch. 20): squared uv-space footprint of a cone of width `coneWidth` on a triangle =
// (uv double area / world double area) * (coneWidth / |n.d|)^2.  A texture of WxH texels then takes
// lambda = 0.5 * log2(footprint2 * W * H).
VKRT_HD float coneFootprint2(F3 p0, F3 p1, F3 p2, float u0, float v0, float u1, float v1, float u2, float v2, F3 rayDir, float coneWidth)
{
    const F3 c = cross3(p1 - p0, p2 - p0);
    const float worldArea2 = length3(c);
    const float uvArea2 = fabsf((u1 - u0) * (v2 - v0) - (u2 - u0) * (v1 - v0));
    const float ndotd = fabsf(dot3(c, rayDir)) / (worldArea2 * length3(rayDir));
    const float wn = coneWidth / ndotd;
    return (uvArea2 / worldArea2) * (wn * wn);
}

// Trilinear sample (mip LINEAR, minLod 0, no maxLod: src/Raytracer.cpp:436-458, src/Rasterizer.cpp:484-498) for a squared
// footprint of `texels2` level-0 texels: lambda = 0.5 * log2(texels2), clamped to the chain.  A footprint that is not
// > 1 texel (magnification, degenerate triangles or uv, NaN, inf) samples level 0 exactly as sampleTexture does.
VKRT_HD void sampleTrilinear(const TextureRecord& tex, const TextureMips& mips, float u, float v, float texels2, float rgba[4],
                             const float* unorm = nullptr)
{
    if (!(texels2 > 1.0f) || !finite1(texels2) || mips.levels <= 1u)
    {
        sampleLevel(tex.texels, tex.width, tex.height, u, v, rgba, unorm);
        return;
    }
    float lambda = 0.5f * lodLog2(texels2);
    const float top = (float)(mips.levels - 1u);
    if (lambda > top)
    {
        lambda = top;
    }
    const float fl = floorf(lambda);
    const float frac = lambda - fl;
    const uint32_t l0 = (uint32_t)fl;
    // byte offset of level l0 inside mips.texels (levels 1.. packed)
    size_t offset = 0;
    for (uint32_t k = 1; k < l0; ++k)
    {
        offset += (size_t)mipExtent(tex.width, k) * mipExtent(tex.height, k) * 4;
    }
    const uint8_t* base0 = l0 == 0 ? tex.texels : mips.texels + offset;
    const uint32_t w0 = mipExtent(tex.width, l0), h0 = mipExtent(tex.height, l0);
    sampleLevel(base0, w0, h0, u, v, rgba, unorm);
    if (frac > 0.0f) // l0 < levels - 1 here, because lambda <= levels - 1
    {
        const uint8_t* base1 = l0 == 0 ? mips.texels : base0 + (size_t)w0 * h0 * 4;
        float hi[4];
        sampleLevel(base1, mipExtent(tex.width, l0 + 1), mipExtent(tex.height, l0 + 1), u, v, hi, unorm);
        for (int ch = 0; ch < 4; ++ch)
        {
            rgba[ch] = rgba[ch] * (1.0f - frac) + hi[ch] * frac;
        }
    }
}

// Ray-cone level of detail: the squared uv footprint covers footprint2 * W * H texels of this texture.
VKRT_HD void sampleTextureLod(const TextureRecord& tex, const TextureMips& mips, float u, float v, float footprint2, float rgba[4],
                              const float* unorm = nullptr)
{
    sampleTrilinear(tex, mips, u, v, (footprint2 * (float)tex.width) * (float)tex.height, rgba, unorm);
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

// shader.rchit:88-108.  LOD = false is the reference's behaviour (every texture() call samples level 0).
template<bool LOD = false>
VKRT_HD SurfacePoint shadeSurface(const InstanceRecord& inst, const VkrtSubmeshInfo* rows, const TextureRecord* textures,
                                  uint32_t geometry, uint32_t primitive, float b1, float b2, const LodInput* lod = nullptr,
                                  const float* unorm = nullptr)
{
    SurfacePoint sp;
    sp.geometry = geometry;
    sp.footprint2 = 0.0f;
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
    if (LOD)
    {
        sp.footprint2 = coneFootprint2(affinePoint(inst.o2w, f3(p0.x, p0.y, p0.z)), affinePoint(inst.o2w, f3(p1.x, p1.y, p1.z)),
                                       affinePoint(inst.o2w, f3(p2.x, p2.y, p2.z)), uv0.x, uv0.y, uv1.x, uv1.y, uv2.x, uv2.y,
                                       lod->worldRayDir, lod->coneWidth);
        const uint32_t ti = (uint32_t)row.normalTextureIndex;
        sampleTextureLod(textures[ti], lod->mips[ti], sp.u, sp.v, sp.footprint2, texel, unorm);
    }
    else
    {
        sampleTexture(textures[(uint32_t)row.normalTextureIndex], sp.u, sp.v, texel, unorm); // rchit:106-107
    }
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
template<bool LOD = false>
VKRT_HD ResolveOut shadeResolve(const VkrtUniforms& u, const VkrtSubmeshInfo* rows, const TextureRecord* textures,
                                const SurfacePoint& sp, uint32_t shadowBits, float attenuation, F3 worldRayDir,
                                const TextureMips* mips = nullptr, const float* unorm = nullptr)
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
    if (LOD)
    {
        const uint32_t ti = (uint32_t)row.baseColorTextureIndex;
        sampleTextureLod(textures[ti], mips[ti], sp.u, sp.v, sp.footprint2, texel, unorm);
    }
    else
    {
        sampleTexture(textures[(uint32_t)row.baseColorTextureIndex], sp.u, sp.v, texel, unorm); // rchit:156-157
    }
    const F3 baseColor = f3(texel[0], texel[1], texel[2]);
    out.hitValue = (baseColor * totalLightAmount) * attenuation + baseColor * ambient; // rchit:158
    if (LOD)
    {
        const uint32_t ti = (uint32_t)row.metallicRoughnessTextureIndex;
        sampleTextureLod(textures[ti], mips[ti], sp.u, sp.v, sp.footprint2, texel, unorm);
    }
    else
    {
        sampleTexture(textures[(uint32_t)row.metallicRoughnessTextureIndex], sp.u, sp.v, texel, unorm); // rchit:161-162
    }
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
