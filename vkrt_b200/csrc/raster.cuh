// raster.cuh — the reference's OTHER renderer (`Rasterizer`, disabled in src/main.cpp:8) as a software rasteriser:
//   shaders/shader.vert:18-23     gl_Position = wvp * (position, 1); uv passed through
//   shaders/shader.frag:13-22     colour = texture(baseColor, uv); discard when alpha < 0.1
//   src/Rasterizer.cpp:779-829    viewport (0,0,W,H) depth [0,1]; back-face culling, counter-clockwise front faces; 8x MSAA
//                                 without sample shading; depth test LESS with writes; no blending
//   src/Rasterizer.cpp:119-120    clear colour (0,0,0.2,1), clear depth 1;  src/VulkanUtils.hpp:25 D24 depth
//   src/Rasterizer.cpp:484-498    sampler: LINEAR / REPEAT / mip LINEAR, all levels (implicit derivatives: the mip chain IS used here)
//
// Rasterisation is done in homogeneous coordinates (Olano & Greer, "Triangle Scan Conversion using 2D Homogeneous
// Coordinates", 1997): the three edge functions are the rows of the adjoint of the matrix of pixel-space homogeneous
// vertices, so triangles that cross the near plane need no clipping, and all attributes are perspective-correct ratios
// of planes.  An edge shared by two triangles gets exactly negated coefficients (cross3(a,b) == -cross3(b,a) bit for
// bit), which together with the top-left rule makes the coverage watertight without a fixed-point snap.
// Visibility is resolved through a 64-bit key per sample, (24-bit depth << 32) | triangle index in draw order, kept with
// atomicMin: the smallest depth wins and equal depths go to the earlier triangle, which is what depth test LESS does
// when triangles are drawn in order - so the result does not depend on the order in which threads get there.
// FP32, fixed evaluation order, no FMA contraction (see device_math.cuh): bit-comparable with the CPU oracle.
#pragma once

#include "shading.cuh"

namespace vkrt
{

constexpr int kRasterSamples = 8;        // c_msaaSampleCount, src/Rasterizer.cpp:17
constexpr int kRasterTile = 16;          // work-item granularity of the large-triangle kernel (pixels)
constexpr uint64_t kRasterEmpty = ~0ull; // cleared sample: depth 1.0 never passes LESS against it, no triangle

// Vulkan standard sample locations for VK_SAMPLE_COUNT_8_BIT (spec table "Standard sample locations"), in 1/16 pixel.
// x = {9, 7, 13, 5, 3, 1, 11, 15} / 16, y = {5, 11, 9, 3, 13, 7, 15, 1} / 16, one nibble per sample.
VKRT_HD float rasterSampleX(int s)
{
    return (float)((0xFB135D79u >> (4 * s)) & 15u) * 0.0625f;
}
VKRT_HD float rasterSampleY(int s)
{
    return (float)((0x1F7D39B5u >> (4 * s)) & 15u) * 0.0625f;
}

// Per-triangle planes, 96 bytes.  Everything is a plane p(x,y) = (a*x + b*y) + c over framebuffer coordinates.
struct alignas(16) RasterSetup
{
    float ea[3], eb[3], ec[3]; // edge function i (opposite vertex i), >= 0 inside
    float za, zb, zc;          // depth z_clip / w_clip
    float ua, ub, uc;          // numerator of u
    float va, vb, vc;          // numerator of v
    float sa, sb, sc;          // common denominator (sum of the edge functions = D / w)
    uint32_t texture;          // base colour image of the submesh's material
    uint32_t bboxX, bboxY;     // first | last << 16, inclusive pixel range; bboxX == kInvalid: not drawn
};
static_assert(sizeof(RasterSetup) == 96, "RasterSetup must be 96 bytes");

struct Clip4
{
    float x, y, z, w;
};

VKRT_HD Clip4 rasterVertex(const float* wvp, F3 p) // shader.vert:20, column-major mat4 * (p, 1)
{
    Clip4 r;
    r.x = ((wvp[0] * p.x + wvp[4] * p.y) + wvp[8] * p.z) + wvp[12];
    r.y = ((wvp[1] * p.x + wvp[5] * p.y) + wvp[9] * p.z) + wvp[13];
    r.z = ((wvp[2] * p.x + wvp[6] * p.y) + wvp[10] * p.z) + wvp[14];
    r.w = ((wvp[3] * p.x + wvp[7] * p.y) + wvp[11] * p.z) + wvp[15];
    return r;
}

VKRT_HD float plane3(float a, float b, float c, float x, float y)
{
    return (a * x + b * y) + c;
}

// Triangle set-up.  Returns false (and marks the record) when nothing of the triangle can be drawn: outside one frustum
// plane, back-facing or degenerate (front faces are counter-clockwise, VK_CULL_MODE_BACK_BIT), or off screen.
VKRT_HD bool rasterSetup(const float* wvp, F3 p0, F3 p1, F3 p2, float u0, float v0, float u1, float v1, float u2, float v2, uint32_t texture,
                         uint32_t width, uint32_t height, RasterSetup& out)
{
    out.bboxX = kInvalid, out.bboxY = 0u, out.texture = texture;
    const Clip4 c0 = rasterVertex(wvp, p0), c1 = rasterVertex(wvp, p1), c2 = rasterVertex(wvp, p2);
    if ((c0.x < -c0.w && c1.x < -c1.w && c2.x < -c2.w) || (c0.x > c0.w && c1.x > c1.w && c2.x > c2.w) ||
        (c0.y < -c0.w && c1.y < -c1.w && c2.y < -c2.w) || (c0.y > c0.w && c1.y > c1.w && c2.y > c2.w) ||
        (c0.z < 0.0f && c1.z < 0.0f && c2.z < 0.0f) || (c0.z > c0.w && c1.z > c1.w && c2.z > c2.w))
    {
        return false;
    }
    // homogeneous framebuffer coordinates: (x_fb * w, y_fb * w, w)
    const float hx = 0.5f * (float)width, hy = 0.5f * (float)height;
    const F3 V0 = f3((c0.x + c0.w) * hx, (c0.y + c0.w) * hy, c0.w);
    const F3 V1 = f3((c1.x + c1.w) * hx, (c1.y + c1.w) * hy, c1.w);
    const F3 V2 = f3((c2.x + c2.w) * hx, (c2.y + c2.w) * hy, c2.w);
    const F3 n0 = cross3(V2, V1), n1 = cross3(V0, V2), n2 = cross3(V1, V0);
    const float D = dot3(V0, n0); // > 0 for counter-clockwise triangles in Vulkan's y-down framebuffer
    if (!(D > 0.0f) || !finite1(D))
    {
        return false;
    }
    out.ea[0] = n0.x, out.eb[0] = n0.y, out.ec[0] = n0.z;
    out.ea[1] = n1.x, out.eb[1] = n1.y, out.ec[1] = n1.z;
    out.ea[2] = n2.x, out.eb[2] = n2.y, out.ec[2] = n2.z;
    out.za = ((c0.z * n0.x + c1.z * n1.x) + c2.z * n2.x) / D;
    out.zb = ((c0.z * n0.y + c1.z * n1.y) + c2.z * n2.y) / D;
    out.zc = ((c0.z * n0.z + c1.z * n1.z) + c2.z * n2.z) / D;
    out.ua = (u0 * n0.x + u1 * n1.x) + u2 * n2.x, out.ub = (u0 * n0.y + u1 * n1.y) + u2 * n2.y, out.uc = (u0 * n0.z + u1 * n1.z) + u2 * n2.z;
    out.va = (v0 * n0.x + v1 * n1.x) + v2 * n2.x, out.vb = (v0 * n0.y + v1 * n1.y) + v2 * n2.y, out.vc = (v0 * n0.z + v1 * n1.z) + v2 * n2.z;
    out.sa = (n0.x + n1.x) + n2.x, out.sb = (n0.y + n1.y) + n2.y, out.sc = (n0.z + n1.z) + n2.z;
    // pixel bounding box; a triangle with a vertex behind the eye takes the whole screen
    float x0 = 0.0f, y0 = 0.0f, x1 = (float)(width - 1u), y1 = (float)(height - 1u);
    if (c0.w > 0.0f && c1.w > 0.0f && c2.w > 0.0f)
    {
        const float ax = V0.x / V0.z, bx = V1.x / V1.z, cx = V2.x / V2.z;
        const float ay = V0.y / V0.z, by = V1.y / V1.z, cy = V2.y / V2.z;
        const float lox = fminf(fminf(ax, bx), cx), hix = fmaxf(fmaxf(ax, bx), cx);
        const float loy = fminf(fminf(ay, by), cy), hiy = fmaxf(fmaxf(ay, by), cy);
        if (!(hix >= 0.0f) || !(lox < (float)width) || !(hiy >= 0.0f) || !(loy < (float)height))
        {
            return false;
        }
        x0 = fmaxf(floorf(lox), 0.0f), y0 = fmaxf(floorf(loy), 0.0f);
        x1 = fminf(floorf(hix), x1), y1 = fminf(floorf(hiy), y1);
    }
    out.bboxX = (uint32_t)x0 | ((uint32_t)x1 << 16);
    out.bboxY = (uint32_t)y0 | ((uint32_t)y1 << 16);
    return true;
}

// Coverage of the 8 samples of pixel (px, py): inside all three edges, ties by the top-left rule (an edge whose
// interior lies to its right, or a horizontal edge whose interior lies below it, owns its samples).
VKRT_HD uint32_t rasterCoverage(const RasterSetup& t, uint32_t px, uint32_t py)
{
    uint32_t mask = 0;
    for (int s = 0; s < kRasterSamples; ++s)
    {
        const float x = (float)px + rasterSampleX(s), y = (float)py + rasterSampleY(s);
        bool inside = true;
        for (int i = 0; i < 3; ++i)
        {
            const float e = plane3(t.ea[i], t.eb[i], t.ec[i], x, y);
            const bool owns = t.ea[i] > 0.0f || (t.ea[i] == 0.0f && t.eb[i] > 0.0f);
            inside = inside && (e > 0.0f || (e == 0.0f && owns));
        }
        mask |= inside ? 1u << s : 0u;
    }
    return mask;
}

// 24-bit depth of sample s (VK_FORMAT_D24_UNORM_S8_UINT); false when the sample is clipped by the near or far plane or
// cannot pass the depth test against the cleared buffer.
VKRT_HD bool rasterDepth(const RasterSetup& t, uint32_t px, uint32_t py, int s, uint32_t& depth24)
{
    const float z = plane3(t.za, t.zb, t.zc, (float)px + rasterSampleX(s), (float)py + rasterSampleY(s));
    if (!(z >= 0.0f && z <= 1.0f))
    {
        return false;
    }
    depth24 = (uint32_t)(z * 16777215.0f + 0.5f);
    return depth24 < 0xFFFFFFu; // LESS against the clear value 1.0 = 0xFFFFFF
}

// shader.frag:15-21 for the fragment of pixel (px, py): attributes at the pixel centre (no sample shading, no centroid),
// implicit level of detail from the analytic screen-space derivatives of uv = (U/S, V/S):
// rho^2 = max(|d(uv*size)/dx|^2, |d(uv*size)/dy|^2).  Returns false when the fragment is discarded.
VKRT_HD bool rasterFragment(const RasterSetup& t, const TextureRecord* textures, const TextureMips* mips, uint32_t px, uint32_t py, float rgba[4],
                            const float* unorm = nullptr)
{
    const float x = (float)px + 0.5f, y = (float)py + 0.5f;
    const float U = plane3(t.ua, t.ub, t.uc, x, y), V = plane3(t.va, t.vb, t.vc, x, y), S = plane3(t.sa, t.sb, t.sc, x, y);
    const float u = U / S, v = V / S;
    const float S2 = S * S;
    const float dudx = (t.ua * S - U * t.sa) / S2, dudy = (t.ub * S - U * t.sb) / S2;
    const float dvdx = (t.va * S - V * t.sa) / S2, dvdy = (t.vb * S - V * t.sb) / S2;
    const TextureRecord& tex = textures[t.texture];
    const float w = (float)tex.width, h = (float)tex.height;
    const float rx2 = (dudx * w) * (dudx * w) + (dvdx * h) * (dvdx * h);
    const float ry2 = (dudy * w) * (dudy * w) + (dvdy * h) * (dvdy * h);
    sampleTrilinear(tex, mips[t.texture], u, v, fmaxf(rx2, ry2), rgba, unorm);
    return !(rgba[3] < 0.1f);
}

VKRT_HD uint32_t rasterQuantise(const float rgba[4]) // UNORM8 store of one sample, R | G<<8 | B<<16 | A<<24
{
    uint32_t q = 0;
    for (int ch = 0; ch < 4; ++ch)
    {
        q |= (uint32_t)(clamp01(rgba[ch]) * 255.0f + 0.5f) << (8 * ch);
    }
    return q;
}

// One pixel of one triangle: coverage, early depth rejection against the keys already there, fragment shading, then the
// visibility keys of the surviving samples.  `Min` performs key = min(key, candidate) (atomicMin on the device).
template<typename Min>
VKRT_HD void rasterPixel(const RasterSetup& t, uint32_t tri, const TextureRecord* textures, const TextureMips* mips, uint32_t width,
                         uint32_t px, uint32_t py, uint64_t* visibility, Min minKey, const float* unorm = nullptr)
{
    const uint32_t coverage = rasterCoverage(t, px, py);
    if (!coverage)
    {
        return;
    }
    uint64_t* keys = visibility + ((size_t)py * width + px) * kRasterSamples;
    uint32_t alive = 0;
    for (int s = 0; s < kRasterSamples; ++s)
    {
        uint32_t depth24;
        if (((coverage >> s) & 1u) && rasterDepth(t, px, py, s, depth24))
        {
            // keys only ever decrease, so a sample that already holds a smaller key can never be won by this triangle
            alive |= ((((uint64_t)depth24) << 32) | tri) < keys[s] ? 1u << s : 0u;
        }
    }
    if (!alive)
    {
        return;
    }
    float rgba[4];
    if (!rasterFragment(t, textures, mips, px, py, rgba, unorm))
    {
        return; // discard: no depth write either
    }
    for (int s = 0; s < kRasterSamples; ++s)
    {
        uint32_t depth24; // evaluated again rather than kept: eight 64-bit candidates cost 16 registers per thread
        if (((alive >> s) & 1u) && rasterDepth(t, px, py, s, depth24))
        {
            minKey(&keys[s], (((uint64_t)depth24) << 32) | tri);
        }
    }
}

// true when the tile [tx0, tx1] x [ty0, ty1] (pixel indices, inclusive) lies entirely outside one edge
VKRT_HD bool rasterTileOutside(const RasterSetup& s, uint32_t tx0, uint32_t ty0, uint32_t tx1, uint32_t ty1)
{
    // samples lie strictly inside (px, px + 1): the extreme sample coordinates are bounded by the tile's outer corners;
    // the small pad keeps the test conservative against the rounding of the plane evaluation
    const float lx = (float)tx0, hx = (float)(tx1 + 1u), ly = (float)ty0, hy = (float)(ty1 + 1u);
    for (int i = 0; i < 3; ++i)
    {
        const float a = s.ea[i], b = s.eb[i], c = s.ec[i];
        const float best = (a * (a > 0.0f ? hx : lx) + b * (b > 0.0f ? hy : ly)) + c;
        const float pad = (fabsf(a) * hx + fabsf(b) * hy + fabsf(c)) * 1e-5f;
        if (best < -pad)
        {
            return true;
        }
    }
    return false;
}

// Conservative 24-bit depth range of a triangle over the pixel rectangle [x0, x1] x [y0, y1] (inclusive; all sample
// positions of those pixels): the depth plane is linear, so its extremes are at the corners; the pad covers the rounding
// of the plane evaluation and of the quantisation.  Returns false when the plane leaves [0, 1] there (clipped samples).
VKRT_HD bool rasterDepthRange(const RasterSetup& t, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t& zmin24, uint32_t& zmax24)
{
    const float lx = (float)x0, hx = (float)(x1 + 1u), ly = (float)y0, hy = (float)(y1 + 1u);
    const float z00 = plane3(t.za, t.zb, t.zc, lx, ly), z10 = plane3(t.za, t.zb, t.zc, hx, ly);
    const float z01 = plane3(t.za, t.zb, t.zc, lx, hy), z11 = plane3(t.za, t.zb, t.zc, hx, hy);
    const float pad = (fabsf(t.za) * hx + fabsf(t.zb) * hy + fabsf(t.zc)) * 4e-7f + 1e-7f;
    const float lo = fminf(fminf(z00, z10), fminf(z01, z11)) - pad, hi = fmaxf(fmaxf(z00, z10), fmaxf(z01, z11)) + pad;
    if (!(lo >= 0.0f && hi <= 1.0f)) // also false for NaN
    {
        return false;
    }
    zmin24 = (uint32_t)(lo * 16777215.0f + 0.5f);
    zmax24 = (uint32_t)(hi * 16777215.0f + 0.5f);
    zmin24 = zmin24 > 0u ? zmin24 - 1u : 0u;
    zmax24 += 1u;
    return true;
}

// true when every sample of the pixel rectangle lies strictly inside all three edges (no tie rule involved)
VKRT_HD bool rasterTileInside(const RasterSetup& s, uint32_t tx0, uint32_t ty0, uint32_t tx1, uint32_t ty1)
{
    const float lx = (float)tx0, hx = (float)(tx1 + 1u), ly = (float)ty0, hy = (float)(ty1 + 1u);
    for (int i = 0; i < 3; ++i)
    {
        const float a = s.ea[i], b = s.eb[i], c = s.ec[i];
        const float worst = (a * (a > 0.0f ? lx : hx) + b * (b > 0.0f ? ly : hy)) + c;
        const float pad = (fabsf(a) * hx + fabsf(b) * hy + fabsf(c)) * 1e-5f;
        if (!(worst > pad))
        {
            return false;
        }
    }
    return true;
}

// Resolve of pixel (px, py): every sample takes the colour of the triangle that won it (shaded at the pixel centre,
// quantised to UNORM8 as the multisampled colour attachment stores it) or the clear colour; the 8 samples are averaged
// (vkCmdResolveImage / resolve attachment) and stored as B8G8R8A8.  Returns B | G<<8 | R<<16 | A<<24.
VKRT_HD uint32_t rasterResolve(const RasterSetup* setups, const TextureRecord* textures, const TextureMips* mips, uint32_t width, uint32_t px,
                               uint32_t py, const uint64_t* visibility, uint32_t& firstTriangle, const float* unorm = nullptr)
{
    const uint64_t* keys = visibility + ((size_t)py * width + px) * kRasterSamples;
    uint32_t sum[4] = {0, 0, 0, 0};
    uint32_t lastTri = kInvalid, lastColour = 0;
    firstTriangle = (uint32_t)(keys[0] & 0xFFFFFFFFull);
    for (int s = 0; s < kRasterSamples; ++s)
    {
        const uint64_t key = keys[s];
        uint32_t q;
        if (key == kRasterEmpty)
        {
            q = 0u | (0u << 8) | (51u << 16) | (255u << 24); // (0, 0, 0.2, 1): 0.2 * 255 + 0.5 -> 51
        }
        else
        {
            const uint32_t tri = (uint32_t)(key & 0xFFFFFFFFull);
            if (tri != lastTri)
            {
                float rgba[4];
                rasterFragment(setups[tri], textures, mips, px, py, rgba, unorm);
                lastColour = rasterQuantise(rgba);
                lastTri = tri;
            }
            q = lastColour;
        }
        for (int ch = 0; ch < 4; ++ch)
        {
            sum[ch] += (q >> (8 * ch)) & 0xFFu;
        }
    }
    uint32_t out[4];
    for (int ch = 0; ch < 4; ++ch)
    {
        out[ch] = (uint32_t)((float)sum[ch] / (float)kRasterSamples + 0.5f);
    }
    return out[2] | (out[1] << 8) | (out[0] << 16) | (out[3] << 24);
}

} // namespace vkrt
