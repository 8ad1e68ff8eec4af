/*
 * vkrt_oracle.cpp — CPU ORACLE (test infrastructure, see vkrt_oracle.h).  PARITY UNPINNED (see header).
 *
 * Restates, in scalar FP32 C++ with no FMA contraction (build with -ffp-contract=off):
 *   shader.rgen:30-77, shader.rchit:78-172, shader.rmiss:15-18, shader_shadow.rmiss:6-9
 * and the traceRayEXT semantics the Vulkan driver provides to them.  Every arithmetic expression
 * below fixes an evaluation order (left-to-right, no fused multiply-add); the CUDA kernels restate
 * the same order independently so the two can be compared bit-for-bit.
 *
 * Conventions fixed here (the GLSL/Vulkan specs leave them open):
 *   normalize(v) = v / sqrt(dot(v,v));  length(v) = sqrt(dot(v,v));  dot = (x*x' + y*y') + z*z'
 *   clamp(x,0,1) = fmin(fmax(x,0),1)   (NaN -> 0, CUDA fmaxf/fminf semantics)
 *   mat*vec      = ((c0*x + c1*y) + c2*z) + c3*w  per row
 *   ray/triangle = Woop, Benthin, Wald 2013 "Watertight Ray/Triangle Intersection", no culling,
 *                  hit interval tmin < t < tmax exclusive (Vulkan spec), equal-t ties -> lower id
 *   non-finite or zero-direction rays miss everything
 */
#include "vkrt_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace
{
struct V3
{
    float x, y, z;
};
inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline float length(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalize(V3 a) { return a / length(a); }
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline float clamp01(float x) { return std::fmin(std::fmax(x, 0.0f), 1.0f); }
inline float comp(V3 v, int k) { return k == 0 ? v.x : (k == 1 ? v.y : v.z); }
inline bool finite3(V3 v) { return std::isfinite(v.x) && std::isfinite(v.y) && std::isfinite(v.z); }

/* column-major mat4 * vec4, rows 0..2 */
inline V3 mat4MulVec(const float* m, float x, float y, float z, float w)
{
    V3 r;
    r.x = ((m[0] * x + m[4] * y) + m[8] * z) + m[12] * w;
    r.y = ((m[1] * x + m[5] * y) + m[9] * z) + m[13] * w;
    r.z = ((m[2] * x + m[6] * y) + m[10] * z) + m[14] * w;
    return r;
}

/* row-major 3x4 affine * (p,1) and * (d,0) */
inline V3 affinePoint(const float* m, V3 p)
{
    V3 r;
    r.x = ((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3];
    r.y = ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7];
    r.z = ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11];
    return r;
}
inline V3 affineDir(const float* m, V3 d)
{
    V3 r;
    r.x = (m[0] * d.x + m[1] * d.y) + m[2] * d.z;
    r.y = (m[4] * d.x + m[5] * d.y) + m[6] * d.z;
    r.z = (m[8] * d.x + m[9] * d.y) + m[10] * d.z;
    return r;
}

void affineInverse(const float m[12], float inv[12])
{
    const double a = m[0], b = m[1], c = m[2], tx = m[3];
    const double d = m[4], e = m[5], f = m[6], ty = m[7];
    const double g = m[8], h = m[9], i = m[10], tz = m[11];
    const double c00 = e * i - f * h;
    const double c01 = f * g - d * i;
    const double c02 = d * h - e * g;
    const double det = (a * c00 + b * c01) + c * c02;
    const double r00 = c00 / det, r01 = (c * h - b * i) / det, r02 = (b * f - c * e) / det;
    const double r10 = c01 / det, r11 = (a * i - c * g) / det, r12 = (c * d - a * f) / det;
    const double r20 = c02 / det, r21 = (b * g - a * h) / det, r22 = (a * e - b * d) / det;
    inv[0] = (float)r00, inv[1] = (float)r01, inv[2] = (float)r02;
    inv[4] = (float)r10, inv[5] = (float)r11, inv[6] = (float)r12;
    inv[8] = (float)r20, inv[9] = (float)r21, inv[10] = (float)r22;
    inv[3] = (float)(-((r00 * tx + r01 * ty) + r02 * tz));
    inv[7] = (float)(-((r10 * tx + r11 * ty) + r12 * tz));
    inv[11] = (float)(-((r20 * tx + r21 * ty) + r22 * tz));
}

struct Tri
{
    V3 v0, v1, v2;
    uint32_t geometry, primitive, globalTri;
};

struct Box
{
    float lo[3], hi[3];
    void reset()
    {
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = std::numeric_limits<float>::infinity();
            hi[k] = -std::numeric_limits<float>::infinity();
        }
    }
    void grow(V3 p)
    {
        lo[0] = std::min(lo[0], p.x), lo[1] = std::min(lo[1], p.y), lo[2] = std::min(lo[2], p.z);
        hi[0] = std::max(hi[0], p.x), hi[1] = std::max(hi[1], p.y), hi[2] = std::max(hi[2], p.z);
    }
    void grow(const Box& b)
    {
        for (int k = 0; k < 3; ++k)
        {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    float halfArea() const
    {
        const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
};

struct Node
{
    Box box;
    uint32_t left;  /* inner: index of left child (right = left+1); leaf: first triangle slot */
    uint32_t count; /* 0 = inner, else number of triangles */
};

struct Image
{
    uint32_t w, h;
    std::vector<uint8_t> px;
};

/* One generated mip level (levels 1.. of an Image; level 0 is the Image itself). */
struct MipLevel
{
    uint32_t w, h;
    std::vector<uint8_t> px;
};

struct Instance
{
    float o2w[12];
    float w2o[12];
    uint32_t mask = 0xFFu; /* VkAccelerationStructureInstanceKHR::mask, 0xFF in the reference (Raytracer.cpp:1174) */
};

/* Watertight per-ray precomputation (Woop et al. 2013, section 3). */
struct RayPre
{
    int kx, ky, kz;
    float Sx, Sy, Sz;
    V3 mx, my, mz; /* shear as three vectors: mx = 1 at kx, -Sx at kz; my = 1 at ky, -Sy at kz; mz = Sz at kz; 0 elsewhere */
};

inline void setComp(V3& v, int k, float value)
{
    if (k == 0)
    {
        v.x = value;
    }
    else if (k == 1)
    {
        v.y = value;
    }
    else
    {
        v.z = value;
    }
}

/* NUMERICAL CONTRACT of the shear (Woop et al. 2013, eq. 2-4: A' = A[kx] - Sx A[kz], ...): one fixed sequence of fused
 * operations, fma(a.x, m.x, fma(a.y, m.y, a.z * m.z)).  GLSL leaves contraction to the implementation (no `precise`
 * qualifier in the reference's shaders, and the driver's intersector is opaque anyway); fixing it to this form keeps the
 * result a function of (vertex, ray) only - the property watertightness needs - and lets the CUDA path spend 3
 * instead of 5 instructions per sheared coordinate.  The edge functions below stay unfused on purpose. */
inline float shearDot(V3 a, V3 m)
{
    return std::fmaf(a.x, m.x, std::fmaf(a.y, m.y, a.z * m.z));
}

inline RayPre rayPre(V3 d)
{
    RayPre r;
    const float ax = std::fabs(d.x), ay = std::fabs(d.y), az = std::fabs(d.z);
    r.kz = 0;
    float best = ax;
    if (ay > best)
    {
        r.kz = 1;
        best = ay;
    }
    if (az > best)
    {
        r.kz = 2;
    }
    r.kx = (r.kz + 1) % 3;
    r.ky = (r.kx + 1) % 3;
    if (comp(d, r.kz) < 0.0f)
    {
        std::swap(r.kx, r.ky);
    }
    const float dz = comp(d, r.kz);
    r.Sx = comp(d, r.kx) / dz;
    r.Sy = comp(d, r.ky) / dz;
    r.Sz = 1.0f / dz;
    r.mx = r.my = r.mz = mk(0.0f, 0.0f, 0.0f);
    setComp(r.mx, r.kx, 1.0f);
    setComp(r.mx, r.kz, -r.Sx);
    setComp(r.my, r.ky, 1.0f);
    setComp(r.my, r.kz, -r.Sy);
    setComp(r.mz, r.kz, r.Sz);
    return r;
}

/* Returns true and updates (t,b1,b2) if the triangle is hit with tmin < t < tmax. */
inline bool intersectTri(const Tri& tri, V3 o, const RayPre& rp, float tmin, float tmax, float& tOut, float& b1,
                         float& b2)
{
    const V3 A = tri.v0 - o, B = tri.v1 - o, C = tri.v2 - o;
    const float Ax = shearDot(A, rp.mx), Ay = shearDot(A, rp.my);
    const float Bx = shearDot(B, rp.mx), By = shearDot(B, rp.my);
    const float Cx = shearDot(C, rp.mx), Cy = shearDot(C, rp.my);
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f)
    {
        U = (float)((double)Cx * (double)By - (double)Cy * (double)Bx);
        V = (float)((double)Ax * (double)Cy - (double)Ay * (double)Cx);
        W = (float)((double)Bx * (double)Ay - (double)By * (double)Ax);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f))
    {
        return false;
    }
    const float det = (U + V) + W;
    if (det == 0.0f)
    {
        return false;
    }
    const float Az = shearDot(A, rp.mz), Bz = shearDot(B, rp.mz), Cz = shearDot(C, rp.mz);
    const float T = std::fmaf(W, Cz, std::fmaf(V, Bz, U * Az));
    const float rcp = 1.0f / det;
    const float t = T * rcp;
    if (!(t > tmin && t < tmax))
    {
        return false;
    }
    tOut = t;
    b1 = V * rcp;
    b2 = W * rcp;
    return true;
}

struct Hit
{
    float t, b1, b2;
    uint32_t instance, geometry, primitive, globalTri;
    bool valid;
};

struct Counters
{
    uint64_t raysPrimary = 0, raysShadow = 0, raysReflection = 0, nodes = 0, tris = 0;
    void add(const Counters& o)
    {
        raysPrimary += o.raysPrimary, raysShadow += o.raysShadow, raysReflection += o.raysReflection;
        nodes += o.nodes, tris += o.tris;
    }
};
} // namespace

struct OrcScene
{
    std::vector<float> vertices; /* 16 floats per vertex */
    std::vector<uint32_t> indices;
    std::vector<OrcGeometryRange> ranges;
    std::vector<OrcSubmeshInfo> rows;
    std::vector<Image> images;
    std::vector<std::vector<MipLevel>> mips; /* per image: levels 1.. (orc_scene_generate_mips) */
    int lodMode = 0;                         /* 0 = LOD 0 (the reference's ray-tracing path), 1 = ray cones */
    std::vector<Instance> instances;
    std::vector<Tri> tris;       /* in (geometry, primitive) order */
    std::vector<uint32_t> order; /* BVH leaf order -> index into tris */
    std::vector<Node> nodes;
    double buildSeconds = 0.0;

    void buildTriangles();
    void buildBvh();
    uint32_t buildRange(std::vector<Box>& tb, std::vector<V3>& cent, uint32_t first, uint32_t count);
    void subdivide(uint32_t nodeIdx, std::vector<Box>& tb, std::vector<V3>& cent);

    bool traceObject(V3 o, V3 d, float tmin, float& tmax, bool anyHit, int mode, uint32_t instance, Hit& hit,
                     Counters& c) const;
    Hit trace(V3 o, V3 d, float tmin, float tmax, bool anyHit, int mode, Counters& c) const;
    void sample(uint32_t index, float u, float v, float rgba[4]) const;
    void sampleLod(uint32_t index, float u, float v, float footprint2, float rgba[4]) const;
    void sampleTexels2(uint32_t index, float u, float v, float texels2, float rgba[4]) const;
    void generateMips();
};

void OrcScene::buildTriangles()
{
    tris.clear();
    uint32_t globalTri = 0;
    for (uint32_t g = 0; g < ranges.size(); ++g)
    {
        const uint64_t firstIndex = ranges[g].indexByteOffset / 4;
        for (uint32_t p = 0; p < ranges[g].primitiveCount; ++p)
        {
            Tri t;
            const uint32_t i0 = indices[firstIndex + 3 * (uint64_t)p + 0];
            const uint32_t i1 = indices[firstIndex + 3 * (uint64_t)p + 1];
            const uint32_t i2 = indices[firstIndex + 3 * (uint64_t)p + 2];
            const float* a = &vertices[(size_t)i0 * 16];
            const float* b = &vertices[(size_t)i1 * 16];
            const float* c = &vertices[(size_t)i2 * 16];
            t.v0 = mk(a[0], a[1], a[2]);
            t.v1 = mk(b[0], b[1], b[2]);
            t.v2 = mk(c[0], c[1], c[2]);
            t.geometry = g;
            t.primitive = p;
            /* global id as the closest-hit shader forms it: indexBufferOffset + gl_PrimitiveID (rchit:88-89) */
            t.globalTri = (uint32_t)rows[g].indexBufferOffset + p;
            (void)globalTri;
            tris.push_back(t);
        }
    }
}

void OrcScene::subdivide(uint32_t nodeIdx, std::vector<Box>& tb, std::vector<V3>& cent)
{
    /* iterative binned-SAH build; explicit stack to survive deep trees */
    std::vector<uint32_t> stack{nodeIdx};
    while (!stack.empty())
    {
        const uint32_t ni = stack.back();
        stack.pop_back();
        const uint32_t first = nodes[ni].left, count = nodes[ni].count;
        if (count <= 2)
        {
            continue;
        }
        Box cb;
        cb.reset();
        for (uint32_t i = first; i < first + count; ++i)
        {
            cb.grow(cent[order[i]]);
        }
        constexpr int BINS = 16;
        float bestCost = std::numeric_limits<float>::infinity();
        int bestAxis = -1, bestBin = -1;
        for (int axis = 0; axis < 3; ++axis)
        {
            const float lo = cb.lo[axis], hi = cb.hi[axis];
            if (!(hi > lo))
            {
                continue;
            }
            Box bins[BINS];
            uint32_t cnt[BINS] = {};
            for (auto& b : bins)
            {
                b.reset();
            }
            const float scale = (float)BINS / (hi - lo);
            for (uint32_t i = first; i < first + count; ++i)
            {
                const uint32_t ti = order[i];
                int b = (int)((comp(cent[ti], axis) - lo) * scale);
                b = std::min(std::max(b, 0), BINS - 1);
                bins[b].grow(tb[ti]);
                ++cnt[b];
            }
            float rightArea[BINS];
            uint32_t rightCnt[BINS];
            Box acc;
            acc.reset();
            uint32_t n = 0;
            for (int b = BINS - 1; b > 0; --b)
            {
                acc.grow(bins[b]);
                n += cnt[b];
                rightArea[b] = n ? acc.halfArea() : 0.0f;
                rightCnt[b] = n;
            }
            acc.reset();
            n = 0;
            for (int b = 0; b < BINS - 1; ++b)
            {
                acc.grow(bins[b]);
                n += cnt[b];
                if (n == 0 || rightCnt[b + 1] == 0)
                {
                    continue;
                }
                const float cost = acc.halfArea() * (float)n + rightArea[b + 1] * (float)rightCnt[b + 1];
                if (cost < bestCost)
                {
                    bestCost = cost, bestAxis = axis, bestBin = b;
                }
            }
        }
        uint32_t mid;
        if (bestAxis < 0)
        {
            if (count <= 4)
            {
                continue;
            }
            mid = first + count / 2; /* all centroids equal: split by order */
        }
        else
        {
            const float leafCost = nodes[ni].box.halfArea() * (float)count;
            if (count <= 4 && bestCost >= leafCost)
            {
                continue;
            }
            const float lo = cb.lo[bestAxis], hi = cb.hi[bestAxis];
            const float scale = (float)BINS / (hi - lo);
            auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t ti) {
                int b = (int)((comp(cent[ti], bestAxis) - lo) * scale);
                b = std::min(std::max(b, 0), BINS - 1);
                return b <= bestBin;
            });
            mid = (uint32_t)(it - order.begin());
            if (mid == first || mid == first + count)
            {
                mid = first + count / 2;
            }
        }
        const uint32_t l = (uint32_t)nodes.size();
        nodes.push_back(Node{});
        nodes.push_back(Node{});
        nodes[l].left = first, nodes[l].count = mid - first;
        nodes[l + 1].left = mid, nodes[l + 1].count = first + count - mid;
        for (int s = 0; s < 2; ++s)
        {
            Node& n = nodes[l + s];
            n.box.reset();
            for (uint32_t i = n.left; i < n.left + n.count; ++i)
            {
                n.box.grow(tb[order[i]]);
            }
        }
        nodes[ni].left = l;
        nodes[ni].count = 0;
        stack.push_back(l);
        stack.push_back(l + 1);
    }
}

void OrcScene::buildBvh()
{
    const auto t0 = std::chrono::steady_clock::now();
    const uint32_t n = (uint32_t)tris.size();
    std::vector<Box> tb(n);
    std::vector<V3> cent(n);
    order.resize(n);
    nodes.clear();
    nodes.reserve(2 * (size_t)n + 1);
    Node root;
    root.box.reset();
    for (uint32_t i = 0; i < n; ++i)
    {
        tb[i].reset();
        tb[i].grow(tris[i].v0), tb[i].grow(tris[i].v1), tb[i].grow(tris[i].v2);
        cent[i] = mk(0.5f * (tb[i].lo[0] + tb[i].hi[0]), 0.5f * (tb[i].lo[1] + tb[i].hi[1]),
                     0.5f * (tb[i].lo[2] + tb[i].hi[2]));
        order[i] = i;
        root.box.grow(tb[i]);
    }
    root.left = 0;
    root.count = n;
    nodes.push_back(root);
    if (n > 0)
    {
        subdivide(0, tb, cent);
    }
    buildSeconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

/* Conservative slab test in double precision with relative slack, so that BVH culling can never
 * reject a triangle the brute-force loop would accept. */
static inline bool slab(const Box& b, const double o[3], const double inv[3], double tmin, double tmax, double& tnear)
{
    double tn = tmin, tf = tmax;
    for (int k = 0; k < 3; ++k)
    {
        double t0 = ((double)b.lo[k] - o[k]) * inv[k];
        double t1 = ((double)b.hi[k] - o[k]) * inv[k];
        if (t0 != t0 || t1 != t1)
        {
            continue; /* 0 * inf: the ray lies in the plane; do not cull on this axis */
        }
        if (t0 > t1)
        {
            std::swap(t0, t1);
        }
        t0 -= std::fabs(t0) * 1e-4;
        t1 += std::fabs(t1) * 1e-4;
        tn = std::max(tn, t0);
        tf = std::min(tf, t1);
    }
    tnear = tn;
    return tn <= tf;
}

/* Intersect one instance's BLAS with an object-space ray. `tmax` shrinks on closest hits. */
bool OrcScene::traceObject(V3 o, V3 d, float tmin, float& tmax, bool anyHit, int mode, uint32_t instance, Hit& hit,
                           Counters& c) const
{
    bool found = false;
    const RayPre rp = rayPre(d);
    auto test = [&](uint32_t ti) -> bool {
        const Tri& tri = tris[ti];
        float t, b1, b2;
        ++c.tris;
        /* accept strictly closer hits; on an exact tie in t prefer the lower (instance, triangle) id */
        if (intersectTri(tri, o, rp, tmin, std::nextafter(tmax, std::numeric_limits<float>::infinity()), t, b1, b2))
        {
            if (!(t < tmax))
            {
                /* t == tmax (previous best or the ray's tmax): only a tie with a previous hit counts */
                if (!(hit.valid && t == hit.t))
                {
                    return false;
                }
                const uint64_t idNew = ((uint64_t)instance << 32) | tri.globalTri;
                const uint64_t idOld = ((uint64_t)hit.instance << 32) | hit.globalTri;
                if (idNew >= idOld)
                {
                    return false;
                }
            }
            hit.t = t, hit.b1 = b1, hit.b2 = b2;
            hit.instance = instance, hit.geometry = tri.geometry, hit.primitive = tri.primitive;
            hit.globalTri = tri.globalTri;
            hit.valid = true;
            tmax = t;
            found = true;
            return true;
        }
        return false;
    };

    if (mode == ORC_MODE_BRUTE)
    {
        for (uint32_t i = 0; i < tris.size(); ++i)
        {
            if (test(i) && anyHit)
            {
                return true;
            }
        }
        return found;
    }

    if (nodes.empty() || tris.empty())
    {
        return false;
    }
    const double od[3] = {o.x, o.y, o.z};
    const double inv[3] = {1.0 / (double)d.x, 1.0 / (double)d.y, 1.0 / (double)d.z};
    const double slackLo = (double)tmin * (1.0 - 1e-4);
    uint32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0)
    {
        const Node& n = nodes[stack[--sp]];
        double tn;
        ++c.nodes;
        if (!slab(n.box, od, inv, slackLo, (double)tmax * (1.0 + 1e-4), tn))
        {
            continue;
        }
        if (n.count)
        {
            for (uint32_t i = n.left; i < n.left + n.count; ++i)
            {
                if (test(order[i]) && anyHit)
                {
                    return true;
                }
            }
        }
        else
        {
            /* near child last so it is popped first */
            double t0, t1;
            const bool h0 = slab(nodes[n.left].box, od, inv, slackLo, (double)tmax * (1.0 + 1e-4), t0);
            const bool h1 = slab(nodes[n.left + 1].box, od, inv, slackLo, (double)tmax * (1.0 + 1e-4), t1);
            if (h0 && h1)
            {
                if (t0 <= t1)
                {
                    stack[sp++] = n.left + 1, stack[sp++] = n.left;
                }
                else
                {
                    stack[sp++] = n.left, stack[sp++] = n.left + 1;
                }
            }
            else if (h0)
            {
                stack[sp++] = n.left;
            }
            else if (h1)
            {
                stack[sp++] = n.left + 1;
            }
        }
    }
    return found;
}

/* traceRayEXT over the TLAS: every instance (mask 0xFF, no culling; Raytracer.cpp:1172-1177). */
Hit OrcScene::trace(V3 o, V3 d, float tmin, float tmax, bool anyHit, int mode, Counters& c) const
{
    Hit hit{};
    hit.valid = false;
    hit.instance = 0xFFFFFFFFu;
    if (!finite3(o) || !finite3(d) || (d.x == 0.0f && d.y == 0.0f && d.z == 0.0f) || !(tmax > tmin))
    {
        return hit;
    }
    for (uint32_t i = 0; i < instances.size(); ++i)
    {
        /* instance mask against the cullMask of traceRayEXT, 0xFF at both call sites (shader.rgen:52, shader.rchit:133) */
        if ((instances[i].mask & 0xFFu) == 0u)
        {
            continue;
        }
        /* object-space ray: origin and direction through world->object, t shared (Vulkan spec) */
        const V3 oo = affinePoint(instances[i].w2o, o);
        const V3 dd = affineDir(instances[i].w2o, d);
        if (!finite3(oo) || !finite3(dd) || (dd.x == 0.0f && dd.y == 0.0f && dd.z == 0.0f))
        {
            continue;
        }
        if (traceObject(oo, dd, tmin, tmax, anyHit, mode, i, hit, c) && anyHit)
        {
            return hit;
        }
    }
    return hit;
}

/* Bilinear, REPEAT, texel = u8/255 (UNORM). Vulkan texel addressing: u*W - 0.5. */
static void bilinearRepeat(const uint8_t* px, uint32_t iw, uint32_t ih, float u, float v, float rgba[4])
{
    float x = u * (float)iw - 0.5f;
    float y = v * (float)ih - 0.5f;
    if (!std::isfinite(x))
    {
        x = -0.5f;
    }
    if (!std::isfinite(y))
    {
        y = -0.5f;
    }
    const float fx = std::floor(x), fy = std::floor(y);
    const float a = x - fx, b = y - fy;
    auto wrap = [](float f, uint32_t n) -> uint32_t {
        /* floor value -> [0,n) with REPEAT; f is integral. Large magnitudes go through fmod first. */
        float m = std::fmod(f, (float)n);
        if (m < 0.0f)
        {
            m += (float)n;
        }
        uint32_t i = (uint32_t)m;
        return i >= n ? 0u : i;
    };
    const uint32_t x0 = wrap(fx, iw), y0 = wrap(fy, ih);
    const uint32_t x1 = (x0 + 1 == iw) ? 0u : x0 + 1, y1 = (y0 + 1 == ih) ? 0u : y0 + 1;
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    const uint8_t* p00 = &px[((size_t)y0 * iw + x0) * 4];
    const uint8_t* p10 = &px[((size_t)y0 * iw + x1) * 4];
    const uint8_t* p01 = &px[((size_t)y1 * iw + x0) * 4];
    const uint8_t* p11 = &px[((size_t)y1 * iw + x1) * 4];
    for (int ch = 0; ch < 4; ++ch)
    {
        const float t00 = (float)p00[ch] / 255.0f, t10 = (float)p10[ch] / 255.0f;
        const float t01 = (float)p01[ch] / 255.0f, t11 = (float)p11[ch] / 255.0f;
        rgba[ch] = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
    }
}

/* texture() at LOD 0: what rchit:107,157,162 get on the ray-tracing path (no derivatives). */
void OrcScene::sample(uint32_t index, float u, float v, float rgba[4]) const
{
    const Image& img = images[index];
    bilinearRepeat(img.px.data(), img.w, img.h, u, v, rgba);
}

/*
 * Mip chain of Raytracer::createMipmaps (src/Raytracer.cpp:572-640): floor(log2(max(w,h)))+1 levels (:467-468), level k
 * = vkCmdBlitImage(level k-1 -> k, VK_FILTER_LINEAR) with extents max(w/2,1) x max(h/2,1) (:612,629-630).
 * Blit addressing (Vulkan spec, "Image Copies with Scaling"): source coordinate of destination texel centre i is
 * (i + 0.5) * srcExtent/dstExtent; linear filter = bilinear around coordinate - 0.5, clamp to edge; UNORM8 store rounds
 * to nearest.
 */
void OrcScene::generateMips()
{
    mips.assign(images.size(), {});
    for (size_t n = 0; n < images.size(); ++n)
    {
        const Image& img = images[n];
        const uint32_t levelCount = (uint32_t)std::floor(std::log2((double)std::max(img.w, img.h))) + 1u;
        const uint8_t* src = img.px.data();
        uint32_t sw = img.w, sh = img.h;
        for (uint32_t level = 1; level < levelCount; ++level)
        {
            MipLevel out;
            out.w = sw > 1 ? sw / 2 : 1, out.h = sh > 1 ? sh / 2 : 1;
            out.px.resize((size_t)out.w * out.h * 4);
            const float scaleX = (float)sw / (float)out.w, scaleY = (float)sh / (float)out.h;
            for (uint32_t j = 0; j < out.h; ++j)
            {
                for (uint32_t i = 0; i < out.w; ++i)
                {
                    const float cx = ((float)i + 0.5f) * scaleX - 0.5f, cy = ((float)j + 0.5f) * scaleY - 0.5f;
                    const float fx = std::floor(cx), fy = std::floor(cy);
                    const float a = cx - fx, b = cy - fy;
                    auto edge = [](long c, uint32_t n) -> uint32_t { return (uint32_t)std::min<long>(std::max<long>(c, 0), (long)n - 1); };
                    const uint32_t x0 = edge((long)fx, sw), x1 = edge((long)fx + 1, sw);
                    const uint32_t y0 = edge((long)fy, sh), y1 = edge((long)fy + 1, sh);
                    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
                    for (int ch = 0; ch < 4; ++ch)
                    {
                        const float t00 = (float)src[((size_t)y0 * sw + x0) * 4 + ch] / 255.0f;
                        const float t10 = (float)src[((size_t)y0 * sw + x1) * 4 + ch] / 255.0f;
                        const float t01 = (float)src[((size_t)y1 * sw + x0) * 4 + ch] / 255.0f;
                        const float t11 = (float)src[((size_t)y1 * sw + x1) * 4 + ch] / 255.0f;
                        const float c = ((t00 * w00 + t10 * w10) + t01 * w01) + t11 * w11;
                        out.px[((size_t)j * out.w + i) * 4 + ch] = (uint8_t)(clamp01(c) * 255.0f + 0.5f);
                    }
                }
            }
            mips[n].push_back(std::move(out));
            const MipLevel& made = mips[n].back();
            src = made.px.data(), sw = made.w, sh = made.h;
        }
    }
}

/*
 * log2 for the level-of-detail computation: exponent + cubic-corrected mantissa term, + and * only in a fixed order
 * (|error| < 1.5e-4).  Part of the numerical contract of the ray-cone mode: libm's log2f and the
 * GPU's differ in the last bits, this does not.
 */
static float lodLog2(float x)
{
    uint32_t bits;
    std::memcpy(&bits, &x, 4);
    const float e = (float)((int)(bits >> 23) - 127);
    const uint32_t mantissaBits = (bits & 0x007FFFFFu) | 0x3F800000u;
    float m;
    std::memcpy(&m, &mantissaBits, 4);
    const float f = m - 1.0f;
    const float poly = 0.4380733f + f * (-0.2366934f + f * 0.0803073f);
    return e + (f + (f * (1.0f - f)) * poly);
}

/*
 * Trilinear sample at the ray-cone level of detail (Akenine-Moller et al., Ray Tracing Gems 2019, ch. 20):
 * lambda = 0.5 * log2(footprint2 * W * H), clamped to the chain; footprints of at most one texel (and NaN) take level 0.
 */
void OrcScene::sampleLod(uint32_t index, float u, float v, float footprint2, float rgba[4]) const
{
    const Image& img = images[index];
    sampleTexels2(index, u, v, (footprint2 * (float)img.w) * (float)img.h, rgba);
}

/* Trilinear sample for a squared footprint of `texels2` level-0 texels (mip LINEAR, minLod 0, all levels). */
void OrcScene::sampleTexels2(uint32_t index, float u, float v, float texels2, float rgba[4]) const
{
    const Image& img = images[index];
    static const std::vector<MipLevel> none;
    const std::vector<MipLevel>& chain = mips.size() == images.size() ? mips[index] : none;
    if (!(texels2 > 1.0f) || !std::isfinite(texels2) || chain.empty())
    {
        bilinearRepeat(img.px.data(), img.w, img.h, u, v, rgba);
        return;
    }
    float lambda = 0.5f * lodLog2(texels2);
    const float top = (float)chain.size(); /* index of the last level */
    if (lambda > top)
    {
        lambda = top;
    }
    const float fl = std::floor(lambda);
    const float frac = lambda - fl;
    const uint32_t l0 = (uint32_t)fl;
    auto level = [&](uint32_t l, float out[4]) {
        if (l == 0)
        {
            bilinearRepeat(img.px.data(), img.w, img.h, u, v, out);
        }
        else
        {
            bilinearRepeat(chain[l - 1].px.data(), chain[l - 1].w, chain[l - 1].h, u, v, out);
        }
    };
    level(l0, rgba);
    if (frac > 0.0f)
    {
        float hi[4];
        level(l0 + 1, hi);
        for (int ch = 0; ch < 4; ++ch)
        {
            rgba[ch] = rgba[ch] * (1.0f - frac) + hi[ch] * frac;
        }
    }
}

namespace
{
float radicalInverse(uint32_t s, uint32_t base)
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

struct Payload
{
    V3 hitValue;
    int done;
    int depth;
    float attenuation;
    V3 rayOrigin;
    V3 rayDir;
};

/* shader.rchit main(), lines 86-172. Returns the 8-bit shadow mask of this hit. */
uint8_t closestHit(const OrcScene& s, const OrcUniforms& u, const Hit& hit, V3 worldRayDir, Payload& payload, int mode,
                   Counters& c, float coneWidth)
{
    const OrcSubmeshInfo& mat = s.rows[hit.geometry];
    const uint64_t tri = (uint64_t)(uint32_t)mat.indexBufferOffset + hit.primitive; /* rchit:88-89 */
    const float* v0 = &s.vertices[(size_t)s.indices[3 * tri + 0] * 16];
    const float* v1 = &s.vertices[(size_t)s.indices[3 * tri + 1] * 16];
    const float* v2 = &s.vertices[(size_t)s.indices[3 * tri + 2] * 16];
    const Instance& inst = s.instances[hit.instance];

    const float bx = 1.0f - hit.b1 - hit.b2, by = hit.b1, bz = hit.b2; /* rchit:94 */
    auto interp3 = [&](int off) {
        return (mk(v0[off], v0[off + 1], v0[off + 2]) * bx + mk(v1[off], v1[off + 1], v1[off + 2]) * by) +
               mk(v2[off], v2[off + 1], v2[off + 2]) * bz;
    };
    const float uvx = (v0[8] * bx + v1[8] * by) + v2[8] * bz; /* rchit:95 */
    const float uvy = (v0[9] * bx + v1[9] * by) + v2[9] * bz;
    const V3 position = interp3(0);                         /* rchit:97 */
    const V3 worldPos = affinePoint(inst.o2w, position);    /* rchit:98 */
    const V3 normal = interp3(4);                           /* rchit:100 */
    /* normal * gl_WorldToObjectEXT: row vector times the 3x3 part (rchit:101) */
    const float* w = inst.w2o;
    const V3 nW = mk((normal.x * w[0] + normal.y * w[4]) + normal.z * w[8],
                     (normal.x * w[1] + normal.y * w[5]) + normal.z * w[9],
                     (normal.x * w[2] + normal.y * w[6]) + normal.z * w[10]);
    const V3 worldNormal = normalize(nW);
    const V3 tangent = interp3(12); /* rchit:103 */

    /* getTBN (rchit:78-84) with M = identity */
    const V3 N = worldNormal;
    const V3 T = normalize(tangent);
    const V3 B = cross(T, N);
    float texel[4];
    /* Ray-cone mode (not in the reference): squared uv footprint of the cone on this triangle =
     * (uv double area / world double area) * (coneWidth / |n.d|)^2, n the geometric normal. */
    float footprint2 = 0.0f;
    const bool lod = s.lodMode == 1;
    if (lod)
    {
        const V3 q0 = affinePoint(inst.o2w, mk(v0[0], v0[1], v0[2]));
        const V3 q1 = affinePoint(inst.o2w, mk(v1[0], v1[1], v1[2]));
        const V3 q2 = affinePoint(inst.o2w, mk(v2[0], v2[1], v2[2]));
        const V3 areaVec = cross(q1 - q0, q2 - q0);
        const float worldArea2 = length(areaVec);
        const float uvArea2 = std::fabs((v1[8] - v0[8]) * (v2[9] - v0[9]) - (v2[8] - v0[8]) * (v1[9] - v0[9]));
        const float ndotd = std::fabs(dot(areaVec, worldRayDir)) / (worldArea2 * length(worldRayDir));
        const float wn = coneWidth / ndotd;
        footprint2 = (uvArea2 / worldArea2) * (wn * wn);
    }
    auto fetch = [&](int32_t textureIndex, float out[4]) {
        /* textures[-1] (a material without that texture, which the reference does not clamp for base colour,
         * Raytracer.cpp:1424-1426) or an index past the array is undefined in Vulkan; both sample image 0 here. */
        if (textureIndex < 0 || (size_t)textureIndex >= s.images.size())
        {
            textureIndex = 0;
        }
        if (lod)
        {
            s.sampleLod((uint32_t)textureIndex, uvx, uvy, footprint2, out);
        }
        else
        {
            s.sample((uint32_t)textureIndex, uvx, uvy, out);
        }
    };
    fetch(mat.normalTextureIndex, texel); /* rchit:106-107 */
    const V3 mapN = mk(texel[0] * 2.0f - 1.0f, texel[1] * 2.0f - 1.0f, texel[2] * 2.0f - 1.0f);
    const V3 mn = normalize(mapN);
    const V3 perturbed = normalize((T * mn.x + B * mn.y) + N * mn.z); /* rchit:108 */

    float totalLightAmount = 0.0f;
    const float lightIntensity = 10.0f;
    uint8_t mask = 0;
    for (int i = 0; i < 4; ++i) /* rchit:119-152 */
    {
        const V3 lightVec = mk(u.lightPositions[i][0], u.lightPositions[i][1], u.lightPositions[i][2]) - worldPos;
        const float lightDistance = length(lightVec);
        const V3 lightDir = lightVec / lightDistance;
        const float ndl = dot(perturbed, lightDir);
        const float diffuse = clamp01(ndl);
        const float lightPower = lightIntensity / (lightDistance * lightDistance);
        float shadowMultiplier = 1.0f;
        if (ndl > 0.0f)
        {
            ++c.raysShadow;
            mask |= (uint8_t)(0x10u << i);
            const Hit sh = s.trace(worldPos, lightDir, 0.001f, lightDistance, true, mode, c);
            if (sh.valid)
            {
                shadowMultiplier = 0.3f;
                mask |= (uint8_t)(1u << i);
            }
        }
        totalLightAmount += (diffuse * lightPower) * shadowMultiplier;
    }

    const float ambient = 0.1f;
    fetch(mat.baseColorTextureIndex, texel); /* rchit:156-157 */
    const V3 baseColor = mk(texel[0], texel[1], texel[2]);
    payload.hitValue = (baseColor * totalLightAmount) * payload.attenuation + baseColor * ambient; /* rchit:158 */

    fetch(mat.metallicRoughnessTextureIndex, texel); /* rchit:161-162 */
    const float metallic = texel[2];
    if (metallic > 0.1f) /* rchit:163-171 */
    {
        const float reflectAmount = 0.5f * metallic;
        payload.attenuation *= reflectAmount;
        payload.hitValue = payload.hitValue * (1.0f - payload.attenuation);
        payload.done = 0;
        payload.rayOrigin = worldPos;
        /* reflect(I,N) = I - 2*dot(N,I)*N */
        const float k = 2.0f * dot(perturbed, worldRayDir);
        payload.rayDir = worldRayDir - perturbed * k;
    }
    return mask;
}

void primaryRay(const OrcUniforms& u, uint32_t width, uint32_t height, uint32_t x, uint32_t y, float jx, float jy,
                V3& origin, V3& direction)
{
    /* shader.rgen:32-38 */
    const float pcx = (float)x + jx, pcy = (float)y + jy;
    const float inx = pcx / (float)width, iny = pcy / (float)height;
    const float ux = inx * 2.0f - 1.0f, uy = iny * 2.0f - 1.0f;
    const V3 target = mat4MulVec(u.projInverse, ux, uy, 1.0f, 1.0f);
    const V3 tn = normalize(target);
    direction = mat4MulVec(u.viewInverse, tn.x, tn.y, tn.z, 0.0f);
    origin = mat4MulVec(u.viewInverse, 0.0f, 0.0f, 0.0f, 1.0f);
}

inline uint8_t unorm8(float x)
{
    return (uint8_t)(clamp01(x) * 255.0f + 0.5f);
}
} // namespace

extern "C" {

OrcScene* orc_scene_create(const float* vertices, uint64_t vertexCount, const uint32_t* indices, uint64_t indexCount,
                           const OrcGeometryRange* ranges, const OrcSubmeshInfo* rows, uint32_t geometryCount,
                           const OrcImage* images, uint32_t imageCount, const float* instances, uint32_t instanceCount)
{
    OrcScene* s = new OrcScene();
    s->vertices.assign(vertices, vertices + vertexCount * 16);
    s->indices.assign(indices, indices + indexCount);
    s->ranges.assign(ranges, ranges + geometryCount);
    s->rows.assign(rows, rows + geometryCount);
    s->images.resize(imageCount);
    for (uint32_t i = 0; i < imageCount; ++i)
    {
        s->images[i].w = images[i].width;
        s->images[i].h = images[i].height;
        s->images[i].px.assign(images[i].rgba8, images[i].rgba8 + (size_t)images[i].width * images[i].height * 4);
    }
    s->instances.resize(instanceCount);
    for (uint32_t i = 0; i < instanceCount; ++i)
    {
        std::memcpy(s->instances[i].o2w, instances + 12 * (size_t)i, sizeof(float) * 12);
        affineInverse(s->instances[i].o2w, s->instances[i].w2o);
    }
    s->buildTriangles();
    s->buildBvh();
    return s;
}

void orc_scene_set_instance_masks(OrcScene* scene, const uint8_t* masks, uint32_t count)
{
    for (uint32_t i = 0; i < count && i < scene->instances.size(); ++i)
    {
        scene->instances[i].mask = masks[i];
    }
}

void orc_scene_destroy(OrcScene* scene)
{
    delete scene;
}

void orc_scene_update_vertices(OrcScene* scene, const float* vertices, uint64_t vertexCount)
{
    scene->vertices.assign(vertices, vertices + vertexCount * 16);
    scene->buildTriangles();
    scene->buildBvh();
}

double orc_scene_build_seconds(const OrcScene* scene)
{
    return scene->buildSeconds;
}

void orc_halton23(uint32_t s, float out[2])
{
    /* 1-based so the degenerate (0,0) corner is skipped */
    out[0] = radicalInverse(s + 1, 2);
    out[1] = radicalInverse(s + 1, 3);
}

void orc_primary_ray(const OrcUniforms* uniforms, uint32_t width, uint32_t height, uint32_t x, uint32_t y, float jx,
                     float jy, OrcRay* ray)
{
    V3 o, d;
    primaryRay(*uniforms, width, height, x, y, jx, jy, o, d);
    ray->origin[0] = o.x, ray->origin[1] = o.y, ray->origin[2] = o.z;
    ray->direction[0] = d.x, ray->direction[1] = d.y, ray->direction[2] = d.z;
    ray->tmin = 0.001f;
    ray->tmax = 1000.0f;
}

void orc_affine_inverse(const float m[12], float inv[12])
{
    affineInverse(m, inv);
}

void orc_scene_generate_mips(OrcScene* scene)
{
    scene->generateMips();
}

void orc_scene_set_texture_lod(OrcScene* scene, int mode)
{
    scene->lodMode = (mode == 1 && scene->mips.size() == scene->images.size()) ? 1 : 0;
}

uint32_t orc_texture_level_count(const OrcScene* scene, uint32_t index)
{
    return scene->mips.size() == scene->images.size() ? (uint32_t)scene->mips[index].size() + 1u : 1u;
}

const uint8_t* orc_texture_level(const OrcScene* scene, uint32_t index, uint32_t level, uint32_t* width, uint32_t* height)
{
    if (level == 0)
    {
        *width = scene->images[index].w, *height = scene->images[index].h;
        return scene->images[index].px.data();
    }
    const MipLevel& m = scene->mips[index][level - 1];
    *width = m.w, *height = m.h;
    return m.px.data();
}

void orc_sample_texture_lod(const OrcScene* scene, uint32_t index, float u, float v, float footprint2, float rgba[4])
{
    scene->sampleLod(index, u, v, footprint2, rgba);
}

void orc_sample_texture(const OrcScene* scene, uint32_t index, float u, float v, float rgba[4])
{
    scene->sample(index, u, v, rgba);
}

void orc_render(OrcScene* scene, const OrcUniforms* uniforms, uint32_t width, uint32_t height, uint32_t spp, int mode,
                int threads, uint32_t x0, uint32_t y0, uint32_t x1, uint32_t y1, uint32_t strideX, uint32_t strideY,
                float* rgb32f, uint8_t* bgra8, uint32_t* triangleId, uint8_t* shadowMask, OrcStats* stats)
{
    const auto tStart = std::chrono::steady_clock::now();
    if (threads <= 0)
    {
        threads = (int)std::max(1u, std::thread::hardware_concurrency());
    }
    strideX = std::max(strideX, 1u), strideY = std::max(strideY, 1u);
    spp = std::max(spp, 1u);
    x1 = std::min(x1, width), y1 = std::min(y1, height);
    const uint32_t rows = y1 > y0 ? (y1 - y0 + strideY - 1) / strideY : 0;
    std::atomic<uint32_t> nextRow{0};
    std::vector<Counters> perThread((size_t)threads);
    const OrcScene& s = *scene;
    const OrcUniforms& u = *uniforms;

    auto worker = [&](int tid) {
        Counters& c = perThread[(size_t)tid];
        for (;;)
        {
            const uint32_t r = nextRow.fetch_add(1);
            if (r >= rows)
            {
                break;
            }
            const uint32_t y = y0 + r * strideY;
            for (uint32_t x = x0; x < x1; x += strideX)
            {
                V3 acc = mk(0, 0, 0);
                uint32_t aovTri = 0xFFFFFFFFu;
                uint8_t aovMask = 0;
                for (uint32_t sIdx = 0; sIdx < spp; ++sIdx)
                {
                    float jx = 0.5f, jy = 0.5f; /* rgen:32 */
                    if (spp > 1)
                    {
                        float h[2];
                        orc_halton23(sIdx, h);
                        jx = h[0], jy = h[1];
                    }
                    V3 origin, direction;
                    primaryRay(u, width, height, x, y, jx, jy, origin, direction);
                    V3 finalHitValue = mk(0, 0, 0);
                    Payload p;
                    p.hitValue = mk(0, 0, 0), p.attenuation = 1.0f, p.depth = 0, p.done = 1; /* rgen:44-47 */
                    p.rayOrigin = mk(0, 0, 0), p.rayDir = mk(0, 0, 0);
                    const int maxDepth = 2;
                    /* ray cones: spread angle 2 tan(fovy/2) / height, tan(fovy/2) = |projInverse[1][1]|; the cone
                     * width at a hit is the spread angle times the distance travelled along the whole path */
                    const float spreadAngle = (2.0f * std::fabs(u.projInverse[5])) / (float)height;
                    float travelled = 0.0f;
                    for (;;) /* rgen:49-74 */
                    {
                        if (p.depth == 0)
                        {
                            ++c.raysPrimary;
                        }
                        else
                        {
                            ++c.raysReflection;
                        }
                        const Hit hit = s.trace(origin, direction, 0.001f, 1000.0f, false, mode, c);
                        uint8_t mask = 0;
                        if (hit.valid)
                        {
                            travelled = p.depth == 0 ? hit.t : travelled + hit.t;
                            mask = closestHit(s, u, hit, direction, p, mode, c, spreadAngle * travelled);
                        }
                        else
                        {
                            p.hitValue = mk(0.8f, 0.8f, 1.0f); /* rmiss:17 */
                        }
                        if (p.depth == 0 && sIdx == 0)
                        {
                            aovTri = hit.valid ? ((hit.instance << 24) + hit.globalTri) : 0xFFFFFFFFu;
                            aovMask = mask;
                        }
                        finalHitValue = finalHitValue + p.hitValue;
                        ++p.depth;
                        if (p.done == 1 || p.depth >= maxDepth)
                        {
                            break;
                        }
                        origin = p.rayOrigin;
                        direction = p.rayDir;
                        p.done = 1;
                    }
                    acc = acc + finalHitValue;
                }
                const V3 out = spp > 1 ? acc / (float)spp : acc;
                const size_t pix = (size_t)y * width + x;
                if (rgb32f)
                {
                    rgb32f[pix * 3 + 0] = out.x, rgb32f[pix * 3 + 1] = out.y, rgb32f[pix * 3 + 2] = out.z;
                }
                if (bgra8)
                {
                    bgra8[pix * 4 + 0] = unorm8(out.z), bgra8[pix * 4 + 1] = unorm8(out.y);
                    bgra8[pix * 4 + 2] = unorm8(out.x), bgra8[pix * 4 + 3] = 255;
                }
                if (triangleId)
                {
                    triangleId[pix] = aovTri;
                }
                if (shadowMask)
                {
                    shadowMask[pix] = aovMask;
                }
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t)
    {
        pool.emplace_back(worker, t);
    }
    worker(0);
    for (auto& t : pool)
    {
        t.join();
    }
    if (stats)
    {
        Counters total;
        for (const Counters& c : perThread)
        {
            total.add(c);
        }
        stats->raysPrimary = total.raysPrimary, stats->raysShadow = total.raysShadow;
        stats->raysReflection = total.raysReflection;
        stats->nodesVisited = total.nodes, stats->trianglesTested = total.tris;
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - tStart).count();
    }
}

/*
 * The reference's second renderer: Rasterizer::render (src/Rasterizer.cpp:100-170) with shaders/shader.vert and
 * shaders/shader.frag, restated as a scalar z-buffer rasteriser.
 *   vertex shader      gl_Position = wvp * (position, 1), uv passed through                      (shader.vert:18-23)
 *   fixed function     viewport (0,0,W,H), depth [0,1]; back faces culled, front = counter-clockwise; 8x MSAA at the
 *                      Vulkan standard sample locations, no sample shading (one fragment per pixel and triangle, attributes
 *                      at the pixel centre); top-left rule; depth test LESS with writes on D24; no blending
 *                                                                                                (Rasterizer.cpp:779-829)
 *   fragment shader    colour = texture(baseColor, uv), discard when alpha < 0.1                 (shader.frag:13-22)
 *   sampler            LINEAR / REPEAT / mip LINEAR over the whole chain, implicit derivatives   (Rasterizer.cpp:484-498)
 *   clears, resolve    colour (0,0,0.2,1), depth 1 (:119-120); 8 samples averaged into B8G8R8A8
 * Triangles are drawn in order (geometry after geometry, :140-146), so at equal depth the earlier triangle keeps the
 * sample.  Numerical contract shared with the product: edge functions in 2D homogeneous
 * coordinates (Olano & Greer 1997), planes evaluated as (a*x + b*y) + c, FP32, no FMA contraction.
 */
namespace
{
struct RasterPlanes
{
    float ea[3], eb[3], ec[3]; /* edge functions, >= 0 inside */
    float z[3], u[3], v[3], s[3];
    uint32_t texture;
    int x0, y0, x1, y1;
    bool drawn;
};

inline float evalPlane(const float p[3], float x, float y)
{
    return (p[0] * x + p[1] * y) + p[2];
}

const int kSampleX16[8] = {9, 7, 13, 5, 3, 1, 11, 15}; /* Vulkan standard sample locations, 8 samples, in 1/16 pixel */
const int kSampleY16[8] = {5, 11, 9, 3, 13, 7, 15, 1};

RasterPlanes rasterPlanes(const float* wvp, const float* a, const float* b, const float* c, uint32_t texture, uint32_t width, uint32_t height)
{
    RasterPlanes r;
    r.drawn = false, r.texture = texture;
    const float* vtx[3] = {a, b, c};
    float cx[3], cy[3], cz[3], cw[3];
    for (int i = 0; i < 3; ++i)
    {
        const float px = vtx[i][0], py = vtx[i][1], pz = vtx[i][2];
        cx[i] = ((wvp[0] * px + wvp[4] * py) + wvp[8] * pz) + wvp[12];
        cy[i] = ((wvp[1] * px + wvp[5] * py) + wvp[9] * pz) + wvp[13];
        cz[i] = ((wvp[2] * px + wvp[6] * py) + wvp[10] * pz) + wvp[14];
        cw[i] = ((wvp[3] * px + wvp[7] * py) + wvp[11] * pz) + wvp[15];
    }
    auto all = [&](auto pred) { return pred(0) && pred(1) && pred(2); };
    if (all([&](int i) { return cx[i] < -cw[i]; }) || all([&](int i) { return cx[i] > cw[i]; }) || all([&](int i) { return cy[i] < -cw[i]; }) ||
        all([&](int i) { return cy[i] > cw[i]; }) || all([&](int i) { return cz[i] < 0.0f; }) || all([&](int i) { return cz[i] > cw[i]; }))
    {
        return r; /* outside one frustum plane */
    }
    const float halfW = 0.5f * (float)width, halfH = 0.5f * (float)height;
    V3 h[3]; /* (x_fb * w, y_fb * w, w) */
    for (int i = 0; i < 3; ++i)
    {
        h[i] = mk((cx[i] + cw[i]) * halfW, (cy[i] + cw[i]) * halfH, cw[i]);
    }
    const V3 n[3] = {cross(h[2], h[1]), cross(h[0], h[2]), cross(h[1], h[0])};
    const float det = dot(h[0], n[0]);
    if (!(det > 0.0f) || !std::isfinite(det))
    {
        return r; /* back-facing (clockwise in the y-down framebuffer) or degenerate */
    }
    const float tu[3] = {a[8], b[8], c[8]}, tv[3] = {a[9], b[9], c[9]};
    for (int k = 0; k < 3; ++k)
    {
        const float nk[3] = {comp(n[0], k), comp(n[1], k), comp(n[2], k)};
        if (k == 0)
        {
            r.ea[0] = nk[0], r.ea[1] = nk[1], r.ea[2] = nk[2];
        }
        else if (k == 1)
        {
            r.eb[0] = nk[0], r.eb[1] = nk[1], r.eb[2] = nk[2];
        }
        else
        {
            r.ec[0] = nk[0], r.ec[1] = nk[1], r.ec[2] = nk[2];
        }
        r.z[k] = ((cz[0] * nk[0] + cz[1] * nk[1]) + cz[2] * nk[2]) / det;
        r.u[k] = (tu[0] * nk[0] + tu[1] * nk[1]) + tu[2] * nk[2];
        r.v[k] = (tv[0] * nk[0] + tv[1] * nk[1]) + tv[2] * nk[2];
        r.s[k] = (nk[0] + nk[1]) + nk[2];
    }
    r.x0 = 0, r.y0 = 0, r.x1 = (int)width - 1, r.y1 = (int)height - 1;
    if (cw[0] > 0.0f && cw[1] > 0.0f && cw[2] > 0.0f)
    {
        float lox = INFINITY, hix = -INFINITY, loy = INFINITY, hiy = -INFINITY;
        for (int i = 0; i < 3; ++i)
        {
            const float fx = h[i].x / h[i].z, fy = h[i].y / h[i].z;
            lox = std::fmin(lox, fx), hix = std::fmax(hix, fx), loy = std::fmin(loy, fy), hiy = std::fmax(hiy, fy);
        }
        if (!(hix >= 0.0f) || !(lox < (float)width) || !(hiy >= 0.0f) || !(loy < (float)height))
        {
            return r;
        }
        r.x0 = (int)std::fmax(std::floor(lox), 0.0f), r.y0 = (int)std::fmax(std::floor(loy), 0.0f);
        r.x1 = (int)std::fmin(std::floor(hix), (float)(width - 1)), r.y1 = (int)std::fmin(std::floor(hiy), (float)(height - 1));
    }
    r.drawn = true;
    return r;
}
} // namespace

/* Fragment colour of triangle planes `t` at pixel (px,py); false = discarded. */
static bool rasterFragmentColour(const OrcScene& s, const RasterPlanes& t, int px, int py, float rgba[4])
{
    const float x = (float)px + 0.5f, y = (float)py + 0.5f;
    const float U = evalPlane(t.u, x, y), V = evalPlane(t.v, x, y), S = evalPlane(t.s, x, y);
    const float u = U / S, v = V / S;
    const float S2 = S * S;
    /* implicit derivatives, analytic: d(U/S)/dx = (U_x S - U S_x) / S^2 */
    const float dudx = (t.u[0] * S - U * t.s[0]) / S2, dudy = (t.u[1] * S - U * t.s[1]) / S2;
    const float dvdx = (t.v[0] * S - V * t.s[0]) / S2, dvdy = (t.v[1] * S - V * t.s[1]) / S2;
    const Image& img = s.images[t.texture];
    const float w = (float)img.w, h = (float)img.h;
    const float rx2 = (dudx * w) * (dudx * w) + (dvdx * h) * (dvdx * h);
    const float ry2 = (dudy * w) * (dudy * w) + (dvdy * h) * (dvdy * h);
    s.sampleTexels2(t.texture, u, v, std::fmax(rx2, ry2), rgba);
    return !(rgba[3] < 0.1f);
}

void orc_rasterize(OrcScene* scene, const float* wvp, uint32_t width, uint32_t height, int threads, uint8_t* bgra8, uint32_t* triangleId)
{
    const OrcScene& s = *scene;
    if (threads <= 0)
    {
        threads = (int)std::max(1u, std::thread::hardware_concurrency());
    }
    /* vertex shader + set-up for every triangle, in draw order */
    std::vector<RasterPlanes> planes(s.tris.size());
    {
        size_t t = 0;
        for (size_t g = 0; g < s.ranges.size(); ++g)
        {
            const uint64_t first = s.ranges[g].indexByteOffset / 4;
            const int32_t baseColor = s.rows[g].baseColorTextureIndex;
            const uint32_t texture = (baseColor < 0 || (size_t)baseColor >= s.images.size()) ? 0u : (uint32_t)baseColor;
            for (uint32_t p = 0; p < s.ranges[g].primitiveCount; ++p, ++t)
            {
                const float* a = &s.vertices[(size_t)s.indices[first + 3 * p + 0] * 16];
                const float* b = &s.vertices[(size_t)s.indices[first + 3 * p + 1] * 16];
                const float* c = &s.vertices[(size_t)s.indices[first + 3 * p + 2] * 16];
                planes[t] = rasterPlanes(wvp, a, b, c, texture, width, height);
            }
        }
    }
    const size_t sampleCount = (size_t)width * height * 8;
    std::vector<uint32_t> depth(sampleCount, 0xFFFFFFu); /* cleared to 1.0 (src/Rasterizer.cpp:120) */
    std::vector<uint32_t> colour(sampleCount, 0u | (0u << 8) | (51u << 16) | (255u << 24)); /* (0,0,0.2,1) as UNORM8 */
    std::vector<uint32_t> owner(sampleCount, 0xFFFFFFFFu);
    auto band = [&](int tid) {
        for (size_t t = 0; t < planes.size(); ++t)
        {
            const RasterPlanes& tp = planes[t];
            if (!tp.drawn)
            {
                continue;
            }
            for (int py = tp.y0; py <= tp.y1; ++py)
            {
                if (py % threads != tid)
                {
                    continue;
                }
                for (int px = tp.x0; px <= tp.x1; ++px)
                {
                    uint32_t covered = 0;
                    for (int k = 0; k < 8; ++k)
                    {
                        const float x = (float)px + (float)kSampleX16[k] * 0.0625f, y = (float)py + (float)kSampleY16[k] * 0.0625f;
                        bool inside = true;
                        for (int e = 0; e < 3; ++e)
                        {
                            const float val = (tp.ea[e] * x + tp.eb[e] * y) + tp.ec[e];
                            const bool topLeft = tp.ea[e] > 0.0f || (tp.ea[e] == 0.0f && tp.eb[e] > 0.0f);
                            inside = inside && (val > 0.0f || (val == 0.0f && topLeft));
                        }
                        covered |= inside ? 1u << k : 0u;
                    }
                    if (!covered)
                    {
                        continue;
                    }
                    float rgba[4];
                    if (!rasterFragmentColour(s, tp, px, py, rgba))
                    {
                        continue; /* discard */
                    }
                    uint32_t q = 0;
                    for (int ch = 0; ch < 4; ++ch)
                    {
                        q |= (uint32_t)unorm8(rgba[ch]) << (8 * ch);
                    }
                    for (int k = 0; k < 8; ++k)
                    {
                        if (!((covered >> k) & 1u))
                        {
                            continue;
                        }
                        const float x = (float)px + (float)kSampleX16[k] * 0.0625f, y = (float)py + (float)kSampleY16[k] * 0.0625f;
                        const float z = evalPlane(tp.z, x, y);
                        if (!(z >= 0.0f && z <= 1.0f))
                        {
                            continue; /* near / far clip */
                        }
                        const uint32_t d24 = (uint32_t)(z * 16777215.0f + 0.5f);
                        const size_t idx = ((size_t)py * width + px) * 8 + k;
                        if (d24 < depth[idx]) /* VK_COMPARE_OP_LESS */
                        {
                            depth[idx] = d24, colour[idx] = q, owner[idx] = (uint32_t)t;
                        }
                    }
                }
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t)
    {
        pool.emplace_back(band, t);
    }
    band(0);
    for (auto& t : pool)
    {
        t.join();
    }
    for (size_t pix = 0; pix < (size_t)width * height; ++pix)
    {
        uint32_t sum[4] = {0, 0, 0, 0};
        for (int k = 0; k < 8; ++k)
        {
            for (int ch = 0; ch < 4; ++ch)
            {
                sum[ch] += (colour[pix * 8 + k] >> (8 * ch)) & 0xFFu;
            }
        }
        if (bgra8)
        {
            const int order[4] = {2, 1, 0, 3}; /* B, G, R, A */
            for (int ch = 0; ch < 4; ++ch)
            {
                bgra8[pix * 4 + ch] = (uint8_t)((float)sum[order[ch]] / 8.0f + 0.5f);
            }
        }
        if (triangleId)
        {
            triangleId[pix] = owner[pix * 8];
        }
    }
}

void orc_trace_rays(OrcScene* scene, const OrcRay* rays, uint64_t count, int anyHit, int mode, int threads,
                    OrcHit* hits, OrcStats* stats)
{
    const auto tStart = std::chrono::steady_clock::now();
    if (threads <= 0)
    {
        threads = (int)std::max(1u, std::thread::hardware_concurrency());
    }
    std::atomic<uint64_t> next{0};
    std::vector<Counters> perThread((size_t)threads);
    const uint64_t chunk = 256;
    auto worker = [&](int tid) {
        Counters& c = perThread[(size_t)tid];
        for (;;)
        {
            const uint64_t b = next.fetch_add(chunk);
            if (b >= count)
            {
                break;
            }
            const uint64_t e = std::min(count, b + chunk);
            for (uint64_t i = b; i < e; ++i)
            {
                const OrcRay& r = rays[i];
                const Hit h = scene->trace(mk(r.origin[0], r.origin[1], r.origin[2]),
                                           mk(r.direction[0], r.direction[1], r.direction[2]), r.tmin, r.tmax,
                                           anyHit != 0, mode, c);
                OrcHit& o = hits[i];
                o.t = h.t, o.b1 = h.b1, o.b2 = h.b2;
                o.instance = h.valid ? h.instance : 0xFFFFFFFFu;
                o.geometry = h.valid ? h.geometry : 0xFFFFFFFFu;
                o.primitive = h.valid ? h.primitive : 0xFFFFFFFFu;
                if (!h.valid)
                {
                    o.t = 0.0f, o.b1 = 0.0f, o.b2 = 0.0f;
                }
                if (anyHit)
                {
                    ++c.raysShadow;
                }
                else
                {
                    ++c.raysPrimary;
                }
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < threads; ++t)
    {
        pool.emplace_back(worker, t);
    }
    worker(0);
    for (auto& t : pool)
    {
        t.join();
    }
    if (stats)
    {
        Counters total;
        for (const Counters& c : perThread)
        {
            total.add(c);
        }
        stats->raysPrimary = total.raysPrimary, stats->raysShadow = total.raysShadow;
        stats->raysReflection = total.raysReflection;
        stats->nodesVisited = total.nodes, stats->trianglesTested = total.tris;
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - tStart).count();
    }
}

} // extern "C"
