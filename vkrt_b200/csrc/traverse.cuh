// traverse.cuh — per-ray traversal of the compressed 8-wide BVH and the watertight ray/triangle test.
//
// Replaces what the Vulkan driver does behind traceRayEXT (shaders/shader.rgen:51-62 closest hit,
// shaders/shader.rchit:132-143 any hit) over the BLAS/TLAS of src/Raytracer.cpp:1027-1283.
//
// Node format and octant-ordered traversal follow Ylitie, Karras, Laine 2017 (compressed wide BVH);
// the triangle test is Woop, Benthin, Wald 2013 (watertight), no culling, exclusive (tmin, tmax).
#pragma once

#include "common.cuh"
#include "device_math.cuh"

#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace vkrt
{

constexpr int kStackSize = 48; // group entries.  A BLAS pushes at most one per level, the top level one per level plus two
                               // when it enters an instance: the builders refuse hierarchies that could exceed this
constexpr int kMaxBlasLevels = kStackSize;

// Traversal stack of one ray.  Default: kStackSize entries of per-thread local memory (they live in L1).  The A/B build
// -DVKRT_SMEM_STACK keeps the kSmemStack entries nearest the top of the tree walk in shared memory (interleaved by thread, no
// bank conflicts) and only deeper ones in local memory: the "short stack" variant, measured in DESIGN.md section 5.2.
#if defined(__CUDACC__) && defined(VKRT_SMEM_STACK)
constexpr int kSmemStack = 8;
struct RayStack
{
    uint2* shared; // this thread's column: entry i at shared[i * blockDim.x]
    uint2 deep[kStackSize - kSmemStack];
    VKRT_HD int stride() const
    {
#if defined(__CUDA_ARCH__)
        return (int)blockDim.x;
#else
        return 1;
#endif
    }
    VKRT_HD void set(int i, uint2 v)
    {
        if (i < kSmemStack)
        {
            shared[i * stride()] = v;
        }
        else
        {
            deep[i - kSmemStack] = v;
        }
    }
    VKRT_HD uint2 get(int i) const { return i < kSmemStack ? shared[i * stride()] : deep[i - kSmemStack]; }
};
#else
struct RayStack
{
    uint2 e[kStackSize];
    VKRT_HD void set(int i, uint2 v) { e[i] = v; }
    VKRT_HD uint2 get(int i) const { return e[i]; }
};
#endif

struct TraceCounters
{
    uint32_t nodes = 0, tris = 0, overflow = 0;
};

// Watertight per-ray constants.  The shear of Woop et al. 2013 is  A' = (A[kx] - Sx*A[kz], A[ky] - Sy*A[kz], Sz*A[kz]).
// Instead of selecting vertex components by the runtime axis indices (two SEL per component on the saturated alu
// pipe), each output is evaluated as a dot product with a per-ray vector whose entries are 1 at kx (ky), -Sx (-Sy) at
// kz and 0 elsewhere (shearDot below): the products by 1 and 0 are exact and adding +-0 is exact, so only the S*a term
// rounds, and the work moves to the fma pipe.
struct RayPre
{
    F3 mx, my, mz;
};

// The shear of the watertight test as ONE fixed sequence of fused operations:  fma(a.x, m.x, fma(a.y, m.y, a.z * m.z)).
// It is part of the numerical contract (the oracle evaluates exactly this expression): a function of the vertex and the
// ray only, so the two triangles that share an edge still see identical sheared vertices, which is what the
// watertightness argument of Woop et al. needs; the edge functions U, V, W stay unfused (a fused a*b - c*d is not
// antisymmetric).  3 instructions instead of 5 per dot product, 20 fewer per triangle test.
VKRT_HD float shearDot(F3 a, F3 m) { return fmaExplicit(a.x, m.x, fmaExplicit(a.y, m.y, a.z * m.z)); }

VKRT_HD void setComp3(F3& v, int k, float value)
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

VKRT_HD RayPre makeRayPre(F3 d)
{
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int kz = 0;
    float best = ax;
    if (ay > best)
    {
        kz = 1;
        best = ay;
    }
    if (az > best)
    {
        kz = 2;
    }
    int kx = kz + 1 == 3 ? 0 : kz + 1;
    int ky = kx + 1 == 3 ? 0 : kx + 1;
    if (comp3(d, kz) < 0.0f)
    {
        const int t = kx;
        kx = ky;
        ky = t;
    }
    const float dz = comp3(d, kz);
    const float Sx = comp3(d, kx) / dz, Sy = comp3(d, ky) / dz, Sz = 1.0f / dz;
    RayPre r;
    r.mx = r.my = r.mz = f3(0.0f, 0.0f, 0.0f);
    setComp3(r.mx, kx, 1.0f);
    setComp3(r.mx, kz, -Sx);
    setComp3(r.my, ky, 1.0f);
    setComp3(r.my, kz, -Sy);
    setComp3(r.mz, kz, Sz);
    return r;
}

// Watertight test. Accepts tmin < t < tmax, or t == tmax when `tieOk` (an earlier hit at exactly tmax with a
// higher id).  Outputs t and the barycentric weights of vertices 1 and 2.
VKRT_HD bool intersectTriangle(F3 v0, F3 v1, F3 v2, F3 o, const RayPre& rp, float tmin, float tmax, bool tieOk, float& tOut,
                               float& b1, float& b2)
{
    const F3 A = v0 - o, B = v1 - o, C = v2 - o;
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
    const float T = fmaExplicit(W, Cz, fmaExplicit(V, Bz, U * Az));
    const float rcp = 1.0f / det;
    const float t = T * rcp;
    if (!(t > tmin))
    {
        return false;
    }
    if (!(t < tmax) && !(tieOk && t == tmax))
    {
        return false;
    }
    tOut = t;
    b1 = V * rcp;
    b2 = W * rcp;
    return true;
}

VKRT_HD uint32_t highestBit(uint32_t v) // index of the most significant set bit, v != 0
{
#if defined(__CUDA_ARCH__)
    uint32_t bit;
    asm("bfind.u32 %0, %1;" : "=r"(bit) : "r"(v)); // FLO.U32 already is the bit index; 31 - clz costs two more ops
    return bit;
#else
    return 31u - (uint32_t)__builtin_clz(v);
#endif
}
VKRT_HD uint32_t popCount(uint32_t v)
{
#if defined(__CUDA_ARCH__)
    return (uint32_t)__popc(v);
#else
    return (uint32_t)__builtin_popcount(v);
#endif
}

// Box-test direction reciprocal.  Zero components give +-inf on purpose: with IEEE arithmetic every slab product is
// then +-inf or NaN, and since fmaxf/fminf drop NaN operands an axis the ray runs parallel to either culls the box
// correctly (ray outside the slab on the node-origin side) or is ignored (conservative).  Clamping to a large
// finite value instead would wrongly cull rays that lie exactly in a box face.
VKRT_HD float safeRcp(float d)
{
    return 1.0f / d;
}

struct NodeWords // one Node8 fetched as five 16-byte vectors
{
    float4 n0;
    uint4 n1, n2, n3, n4;
};

VKRT_HD NodeWords loadNode(const Node8* nodes, uint32_t index)
{
    NodeWords w;
    const float4* p = reinterpret_cast<const float4*>(nodes + index);
#if defined(__CUDA_ARCH__)
    w.n0 = __ldg(p);
    w.n1 = __ldg(reinterpret_cast<const uint4*>(p + 1));
    w.n2 = __ldg(reinterpret_cast<const uint4*>(p + 2));
    w.n3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
    w.n4 = __ldg(reinterpret_cast<const uint4*>(p + 4));
#else
    w.n0 = p[0];
    memcpy(&w.n1, p + 1, 16);
    memcpy(&w.n2, p + 2, 16);
    memcpy(&w.n3, p + 3, 16);
    memcpy(&w.n4, p + 4, 16);
#endif
    return w;
}

VKRT_HD uint32_t byteOf(uint32_t word, int i) { return (word >> (8 * i)) & 0xFFu; }

// Byte i of `word` dropped into mantissa bits 8..15 of 1.0f: the float 1 + q * 2^-15.  One PRMT on the device (alu
// pipe) instead of an I2F conversion; the 2^15 is folded into the per-node scale and the 1 into the per-node bias.
// `one` is the bit pattern of 1.0f, passed in as a kernel parameter (SceneDev::one) so that ptxas cannot fold it: PRMT
// then takes the selector as its immediate operand; with a literal it re-materialises a selector register per PRMT.
template<int I>
VKRT_HD float byteToUnitFloat(uint32_t word, uint32_t one)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(word, one, 0x7604u | (I << 4)));
#else
    return uintAsFloat(one | (((word >> (8 * I)) & 0xFFu) << 8));
#endif
}
// Per-ray padded clamp values for the box tests.  The triangle test computes its own t with its own roundings, so a
// box is only culled against [tmin, tmax] widened by 2^-18 relative: box culling must never reject a watertight hit.
constexpr float kBoxRelPad = 3.8146973e-06f; // 2^-18
VKRT_HD float padDown(float t) { return fmaExplicit(-fabsf(t), kBoxRelPad, t); }
VKRT_HD float padUp(float t) { return fmaExplicit(fabsf(t), kBoxRelPad, t); }

// Slab test of child J (byte J & 3 of its half's plane words); sets bit J of `hitSlots` when the padded interval is
// not empty.
template<int J>
VKRT_HD void testChild(uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz, float ax, float ay, float az,
                       float bnx, float bny, float bnz, float bfx, float bfy, float bfz, float tminP, float tmaxP, uint32_t one,
                       uint32_t& hitSlots)
{
    constexpr int I = J & 3;
    const float tnx = fmaExplicit(byteToUnitFloat<I>(nx, one), ax, bnx), tfx = fmaExplicit(byteToUnitFloat<I>(fx, one), ax, bfx);
    const float tny = fmaExplicit(byteToUnitFloat<I>(ny, one), ay, bny), tfy = fmaExplicit(byteToUnitFloat<I>(fy, one), ay, bfy);
    const float tnz = fmaExplicit(byteToUnitFloat<I>(nz, one), az, bnz), tfz = fmaExplicit(byteToUnitFloat<I>(fz, one), az, bfz);
    const float tn = fmaxf(fmaxf(fmaxf(tnx, tny), tnz), tminP);
    const float tf = fminf(fminf(fminf(tfx, tfy), tfz), tmaxP);
    if (tn <= tf)
    {
        hitSlots |= 1u << J;
    }
}

// Moves bit s of an 8-bit slot mask to bit s ^ octinv (three conditional swap stages, written as variable shifts so
// that no select is needed).  Children are stored in octant order, so after this the highest bit is the nearest child.
VKRT_HD uint32_t permuteSlots(uint32_t x, uint32_t octinv)
{
    const uint32_t s4 = octinv & 4u, s2 = octinv & 2u, s1 = octinv & 1u;
    x = ((x * 0x101u) >> s4) & 0xFFu;
    x = ((x << s2) & 0xCCu) | ((x >> s2) & 0x33u);
    x = ((x << s1) & 0xAAu) | ((x >> s1) & 0x55u);
    return x;
}

// Bit s of an 8-bit slot mask -> bits 3s, 3s+1, 3s+2 (the layout of Node8::leafBits).
VKRT_HD uint32_t spreadSlots(uint32_t x)
{
    x = (x * 0x101u) & 0xF00Fu;
    x = (x * 0x11u) & 0xC30C3u;
    x = (x * 0x5u) & 0x249249u;
    return x * 7u;
}

#if defined(__CUDACC__)
// Device copies of the two functions above as shared-memory tables (3 KB per block): ~20 alu instructions per node
// test become two LDS.  Every kernel that traverses calls initSlotTables() first (all threads of the block).
__shared__ uint32_t sSpreadSlots[256];
__shared__ uint8_t sPermuteSlots[8 * 256]; // [octinv][mask]
__device__ __forceinline__ void initSlotTables()
{
    for (uint32_t i = threadIdx.x; i < 8u * 256u; i += blockDim.x)
    {
        sPermuteSlots[i] = (uint8_t)permuteSlots(i & 255u, i >> 8);
    }
    for (uint32_t i = threadIdx.x; i < 256u; i += blockDim.x)
    {
        sSpreadSlots[i] = spreadSlots(i);
    }
    __syncthreads();
}
#endif

// Intersects the ray with the 8 quantised child boxes of a node.  Returns x = the internal children that are hit,
// as an 8-bit mask in traversal priority order (nearest octant-order child in the highest bit), y = the leafBits of
// the leaf children whose boxes are hit.
//   idir = safeRcp(d), octant bit k set when d_k < 0, tminP/tmaxP = padDown(tmin)/padUp(tmax)
//
// Plane distance of quantised coordinate q on axis x:  t = (px + q*sx - ox) * idir.x.  With m = 1 + q*2^-15 (the PRMT
// float above) this is  t = m*ax + b  with  ax = 2^15*sx*idir.x  and  b = (px-ox)*idir.x - ax : one FFMA per plane.
// Rounding b costs at most 2^-24*(|b|+|ax|) <= 2^-9 of one quantisation step of t, so the near planes are pulled in
// and the far planes pushed out by 2^-7 step (bn/bf below): the test stays conservative, and flat boxes (qlo == qhi)
// keep tn < tf.  NaN planes (0*inf for axis-parallel rays) are dropped by fmaxf/fminf, i.e. that axis never culls.
// Empty slots carry a zero box that a ray may well hit: they drop out because neither imask nor leafBits names them.
#if defined(__CUDA_ARCH__) && !defined(VKRT_NODE_TEST_FP32)
#define VKRT_NODE_TEST_HALF2 1
// ------------------------------------------------------------------------------------------------------------
// Packed-half node test (device only): two children per instruction.
//   * dequantisation: one PRMT drops TWO plane bytes into the low mantissa bytes of the half pair (0x6400 | q) =
//     (1024 + q0, 1024 + q1), exact;
//   * plane distances: one HFMA2 per plane pair,  t' = (1024 + q) * a + b',  with every t measured from `tRef`, the
//     distance at which the ray enters the node's own box (clamped to the ray interval): the values that decide a
//     child's test then lie within the node's own extent along the ray (|t'| <= ~256 |a|), where a half's 11 bits
//     resolve 1/8 of a quantisation step, and  b' = (p - o) * idir - 1024 a - tRef  is formed in FP32;
//   * interval: VHMNMX (three-input packed max/min) + HMNMX2 with the clamp, one HSET2 for both children.
// Rounding, in quantisation steps |a| of the axis: truncating a to a half's 11 bits moves plane q by q (a - a_r) <= 0.25;
// b' -> half <= 2^-11 (1279 + slack) = 0.63; the HFMA2 result <= 0.13 for planes inside the node: 1.01 in total, so near
// planes are pulled in and far planes pushed out by kHalfSlackSteps = 1.125 steps (the FP32 test uses 2^-7 step) and the
// clamps by 2^-10 relative.  Overflow gives +-inf with the sign of the exact value or NaN, which the min/max drop (that
// axis does not cull).  Only nodes whose steps a_k all lie in [2^-14, 32] take this path (the builder keeps the three
// scales of a node within 2^8 of each other, builder.cuh::quantExponents); any other node - rays almost parallel to a box
// face, scenes in very small units - takes the FP32 test below.
// Box tests are not on the parity path: a conservative test cannot change which triangle is hit, only how many nodes
// are visited (measured: DESIGN.md section 5.2).
// ------------------------------------------------------------------------------------------------------------
constexpr float kHalfSlackSteps = 1.125f;

__device__ __forceinline__ __half2 asHalf2(uint32_t u)
{
    return *reinterpret_cast<const __half2*>(&u);
}
__device__ __forceinline__ uint32_t asUint(__half2 h)
{
    return *reinterpret_cast<const uint32_t*>(&h);
}

// children 2*PAIR and 2*PAIR+1: bytes (2*PAIR & 3) and (2*PAIR & 3) + 1 of their half's plane words
template<int PAIR>
__device__ __forceinline__ void testChildPair(uint32_t nx, uint32_t ny, uint32_t nz, uint32_t fx, uint32_t fy, uint32_t fz, __half2 ax,
                                              __half2 ay, __half2 az, __half2 bnx, __half2 bny, __half2 bnz, __half2 bfx, __half2 bfy,
                                              __half2 bfz, __half2 tminH, __half2 tmaxH, uint32_t magic, uint32_t& hitPairs)
{
    constexpr uint32_t sel = ((2 * PAIR) & 3) == 0 ? 0x4140u : 0x4342u; // {q_i, 0x64, q_i+1, 0x64}
    const __half2 tnx = __hfma2(asHalf2(__byte_perm(nx, magic, sel)), ax, bnx), tfx = __hfma2(asHalf2(__byte_perm(fx, magic, sel)), ax, bfx);
    const __half2 tny = __hfma2(asHalf2(__byte_perm(ny, magic, sel)), ay, bny), tfy = __hfma2(asHalf2(__byte_perm(fy, magic, sel)), ay, bfy);
    const __half2 tnz = __hfma2(asHalf2(__byte_perm(nz, magic, sel)), az, bnz), tfz = __hfma2(asHalf2(__byte_perm(fz, magic, sel)), az, bfz);
    const __half2 tn = __hmax2(__hmax2(__hmax2(tnx, tny), tnz), tminH);
    const __half2 tf = __hmin2(__hmin2(__hmin2(tfx, tfy), tfz), tmaxH);
    // child 2*PAIR -> bit 2*PAIR, child 2*PAIR+1 -> bit 16 + 2*PAIR + 1 (folded by the caller)
    hitPairs |= __hle2_mask(tn, tf) & ((1u << (2 * PAIR)) | (0x20000u << (2 * PAIR)));
}
#endif

VKRT_HD uint32_t intersectNodeSlots(const NodeWords& w, F3 o, F3 idir, uint32_t octant, float tminP, float tmaxP, uint32_t one)
{
    // The caller passes idir, tminP, tmaxP multiplied by the ray's power of two kappa (setBoxRay): scaling every distance by
    // one positive power of two changes no comparison, it only moves the plane spacings a_k = scale_k * idir_k of the
    // scene's nodes into the range of the packed-half arithmetic below.
#if defined(VKRT_NODE_TEST_HALF2)
    {
        const uint32_t eb = floatAsUint(w.n0.w);
        // a = scale * idir = the distance between neighbouring quantisation planes along the ray, truncated to 11 significant
        // bits: the half conversion below is then exact, and b' is formed with the SAME a the HFMA2 multiplies by, so the
        // 1024 a offset of the dequantisation cancels exactly (what remains is q (a - a_r))
        const float Ax = uintAsFloat(floatAsUint(uintAsFloat(eb << 16) * idir.x) & 0xFFFFE000u),
                    Ay = uintAsFloat(floatAsUint(uintAsFloat(eb & 0xFFFF0000u) * idir.y) & 0xFFFFE000u),
                    Az = uintAsFloat(floatAsUint(uintAsFloat(w.n1.w << 16) * idir.z) & 0xFFFFE000u);
        const float amax = fmaxf(fmaxf(fabsf(Ax), fabsf(Ay)), fabsf(Az)), amin = fminf(fminf(fabsf(Ax), fabsf(Ay)), fabsf(Az));
        if (amax <= 32.0f && amin >= 6.103515625e-05f) // a in [2^-14, 32] (in practice >= 2^-13, see above); false for inf / NaN
        {
            const float t0x = (w.n0.x - o.x) * idir.x, t0y = (w.n0.y - o.y) * idir.y, t0z = (w.n0.z - o.z) * idir.z;
            // entry distance of the node's own box [p, p + 255 scale] (the near plane of axis k is p_k when a_k > 0), not
            // before the start of the ray: at tRef the ray is inside the box on every axis, so |t0_k - tRef| <= 255 |a_k|
            const float tRef = fmaxf(fmaxf(fmaxf(fmaExplicit(fminf(Ax, 0.0f), 255.0f, t0x), fmaExplicit(fminf(Ay, 0.0f), 255.0f, t0y)),
                                           fmaExplicit(fminf(Az, 0.0f), 255.0f, t0z)), tminP);
            const float bx = fmaExplicit(Ax, -1024.0f, t0x - tRef), by = fmaExplicit(Ay, -1024.0f, t0y - tRef), bz = fmaExplicit(Az, -1024.0f, t0z - tRef);
            const __half2 ax = __float2half2_rn(Ax), ay = __float2half2_rn(Ay), az = __float2half2_rn(Az);
            const __half2 bnx = __float2half2_rn(fmaExplicit(-fabsf(Ax), kHalfSlackSteps, bx)), bfx = __float2half2_rn(fmaExplicit(fabsf(Ax), kHalfSlackSteps, bx));
            const __half2 bny = __float2half2_rn(fmaExplicit(-fabsf(Ay), kHalfSlackSteps, by)), bfy = __float2half2_rn(fmaExplicit(fabsf(Ay), kHalfSlackSteps, by));
            const __half2 bnz = __float2half2_rn(fmaExplicit(-fabsf(Az), kHalfSlackSteps, bz)), bfz = __float2half2_rn(fmaExplicit(fabsf(Az), kHalfSlackSteps, bz));
            const float tlo = tminP - tRef, thi = tmaxP - tRef; // <= 0 and (for a node in reach) >= 0
            const __half2 tminH = __float2half2_rn(fmaExplicit(-fabsf(tlo), 0.0009765625f, tlo) - 1e-7f),
                          tmaxH = __float2half2_rn(fmaExplicit(fabsf(thi), 0.0009765625f, thi) + 1e-7f);
            uint32_t hitPairs = 0;
            const uint32_t m64 = one ^ 0x5BE46464u; // 0x64646464, derived from the kernel parameter so that PRMT keeps an immediate selector
            {
                const uint32_t nx = (octant & 1u) ? w.n3.z : w.n2.x, fx = (octant & 1u) ? w.n2.x : w.n3.z;
                const uint32_t ny = (octant & 2u) ? w.n4.x : w.n2.z, fy = (octant & 2u) ? w.n2.z : w.n4.x;
                const uint32_t nz = (octant & 4u) ? w.n4.z : w.n3.x, fz = (octant & 4u) ? w.n3.x : w.n4.z;
                testChildPair<0>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminH, tmaxH, m64, hitPairs);
                testChildPair<1>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminH, tmaxH, m64, hitPairs);
            }
            {
                const uint32_t nx = (octant & 1u) ? w.n3.w : w.n2.y, fx = (octant & 1u) ? w.n2.y : w.n3.w;
                const uint32_t ny = (octant & 2u) ? w.n4.y : w.n2.w, fy = (octant & 2u) ? w.n2.w : w.n4.y;
                const uint32_t nz = (octant & 4u) ? w.n4.w : w.n3.y, fz = (octant & 4u) ? w.n3.y : w.n4.w;
                testChildPair<2>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminH, tmaxH, m64, hitPairs);
                testChildPair<3>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminH, tmaxH, m64, hitPairs);
            }
            return (hitPairs | (hitPairs >> 16)) & 0xFFu;
        }
    }
#endif
    // n0.w = sx | sy << 16, n1.w = sz | imask << 16: upper halves of the scale floats; the byte-in-mantissa dequantisation
    // (1 + q 2^-15, byteToUnitFloat) wants 2^15 * scale
    const uint32_t e = floatAsUint(w.n0.w);
    const float sx = uintAsFloat(e << 16), sy = uintAsFloat(e & 0xFFFF0000u), sz = uintAsFloat(w.n1.w << 16);
    const float ax = (sx * 32768.0f) * idir.x, ay = (sy * 32768.0f) * idir.y, az = (sz * 32768.0f) * idir.z;
    const float bx = fmaExplicit(w.n0.x - o.x, idir.x, -ax), by = fmaExplicit(w.n0.y - o.y, idir.y, -ay),
                bz = fmaExplicit(w.n0.z - o.z, idir.z, -az);
    const float slack = 2.3841858e-07f; // 2^-22 of ax = 2^-7 quantisation step
    const float bnx = fmaExplicit(-fabsf(ax), slack, bx), bfx = fmaExplicit(fabsf(ax), slack, bx);
    const float bny = fmaExplicit(-fabsf(ay), slack, by), bfy = fmaExplicit(fabsf(ay), slack, by);
    const float bnz = fmaExplicit(-fabsf(az), slack, bz), bfz = fmaExplicit(fabsf(az), slack, bz);
    uint32_t hitSlots = 0;
    // n1 = {childBase, primBase, leafBits, sz | imask};   n2 = {qlox[0..3], qlox[4..7], qloy[0..3], qloy[4..7]}
    // n3 = {qloz lo, qloz hi, qhix lo, qhix hi};          n4 = {qhiy lo, qhiy hi, qhiz lo, qhiz hi}
    {
        // near/far plane selection by ray direction sign
        const uint32_t nx = (octant & 1u) ? w.n3.z : w.n2.x, fx = (octant & 1u) ? w.n2.x : w.n3.z;
        const uint32_t ny = (octant & 2u) ? w.n4.x : w.n2.z, fy = (octant & 2u) ? w.n2.z : w.n4.x;
        const uint32_t nz = (octant & 4u) ? w.n4.z : w.n3.x, fz = (octant & 4u) ? w.n3.x : w.n4.z;
        testChild<0>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<1>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<2>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<3>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
    }
    {
        const uint32_t nx = (octant & 1u) ? w.n3.w : w.n2.y, fx = (octant & 1u) ? w.n2.y : w.n3.w;
        const uint32_t ny = (octant & 2u) ? w.n4.y : w.n2.w, fy = (octant & 2u) ? w.n2.w : w.n4.y;
        const uint32_t nz = (octant & 4u) ? w.n4.w : w.n3.y, fz = (octant & 4u) ? w.n3.y : w.n4.w;
        testChild<4>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<5>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<6>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
        testChild<7>(nx, ny, nz, fx, fy, fz, ax, ay, az, bnx, bny, bnz, bfx, bfy, bfz, tminP, tmaxP, one, hitSlots);
    }
    return hitSlots;
}

// Internal children among `hitSlots` (8 bits) in traversal priority order (nearest octant-order child in the highest
// bit).  Only the low 8 bits of `imask` matter.
VKRT_HD uint32_t orderInnerSlots(uint32_t hitSlots, uint32_t imask, uint32_t octant)
{
#if defined(__CUDA_ARCH__)
    return sPermuteSlots[((7u - octant) << 8) | (hitSlots & imask)];
#else
    return permuteSlots(hitSlots & imask, 7u - octant);
#endif
}

VKRT_HD uint2 intersectNode(const NodeWords& w, F3 o, F3 idir, uint32_t octant, float tminP, float tmaxP, uint32_t one)
{
    const uint32_t hitSlots = intersectNodeSlots(w, o, idir, octant, tminP, tmaxP, one);
#if defined(__CUDA_ARCH__)
    return make_uint2(orderInnerSlots(hitSlots, w.n1.w >> 16, octant), sSpreadSlots[hitSlots] & w.n1.z);
#else
    return make_uint2(orderInnerSlots(hitSlots, w.n1.w >> 16, octant), spreadSlots(hitSlots) & w.n1.z);
#endif
}

VKRT_HD void loadTriangle(const PackedTri* tris, uint32_t index, F3& v0, F3& v1, F3& v2, uint32_t& geometry,
                          uint32_t& primitive, uint32_t& buildIndex)
{
    const float4* p = reinterpret_cast<const float4*>(tris + index);
#if defined(__CUDA_ARCH__)
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
#else
    const float4 a = p[0], b = p[1], c = p[2];
#endif
    v0 = f3(a.x, a.y, a.z);
    v1 = f3(a.w, b.x, b.y);
    v2 = f3(b.z, b.w, c.x);
    geometry = floatAsUint(c.y);
    primitive = floatAsUint(c.z);
    buildIndex = floatAsUint(c.w);
}

// State of the best hit so far for one ray (world-parametrised t, shared across instances).
struct BestHit
{
    float t;             // current tmax (shrinks on closest hits)
    float b1, b2;
    uint32_t tri;        // packed-triangle slot in the instance's BLAS, kInvalid = none
    uint32_t instance;
    uint32_t buildIndex; // tie-break key of the current hit
};

// Per-ray traversal state inside one BLAS (object space).  Kept as plain scalars so it lives in registers.
struct RayCtx
{
    F3 o, idir;   // idir = (1 / d) * kappa
    float kappa;  // power of two that normalises the distances of the box tests (setBoxRay)
    RayPre rp;
    uint32_t octant;
    float tmin, tminP; // tminP = padDown(tmin) * kappa
    uint2 ngroup; // x = base index of the internal children, y = hit bits 24..31 | imask in bits 0..7
    uint2 tgroup; // x = first packed triangle of the node, y = hit bits 0..23
    int sp;
};

// Box-test constants of a ray: reciprocal direction, octant and clamp, all distances scaled by kappa.
// `boxScale` is the hierarchy's power of two 2^(114 - e_lo), e_lo = the smallest scale exponent of any of its nodes
// (Blas::boxScale / Tlas::boxScale): with the smallest |1 / d_k| of the ray normalised to [1, 2), the smallest plane
// spacing a = scale * idir of the hierarchy lands at 2^-13, the bottom of the half range, and the spacings of larger
// nodes and of the ray's other axes have 18 binades above it.
VKRT_HD void setBoxRay(RayCtx& c, F3 d, float tmin, float boxScale)
{
    const F3 id = f3(safeRcp(d.x), safeRcp(d.y), safeRcp(d.z));
    const float m = fminf(fminf(fabsf(id.x), fabsf(id.y)), fabsf(id.z)); // finite for a valid ray (d != 0)
    // 2^(127 - E) for m = 2^(E - 127) * [1, 2); exponent fields 0 (d overflowed 1/x) and 255 cannot arise from finite d != 0
    // except through denormals, for which any power of two is as good as another: kappa = 1
    // 2^(127 - E) for m = 2^(E - 127) * [1, 2); exponent fields 0 and 255 cannot arise from a finite d != 0 except through
    // denormals, for which the normalisation does not matter (such nodes take the FP32 test)
    const uint32_t e = floatAsUint(m) & 0x7F800000u;
    c.kappa = ((e == 0u || e == 0x7F800000u) ? 1.0f : uintAsFloat(0x7F000000u - e)) * boxScale;
    c.idir = id * c.kappa;
    c.octant = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    c.tmin = tmin;
    c.tminP = padDown(tmin) * c.kappa;
}

VKRT_HD void beginTraversal(RayCtx& c, F3 o, F3 d, float tmin, float boxScale)
{
    c.o = o;
    c.rp = makeRayPre(d);
    setBoxRay(c, d, tmin, boxScale);
    c.ngroup = make_uint2(0u, 0x80000000u);
    c.tgroup = make_uint2(0u, 0u);
    c.sp = 0;
}

constexpr int kStepDone = 1;  // nothing left in this BLAS
constexpr int kStepFound = 2; // ANY: an occluder was accepted
constexpr int kStepHit = 4;   // flag: at least one triangle was accepted during this step

// One traversal step: pops the nearest pending internal child (if any), tests its 8 child boxes, then tests the
// triangles the node references.  Returns 0 to continue, or kStepDone / kStepFound, optionally ORed with kStepHit.
// `spFloor` is the stack height at which this BLAS was entered (0 unless a top-level traversal is suspended below);
// `overflowFlag` (may be null) is set when a group had to be dropped because the stack is full: the builders refuse
// hierarchies deeper than the stack (kStackSize), so this is a defensive signal that vkrt surfaces as an error.
template<bool ANY>
VKRT_HD int traversalStep(RayCtx& c, RayStack& stack, const Node8* nodes, const PackedTri* tris, uint32_t instance, BestHit& best,
                          TraceCounters* counters, uint32_t one, int spFloor = 0, uint32_t* overflowFlag = nullptr)
{
    int result = 0;
    uint32_t leafBits = 0;
    if (c.ngroup.y > 0x00FFFFFFu)
    {
        const uint32_t octinv = 7u - c.octant;
        const uint32_t bit = highestBit(c.ngroup.y);
        const uint32_t base = c.ngroup.x;
        c.ngroup.y &= ~(1u << bit);
        if (c.ngroup.y > 0x00FFFFFFu)
        {
            if (c.sp < kStackSize)
            {
                stack.set(c.sp++, c.ngroup);
            }
            else
            {
                if (counters)
                {
                    counters->overflow = 1;
                }
                if (overflowFlag)
                {
                    *overflowFlag = 1u;
                }
            }
        }
        const uint32_t slot = (bit ^ octinv) & 7u; // bit is 24..31
        const uint32_t rel = popCount(c.ngroup.y & ~(0xFFFFFFFFu << slot) & 0xFFu);
        const NodeWords w = loadNode(nodes, base + rel);
        if (counters)
        {
            ++counters->nodes;
        }
        const uint2 hits = intersectNode(w, c.o, c.idir, c.octant, c.tminP, padUp(best.t) * c.kappa, one);
        const uint32_t imask = w.n1.w >> 16;
        c.ngroup = make_uint2(w.n1.x, (hits.x << 24) | imask);
        c.tgroup = make_uint2(w.n1.y, hits.y);
        leafBits = w.n1.z;
    }
    else
    {
        c.tgroup = make_uint2(0u, 0u);
    }

    while (c.tgroup.y != 0u)
    {
        const uint32_t bit = highestBit(c.tgroup.y);
        const uint32_t below = (1u << bit) - 1u;
        c.tgroup.y &= below;
        const uint32_t slotIndex = c.tgroup.x + popCount(leafBits & below); // leaves are stored in slot order
        F3 v0, v1, v2;
        uint32_t geometry, primitive, buildIndex;
        loadTriangle(tris, slotIndex, v0, v1, v2, geometry, primitive, buildIndex);
        if (counters)
        {
            ++counters->tris;
        }
        (void)geometry;
        (void)primitive;
        bool tieOk = false;
        if (!ANY)
        {
            tieOk = best.tri != kInvalid && (instance < best.instance || (instance == best.instance && buildIndex < best.buildIndex));
        }
        float t, b1, b2;
        if (intersectTriangle(v0, v1, v2, c.o, c.rp, c.tmin, best.t, tieOk, t, b1, b2))
        {
            best.t = t, best.b1 = b1, best.b2 = b2;
            best.tri = slotIndex, best.instance = instance, best.buildIndex = buildIndex;
            result |= kStepHit;
            if (ANY)
            {
                return result | kStepFound;
            }
        }
    }

    if (c.ngroup.y <= 0x00FFFFFFu)
    {
        if (c.sp <= spFloor)
        {
            return result | kStepDone;
        }
        c.ngroup = stack.get(--c.sp);
    }
    return result;
}

// Traverses one BLAS with an object-space ray.  ANY: return true at the first accepted triangle.
template<bool ANY>
VKRT_HD bool traverseBlas(const Node8* nodes, const PackedTri* tris, F3 o, F3 d, float tmin, uint32_t instance,
                          BestHit& best, TraceCounters* counters, uint32_t one, float boxScale)
{
    RayCtx c;
    beginTraversal(c, o, d, tmin, boxScale);
    RayStack stack;
    bool found = false;
    for (;;)
    {
        const int r = traversalStep<ANY>(c, stack, nodes, tris, instance, best, counters, one);
        found = found || (r & kStepHit);
        if (r & (kStepDone | kStepFound))
        {
            return found;
        }
    }
}

VKRT_HD bool validRay(F3 o, F3 d) { return finite3(o) && finite3(d) && !(d.x == 0.0f && d.y == 0.0f && d.z == 0.0f); }

// ------------------------------------------------------------------------------------------------------------
// Top level (replaces the TLAS the driver builds in Raytracer::createTLAS, src/Raytracer.cpp:1159-1283).
// The TLAS is a Node8 hierarchy over the instances' world boxes, built on the device by the same builder as a BLAS
// (k_instance_bounds feeds it boxes instead of k_tri_bounds), with exactly one instance per leaf slot:
// Node8::reserved holds the mask of leaf slots, and the instance of leaf slot s is
//     leafInstance[primBase + popcount(reserved & ((1 << s) - 1))].
// One stack serves both levels.  While a ray is inside a BLAS the suspended top-level work lies below `spFloor`:
//   node group      {childBase, hits << 24 | imask}                        (x < 2^31)
//   instance group  {primBase | 2^31, hit leaf slots << 8 | leaf slots}    (x >= 2^31)
// A ray enters an instance only if VkAccelerationStructureInstanceKHR::mask & cullMask != 0 (traceRayEXT's cullMask,
// 0xFF in shaders/shader.rgen:52 and shader.rchit:133; the reference's instance has mask 0xFF, Raytracer.cpp:1174).
// ------------------------------------------------------------------------------------------------------------
struct TlasRefs
{
    const Node8* nodes;            // top-level nodes (null: a single instance, no top level)
    const uint32_t* leafInstance;  // instance index per top-level leaf, in node order
    const InstanceRecord* instances;
    uint32_t instanceCount;
    uint32_t cullMask;
    float boxScale; // of the top-level hierarchy (see setBoxRay)
};

constexpr uint32_t kTopInstanceGroup = 0x80000000u;
constexpr int kTopDone = 1;  // no instance left to enter
constexpr int kTopEnter = 2; // `instanceOut` names the next instance whose box the ray enters

// World-space box-traversal state (no watertight constants: the top level tests boxes only).
VKRT_HD void beginTop(RayCtx& c, F3 o, F3 d, float tmin, int sp, float boxScale)
{
    c.o = o;
    setBoxRay(c, d, tmin, boxScale);
    c.ngroup = make_uint2(0u, 0u);
    c.tgroup = make_uint2(0u, 0u); // holds the pending instance group at the top level
    c.sp = sp;
}

VKRT_HD void pushGroup(RayCtx& c, RayStack& stack, uint2 g, TraceCounters* counters, uint32_t* overflowFlag)
{
    if (c.sp < kStackSize)
    {
        stack.set(c.sp++, g);
    }
    else
    {
        if (counters)
        {
            counters->overflow = 1;
        }
        if (overflowFlag)
        {
            *overflowFlag = 1u;
        }
    }
}

// Advances the top-level traversal until the ray is about to enter an instance (kTopEnter) or nothing is left
// (kTopDone).  `tmax` is the current hit distance.  c.ngroup / c.tgroup hold the pending node and instance groups.
VKRT_HD int topStep(RayCtx& c, RayStack& stack, const TlasRefs& top, float tmax, TraceCounters* counters, uint32_t one,
                    uint32_t* overflowFlag, uint32_t& instanceOut)
{
    for (;;)
    {
        if (c.tgroup.y & 0xFF00u)
        {
            const uint32_t slot = highestBit(c.tgroup.y) - 8u;
            c.tgroup.y &= ~(0x100u << slot);
            const uint32_t idx = top.leafInstance[(c.tgroup.x & ~kTopInstanceGroup) + popCount(c.tgroup.y & 0xFFu & ((1u << slot) - 1u))];
            if (top.instances[idx].mask & top.cullMask)
            {
                instanceOut = idx;
                return kTopEnter;
            }
            continue;
        }
        if (c.ngroup.y > 0x00FFFFFFu)
        {
            const uint32_t octinv = 7u - c.octant;
            const uint32_t bit = highestBit(c.ngroup.y);
            const uint32_t base = c.ngroup.x;
            c.ngroup.y &= ~(1u << bit);
            if (c.ngroup.y > 0x00FFFFFFu)
            {
                pushGroup(c, stack, c.ngroup, counters, overflowFlag);
            }
            const uint32_t slot = (bit ^ octinv) & 7u;
            const uint32_t rel = popCount(c.ngroup.y & ~(0xFFFFFFFFu << slot) & 0xFFu);
            const NodeWords w = loadNode(top.nodes, base + rel);
            if (counters)
            {
                ++counters->nodes;
            }
            const uint32_t hitSlots = intersectNodeSlots(w, c.o, c.idir, c.octant, c.tminP, padUp(tmax) * c.kappa, one);
            const uint32_t imask = (w.n1.w >> 16) & 0xFFu, leafSlots = w.n1.w >> 24;
            c.ngroup = make_uint2(w.n1.x, (orderInnerSlots(hitSlots, imask, c.octant) << 24) | imask);
            c.tgroup = make_uint2(w.n1.y | kTopInstanceGroup, ((hitSlots & leafSlots) << 8) | leafSlots);
            continue;
        }
        if (c.sp == 0)
        {
            return kTopDone;
        }
        const uint2 g = stack.get(--c.sp);
        if (g.x & kTopInstanceGroup)
        {
            c.tgroup = g;
        }
        else
        {
            c.ngroup = g;
        }
    }
}

// Suspends the top level (pending groups go onto the stack) and moves the ray into `inst`.  Returns false when
// the object-space ray is not traceable (non-finite or zero direction after the transform): the top level then
// simply continues.  On success c is the object-space state and `spFloor` the stack height to return at.
VKRT_HD bool enterInstanceTop(RayCtx& c, RayStack& stack, const InstanceRecord& inst, F3 o, F3 d, float tmin, TraceCounters* counters,
                              uint32_t* overflowFlag, int& spFloor)
{
    const F3 oo = affinePoint(inst.w2o, o), dd = affineDir(inst.w2o, d);
    if (!validRay(oo, dd))
    {
        return false;
    }
    if (c.ngroup.y > 0x00FFFFFFu)
    {
        pushGroup(c, stack, c.ngroup, counters, overflowFlag);
    }
    if (c.tgroup.y & 0xFF00u)
    {
        pushGroup(c, stack, c.tgroup, counters, overflowFlag);
    }
    const int sp = c.sp;
    beginTraversal(c, oo, dd, tmin, inst.boxScale);
    c.sp = sp;
    spFloor = sp;
    return true;
}

// traceRayEXT over the whole scene.  Returns true on any accepted hit.  `best.t` must be initialised to the ray's
// tmax and best.tri to kInvalid.  (Sequential form, used by the host emulation; the kernels in wavefront.cu run the same
// steps with the loop turned inside out so that a warp can re-converge between them.)
template<bool ANY>
VKRT_HD bool traceScene(const TlasRefs& top, F3 o, F3 d, float tmin, BestHit& best, TraceCounters* counters, uint32_t one = 0x3F800000u)
{
    if (!validRay(o, d) || !(best.t > tmin))
    {
        return false;
    }
    if (!top.nodes)
    {
        const InstanceRecord& inst = top.instances[0];
        if (!(inst.mask & top.cullMask))
        {
            return false;
        }
        const F3 oo = affinePoint(inst.w2o, o);
        const F3 dd = affineDir(inst.w2o, d);
        if (!validRay(oo, dd))
        {
            return false;
        }
        return traverseBlas<ANY>(inst.nodes, inst.tris, oo, dd, tmin, 0u, best, counters, one, inst.boxScale);
    }
    RayCtx c;
    RayStack stack;
    beginTop(c, o, d, tmin, 0, top.boxScale);
    c.ngroup = make_uint2(0u, 0x80000000u); // the top-level root
    bool found = false;
    for (;;)
    {
        uint32_t idx = 0;
        if (topStep(c, stack, top, best.t, counters, one, nullptr, idx) == kTopDone)
        {
            return found;
        }
        const InstanceRecord& inst = top.instances[idx];
        int spFloor = 0;
        if (!enterInstanceTop(c, stack, inst, o, d, tmin, counters, nullptr, spFloor))
        {
            continue;
        }
        for (;;)
        {
            const int r = traversalStep<ANY>(c, stack, inst.nodes, inst.tris, idx, best, counters, one, spFloor);
            found = found || (r & kStepHit);
            if (ANY && (r & kStepFound))
            {
                return true;
            }
            if (r & kStepDone)
            {
                break;
            }
        }
        beginTop(c, o, d, tmin, spFloor, top.boxScale);
    }
}

// World box of an instance: the 8 corners of the BLAS box through object->world, padded by a few ulps (the top level
// must never cull a ray that the object-space traversal, with its own roundings, would accept).
VKRT_HD void instanceWorldBounds(const InstanceRecord& r, float lo[3], float hi[3])
{
    for (int k = 0; k < 3; ++k)
    {
        lo[k] = INFINITY;
        hi[k] = -INFINITY;
    }
    for (int corner = 0; corner < 8; ++corner)
    {
        const F3 p = f3((corner & 1) ? r.blasHi[0] : r.blasLo[0], (corner & 2) ? r.blasHi[1] : r.blasLo[1], (corner & 4) ? r.blasHi[2] : r.blasLo[2]);
        const F3 q = affinePoint(r.o2w, p);
        lo[0] = fminf(lo[0], q.x), lo[1] = fminf(lo[1], q.y), lo[2] = fminf(lo[2], q.z);
        hi[0] = fmaxf(hi[0], q.x), hi[1] = fmaxf(hi[1], q.y), hi[2] = fmaxf(hi[2], q.z);
    }
    for (int k = 0; k < 3; ++k)
    {
        const float padAbs = (fabsf(lo[k]) + fabsf(hi[k])) * 1e-5f + 1e-30f;
        lo[k] -= padAbs;
        hi[k] += padAbs;
    }
}

} // namespace vkrt
