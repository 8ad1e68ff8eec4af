// builder.cuh — per-item logic of the GPU BVH builder that replaces vkCmdBuildAccelerationStructuresKHR
// (src/Raytracer.cpp:1027-1157 createBLAS).  Pipeline (one kernel per step, see bvh_build.cu):
//
//   triangle bounds + scene bounds -> 30-bit Morton codes -> LSD radix sort -> agglomerative LBVH with
//   fused refit (Apetrei 2014) -> greedy surface-area collapse into 8-wide nodes with octant-ordered child
//   slots -> conservative 8-bit quantisation (80-byte nodes) -> 48-byte packed triangles.
//
// Functions are host+device so the logic can be unit-tested by the no-GPU emulation harness (tests/emu).
#pragma once

#include "common.cuh"
#include "device_math.cuh"

namespace vkrt
{

// (geometry, primitive) lookup for a triangle's build index (= position in geometry-major order).
struct GeometryTable
{
    const uint32_t* firstTri;  // geometryCount+1 prefix sums of primitiveCount
    const uint64_t* indexBase; // geometryCount: first index of the geometry (indexByteOffset / 4)
    uint32_t geometryCount;
};

VKRT_HD uint32_t findGeometry(const GeometryTable& g, uint32_t buildIndex)
{
    uint32_t lo = 0, hi = g.geometryCount; // invariant: firstTri[lo] <= buildIndex < firstTri[hi]
    while (hi - lo > 1)
    {
        const uint32_t mid = (lo + hi) >> 1;
        if (g.firstTri[mid] <= buildIndex)
        {
            lo = mid;
        }
        else
        {
            hi = mid;
        }
    }
    return lo;
}

VKRT_HD void fetchTriangle(const GeometryTable& g, const uint8_t* vertices, uint32_t stride, const uint32_t* indices,
                           uint32_t buildIndex, F3& v0, F3& v1, F3& v2, uint32_t& geometry, uint32_t& primitive)
{
    geometry = findGeometry(g, buildIndex);
    primitive = buildIndex - g.firstTri[geometry];
    const uint64_t base = g.indexBase[geometry] + 3ull * primitive;
    const uint32_t i0 = indices[base + 0], i1 = indices[base + 1], i2 = indices[base + 2];
    const float* a = reinterpret_cast<const float*>(vertices + (size_t)i0 * stride);
    const float* b = reinterpret_cast<const float*>(vertices + (size_t)i1 * stride);
    const float* c = reinterpret_cast<const float*>(vertices + (size_t)i2 * stride);
    v0 = f3(a[0], a[1], a[2]);
    v1 = f3(b[0], b[1], b[2]);
    v2 = f3(c[0], c[1], c[2]);
}

VKRT_HD uint32_t expandBits10(uint32_t v) // 10 bits -> every third bit
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// 30-bit Morton code of p inside [lo, lo+extent]; invExtent = 1/extent (0 for flat axes)
VKRT_HD uint32_t morton30(F3 p, F3 lo, F3 invExtent)
{
    const float fx = fminf(fmaxf((p.x - lo.x) * invExtent.x * 1024.0f, 0.0f), 1023.0f);
    const float fy = fminf(fmaxf((p.y - lo.y) * invExtent.y * 1024.0f, 0.0f), 1023.0f);
    const float fz = fminf(fmaxf((p.z - lo.z) * invExtent.z * 1024.0f, 0.0f), 1023.0f);
    return (expandBits10((uint32_t)fx) << 2) | (expandBits10((uint32_t)fy) << 1) | expandBits10((uint32_t)fz);
}

// ------------------------------------------------------------------------------------------------
// Binary LBVH (intermediate): N-1 internal nodes, children >= 0 internal, < 0 => ~sortedLeafIndex.
// lo.w carries the number of leaves of the subtree (as raw bits); the leaves need not be contiguous in sorted
// order (PLOC merges clusters that are not neighbours), so small subtrees are enumerated by walking them.
// ------------------------------------------------------------------------------------------------
// One 64-byte record per binary node (one DRAM burst per visit; three separate arrays cost three), and one 32-byte
// box per triangle: at 50 M triangles the collapse and the LBVH climb are bound by exactly these random fetches.
struct alignas(64) BinNode
{
    float4 lo; // w = leaf count (raw bits)
    float4 hi;
    int2 children;
    int2 pad0;
    float4 pad1;
};
static_assert(sizeof(BinNode) == 64, "BinNode must be 64 bytes");

// Compact binary node of the fast-build (LBVH + greedy collapse) path: box and children in ONE 32-byte sector, so a
// fetch costs one DRAM sector instead of two.  The leaf count of a subtree does not fit; what the collapse needs of it
// (1, 2, 3 or "more than 3") travels in bits 29..30 of the REFERENCE to the node (kRefClass*): the LBVH climb knows the
// count when it finishes a node and writes the tagged reference into the parent.  Hence at most 2^29 - 1 primitives.
struct alignas(32) BinNode32
{
    float lo[3];
    float hi[3];
    int2 children;
};
static_assert(sizeof(BinNode32) == 32, "BinNode32 must be 32 bytes");
constexpr int kRefIndexMask = 0x1FFFFFFF;
constexpr uint32_t kCompactMaxPrims = 0x1FFFFFFFu;

struct Bvh2
{
    BinNode32* nodes32;   // non-null: compact mode (nodes is unused); references carry the count class
    float4* sortedBox;    // compact mode: the leaf boxes in sorted order, [2i] = min, [2i+1] = max (written by the climb)
    BinNode* nodes;
    int* flags;        // init -1; Apetrei range hand-off
    const uint32_t* keys;  // sorted Morton codes
    const uint32_t* vals;  // sorted build indices
    const float4* triBox;  // per build index: [2i] = min, [2i+1] = max
    uint32_t count;    // number of leaves
    int* root;
    // cost-optimal collapse (dpSolve / dpExpand below); dpChoice == nullptr selects the greedy collapse
    uint32_t* dpChoice; // per binary node: split decisions
    int* dpParent;      // [0, count): parent of sorted leaf i; [count, 2*count-1): parent of binary node i - count
    float dpPrimCost;   // cost of one triangle test relative to one node test
    uint32_t maxLeaf;   // primitives per leaf slot: kMaxLeafTris for a BLAS, 1 for a TLAS (one instance per slot)
};

// References to binary nodes: >= 0 internal (compact mode: index in bits 0..28, count class in bits 29..30: 0 = more than
// 3 leaves, 2 = two, 3 = three), < 0 = ~sortedLeaf.
VKRT_HD int refIndex(const Bvh2& b, int ref) { return b.nodes32 ? (ref & kRefIndexMask) : ref; }
VKRT_HD int tagRef(int node, uint32_t count) { return node | (count <= 3u ? (int)(count << 29) : 0); }
VKRT_HD int2 binChildren(const Bvh2& b, int ref)
{
    return b.nodes32 ? b.nodes32[ref & kRefIndexMask].children : b.nodes[ref].children;
}

VKRT_HD uint64_t lbvhDelta(const Bvh2& b, uint32_t j) // distance between sorted leaves j and j+1
{
    const uint32_t x = b.keys[j] ^ b.keys[j + 1];
    return ((uint64_t)x << 32) | (uint64_t)(j ^ (j + 1)); // index bits break ties between equal codes
}

#if defined(__CUDA_ARCH__)
// exchange with release + acquire semantics at device scope: the writes before it are visible to whoever reads the old
// value, and the reads after it see what the writer of the old value wrote before its own exchange
__device__ __forceinline__ int atomicExchAcqRel(int* p, int v)
{
    int old;
    asm volatile("atom.acq_rel.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
#define VKRT_ATOMIC_EXCH_ACQ_REL(ptr, v) atomicExchAcqRel((ptr), (v))
// arrival counter with the same semantics: what this thread wrote before is visible to the thread that arrives second, and
// that thread sees it after its own increment - one instruction instead of a relaxed add between two fences
__device__ __forceinline__ int atomicAddAcqRel(int* p, int v)
{
    int old;
    asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
#define VKRT_ATOMIC_ADD_ACQ_REL(ptr, v) atomicAddAcqRel((ptr), (v))
#define VKRT_ATOMIC_EXCH(ptr, v) atomicExch((ptr), (v))
#define VKRT_ATOMIC_ADD(ptr, v) atomicAdd((ptr), (v))
#define VKRT_FENCE() __threadfence()
#define VKRT_LOAD_CG_F4(ptr) __ldcg(ptr)
#else
template<typename T>
inline T hostExch(T* p, T v)
{
    T o = *p;
    *p = v;
    return o;
}
template<typename T>
inline T hostAdd(T* p, T v)
{
    T o = *p;
    *p = o + v;
    return o;
}
#define VKRT_ATOMIC_EXCH(ptr, v) hostExch((ptr), (v))
#define VKRT_ATOMIC_EXCH_ACQ_REL(ptr, v) hostExch((ptr), (v))
#define VKRT_ATOMIC_ADD_ACQ_REL(ptr, v) hostAdd((ptr), (v))
#define VKRT_ATOMIC_ADD(ptr, v) hostAdd((ptr), (v))
#define VKRT_FENCE() ((void)0)
#define VKRT_LOAD_CG_F4(ptr) (*(ptr))
#endif

// One thread per sorted leaf: climbs towards the root, the second thread to reach a node merges the boxes.
VKRT_HD void lbvhClimb(const Bvh2& b, uint32_t leaf)
{
    const uint32_t n = b.count;
    uint32_t left = leaf, right = leaf;
    int current = ~(int)leaf;
    const uint32_t tri = b.vals[leaf];
    float4 lo = b.triBox[2 * tri], hi = b.triBox[2 * tri + 1];
    for (;;)
    {
        uint32_t parent;
        int prev;
        bool isLeftChild;
        if (left == 0 || (right != n - 1 && lbvhDelta(b, right) < lbvhDelta(b, left - 1)))
        {
            parent = right;
            isLeftChild = true;
            b.nodes[parent].children.x = current;
            VKRT_FENCE();
            prev = VKRT_ATOMIC_EXCH(&b.flags[parent], (int)left);
            if (prev != -1)
            {
                right = (uint32_t)prev;
            }
        }
        else
        {
            parent = left - 1;
            isLeftChild = false;
            b.nodes[parent].children.y = current;
            VKRT_FENCE();
            prev = VKRT_ATOMIC_EXCH(&b.flags[parent], (int)right);
            if (prev != -1)
            {
                left = (uint32_t)prev;
            }
        }
        if (prev == -1)
        {
            return; // first to arrive: the sibling's thread finishes this node
        }
        VKRT_FENCE();
        // fetch the sibling's box (written before its fence)
        const int sibling = isLeftChild ? ((volatile BinNode*)b.nodes)[parent].children.y : ((volatile BinNode*)b.nodes)[parent].children.x;
        float4 slo, shi;
        if (sibling < 0)
        {
            const uint32_t st = b.vals[~sibling];
            slo = b.triBox[2 * st];
            shi = b.triBox[2 * st + 1];
        }
        else
        {
            slo = VKRT_LOAD_CG_F4(&b.nodes[sibling].lo);
            shi = VKRT_LOAD_CG_F4(&b.nodes[sibling].hi);
        }
        lo.x = fminf(lo.x, slo.x), lo.y = fminf(lo.y, slo.y), lo.z = fminf(lo.z, slo.z);
        hi.x = fmaxf(hi.x, shi.x), hi.y = fmaxf(hi.y, shi.y), hi.z = fmaxf(hi.z, shi.z);
        lo.w = uintAsFloat(right - left + 1u);
        hi.w = 0.0f;
        b.nodes[parent].lo = lo;
        b.nodes[parent].hi = hi;
        current = (int)parent;
        if (left == 0 && right == n - 1)
        {
            *b.root = current;
            return;
        }
    }
}

// The same climb over compact nodes.  Leaf boxes are copied into sorted order on the way (one coalesced 32-byte write per
// leaf), so that the collapse reads a leaf's box with one fetch instead of two dependent ones (vals, then triBox).
// The hand-off at a node is one exchange with release + acquire semantics (18.48 ms against 19.40 ms with a fence on
// either side of a relaxed exchange, 50 M triangles, profiles/r02_q_build_fence_ab.txt); -DVKRT_LBVH_TWO_FENCE and
// -DVKRT_LBVH_ONE_FENCE keep the measured alternatives buildable.
#if !defined(VKRT_LBVH_TWO_FENCE) && !defined(VKRT_LBVH_ONE_FENCE) && !defined(VKRT_LBVH_ACQ_REL)
#define VKRT_LBVH_ACQ_REL 1
#endif
VKRT_HD void lbvhClimbCompact(const Bvh2& b, uint32_t leaf)
{
    const uint32_t n = b.count;
    uint32_t left = leaf, right = leaf;
    int current = ~(int)leaf;
    const uint32_t tri = b.vals[leaf];
    float4 lo = b.triBox[2 * tri], hi = b.triBox[2 * tri + 1];
    b.sortedBox[2 * (size_t)leaf] = lo;
    b.sortedBox[2 * (size_t)leaf + 1] = hi;
    for (;;)
    {
        uint32_t parent;
        int prev;
        bool isLeftChild;
        if (left == 0 || (right != n - 1 && lbvhDelta(b, right) < lbvhDelta(b, left - 1)))
        {
            parent = right;
            isLeftChild = true;
            b.nodes32[parent].children.x = current;
#if defined(VKRT_LBVH_ACQ_REL)
            prev = VKRT_ATOMIC_EXCH_ACQ_REL(&b.flags[parent], (int)left);
#else
            VKRT_FENCE();
            prev = VKRT_ATOMIC_EXCH(&b.flags[parent], (int)left);
#endif
            if (prev != -1)
            {
                right = (uint32_t)prev;
            }
        }
        else
        {
            parent = left - 1;
            isLeftChild = false;
            b.nodes32[parent].children.y = current;
#if defined(VKRT_LBVH_ACQ_REL)
            prev = VKRT_ATOMIC_EXCH_ACQ_REL(&b.flags[parent], (int)right);
#else
            VKRT_FENCE();
            prev = VKRT_ATOMIC_EXCH(&b.flags[parent], (int)right);
#endif
            if (prev != -1)
            {
                left = (uint32_t)prev;
            }
        }
        if (prev == -1)
        {
            return; // first to arrive: the sibling's thread finishes this node
        }
#if !defined(VKRT_LBVH_ONE_FENCE) && !defined(VKRT_LBVH_ACQ_REL)
        VKRT_FENCE();
#endif
        // the sibling's box: its leaf box straight from the triangle boxes (its own copy in sortedBox may not be visible
        // yet), or the finished node's box (written before its fence)
        const volatile BinNode32* pn = b.nodes32 + parent;
        const int sibling = isLeftChild ? pn->children.y : pn->children.x;
        float slo[3], shi[3];
        if (sibling < 0)
        {
            const uint32_t st = b.vals[~sibling];
            const float4 l4 = b.triBox[2 * st], h4 = b.triBox[2 * st + 1];
            slo[0] = l4.x, slo[1] = l4.y, slo[2] = l4.z, shi[0] = h4.x, shi[1] = h4.y, shi[2] = h4.z;
        }
        else
        {
            const float4* q = reinterpret_cast<const float4*>(b.nodes32 + (sibling & kRefIndexMask));
            const float4 q0 = VKRT_LOAD_CG_F4(q), q1 = VKRT_LOAD_CG_F4(q + 1); // lo.xyz hi.x | hi.yz children
            slo[0] = q0.x, slo[1] = q0.y, slo[2] = q0.z, shi[0] = q0.w, shi[1] = q1.x, shi[2] = q1.y;
        }
        lo.x = fminf(lo.x, slo[0]), lo.y = fminf(lo.y, slo[1]), lo.z = fminf(lo.z, slo[2]);
        hi.x = fmaxf(hi.x, shi[0]), hi.y = fmaxf(hi.y, shi[1]), hi.z = fmaxf(hi.z, shi[2]);
        // both children are in place: the whole 32-byte record goes out as two 16-byte stores
        float4* out = reinterpret_cast<float4*>(b.nodes32 + parent);
        const int cx = isLeftChild ? current : sibling, cy = isLeftChild ? sibling : current;
        out[0] = make_float4(lo.x, lo.y, lo.z, hi.x);
        out[1] = make_float4(hi.y, hi.z, uintAsFloat((uint32_t)cx), uintAsFloat((uint32_t)cy));
        current = tagRef((int)parent, right - left + 1u);
        if (left == 0 && right == n - 1)
        {
            *b.root = current;
            return;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Collapse to 8-wide nodes
// ------------------------------------------------------------------------------------------------
#ifndef VKRT_MAX_LEAF_TRIS
#define VKRT_MAX_LEAF_TRIS 3
#endif
constexpr uint32_t kMaxLeafTris = VKRT_MAX_LEAF_TRIS; // a child slot references at most 3 triangles (unary count in 3 bits)

struct CollapseArgs
{
    Bvh2 bvh2;
    Node8* nodes;
    int* taskRef;          // per wide node: the binary subtree it covers (ref encoding as Bvh2 children)
    uint32_t* primSource;  // per packed-triangle slot: build index of the triangle to pack there
    uint32_t* nodeCounter; // next free wide node index
    uint32_t* primCounter; // next free packed-triangle slot
    float* sahAccum;       // [0] += area(node) (internal cost), [1] += area(leaf)*count
    uint32_t nodeCapacity;
    uint32_t* overflow;
};

struct CollapseChild
{
    int ref;
    F3 lo, hi;
    uint32_t count;
};

VKRT_HD CollapseChild collapseFetch(const Bvh2& b, int ref)
{
    CollapseChild c;
    c.ref = ref;
    if (b.nodes32)
    {
        if (ref < 0)
        {
            const uint32_t leaf = (uint32_t)~ref;
            const float4 lo = b.sortedBox[2 * (size_t)leaf], hi = b.sortedBox[2 * (size_t)leaf + 1];
            c.lo = f3(lo.x, lo.y, lo.z), c.hi = f3(hi.x, hi.y, hi.z);
            c.count = 1;
        }
        else
        {
            const float4* q = reinterpret_cast<const float4*>(b.nodes32 + (ref & kRefIndexMask));
            const float4 q0 = q[0], q1 = q[1];
            c.lo = f3(q0.x, q0.y, q0.z), c.hi = f3(q0.w, q1.x, q1.y);
            const uint32_t cls = ((uint32_t)ref >> 29) & 3u;
            c.count = cls ? cls : 4u; // 4 stands for "more than 3": only ever compared with maxLeaf <= 3
        }
        return c;
    }
    if (ref < 0)
    {
        const uint32_t leaf = (uint32_t)~ref;
        const uint32_t tri = b.vals[leaf];
        const float4 lo = b.triBox[2 * tri], hi = b.triBox[2 * tri + 1];
        c.lo = f3(lo.x, lo.y, lo.z), c.hi = f3(hi.x, hi.y, hi.z);
        c.count = 1;
    }
    else
    {
        const float4 lo = b.nodes[ref].lo, hi = b.nodes[ref].hi;
        c.lo = f3(lo.x, lo.y, lo.z), c.hi = f3(hi.x, hi.y, hi.z);
        c.count = floatAsUint(lo.w);
    }
    return c;
}

// Sorted-leaf indices of a subtree with at most kMaxLeafTris leaves, left to right.
VKRT_HD uint32_t collectLeaves(const Bvh2& b, int ref, uint32_t* out)
{
    if (ref < 0)
    {
        out[0] = (uint32_t)~ref;
        return 1;
    }
    // at most 3 leaves: one child is a leaf, the other a leaf or a pair of leaves
    uint32_t n = 0;
    const int2 ch = binChildren(b, ref);
    const int side[2] = {ch.x, ch.y};
    for (int k = 0; k < 2; ++k)
    {
        if (side[k] < 0)
        {
            out[n++] = (uint32_t)~side[k];
        }
        else
        {
            const int2 g = binChildren(b, side[k]);
            out[n++] = (uint32_t)~g.x; // grandchildren of a <= 3-leaf subtree are leaves
            out[n++] = (uint32_t)~g.y;
        }
    }
    return n;
}

VKRT_HD float halfArea(F3 lo, F3 hi)
{
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// ------------------------------------------------------------------------------------------------
// Cost-optimal distribution of a binary subtree over wide-node slots (Ylitie, Karras, Laine 2017, section 3.1).
// c[i], i = 1..7, is the cheapest SAH cost of the subtree represented as a forest of at most i slot entries:
//   c[1]  = leaf (<= 3 triangles): area * count * primCost;  else: area * 1 + dist[8]   (one wide node)
//   c[i]  = min(dist[i], c[i-1]),   dist[j] = min over k of cLeft[k] + cRight[j-k]
// The seven costs live in the padding of the 64-byte BinNode (pad1, pad0, hi.w), the decisions in one word:
//   bits 3(j-2)..3(j-2)+2 = k chosen for dist[j], j = 2..8;  bit 21+(i-2) set: c[i] == c[i-1] (use fewer entries).
// A subtree is an internal wide node exactly when it holds more than 3 triangles, as in the greedy collapse.
// ------------------------------------------------------------------------------------------------
constexpr float kDpPrimCost = 0.5f; // triangle test relative to node test; chosen on the Sponza stand-in (tools/bvh_quality.py)
VKRT_HD void dpFetch(const Bvh2& b, int ref, float* c)
{
    if (ref < 0)
    {
        const uint32_t tri = b.vals[(uint32_t)~ref];
        const float4 lo = b.triBox[2 * tri], hi = b.triBox[2 * tri + 1];
        const float a = halfArea(f3(lo.x, lo.y, lo.z), f3(hi.x, hi.y, hi.z)) * b.dpPrimCost;
        for (int i = 1; i < 8; ++i)
        {
            c[i] = a;
        }
    }
    else
    {
        // written by another thread of the same launch (dpClimb): read past L1
        const BinNode& n = b.nodes[ref];
        const float4 p1 = VKRT_LOAD_CG_F4(&n.pad1), hi = VKRT_LOAD_CG_F4(&n.hi);
        const float4 p0 = VKRT_LOAD_CG_F4(reinterpret_cast<const float4*>(&n.children)); // children, pad0
        c[1] = p1.x, c[2] = p1.y, c[3] = p1.z, c[4] = p1.w;
        c[5] = p0.z, c[6] = p0.w, c[7] = hi.w;
    }
}

VKRT_HD void dpSolve(const Bvh2& b, int ref)
{
    BinNode& n = b.nodes[ref];
    float cl[8], cr[8];
    dpFetch(b, n.children.x, cl);
    dpFetch(b, n.children.y, cr);
    uint32_t choice = 0;
    float dist[9];
    for (int j = 2; j <= 8; ++j)
    {
        float best = INFINITY;
        int bk = j - 7 > 1 ? j - 7 : 1;
        for (int k = bk; k <= 7 && k < j; ++k)
        {
            const float v = cl[k] + cr[j - k];
            if (v < best)
            {
                best = v, bk = k;
            }
        }
        dist[j] = best;
        choice |= (uint32_t)bk << (3 * (j - 2));
    }
    const float4 lo = n.lo, hi = n.hi;
    const float area = halfArea(f3(lo.x, lo.y, lo.z), f3(hi.x, hi.y, hi.z));
    const uint32_t count = floatAsUint(lo.w);
    float c[8];
    c[1] = count <= b.maxLeaf ? area * (float)count * b.dpPrimCost : area + dist[8];
    for (int i = 2; i < 8; ++i)
    {
        if (dist[i] < c[i - 1])
        {
            c[i] = dist[i];
        }
        else
        {
            c[i] = c[i - 1];
            choice |= 1u << (21 + (i - 2));
        }
    }
    n.pad1 = make_float4(c[1], c[2], c[3], c[4]);
    n.pad0 = make_int2((int)floatAsUint(c[5]), (int)floatAsUint(c[6]));
    n.hi.w = c[7];
    b.dpChoice[ref] = choice;
}

// Parent links for the climb: one thread per binary node.
VKRT_HD void dpLinkParents(const Bvh2& b, uint32_t node)
{
    const int2 ch = b.nodes[node].children;
    b.dpParent[ch.x < 0 ? (uint32_t)~ch.x : b.count + (uint32_t)ch.x] = (int)node;
    b.dpParent[ch.y < 0 ? (uint32_t)~ch.y : b.count + (uint32_t)ch.y] = (int)node;
}

// One thread per sorted leaf: climbs towards the root; the second thread to arrive at a node solves it (both children
// are final then).  b.flags must be zero on entry.
VKRT_HD void dpClimb(const Bvh2& b, uint32_t leaf)
{
    const int root = *b.root;
    int node = b.dpParent[leaf];
    for (;;)
    {
        if (VKRT_ATOMIC_ADD_ACQ_REL(&b.flags[node], 1) == 0)
        {
            return;
        }
        dpSolve(b, node);
        if (node == root)
        {
            return;
        }
        node = b.dpParent[b.count + (uint32_t)node];
    }
}

// Slot entries of the wide node that covers binary node `ref` (more than 3 triangles), left to right.
VKRT_HD int dpExpand(const Bvh2& b, int ref, int* out)
{
    int stackRef[8], stackSlots[8];
    int sp = 0, n = 0;
    {
        const int2 ch = b.nodes[ref].children;
        const int k = (int)(b.dpChoice[ref] >> 18) & 7; // dist[8]
        stackRef[sp] = ch.y, stackSlots[sp++] = 8 - k;
        stackRef[sp] = ch.x, stackSlots[sp++] = k;
    }
    while (sp > 0)
    {
        const int m = stackRef[--sp];
        int i = stackSlots[sp];
        if (m < 0)
        {
            out[n++] = m;
            continue;
        }
        const uint32_t choice = b.dpChoice[m];
        while (i > 1 && ((choice >> (21 + (i - 2))) & 1u))
        {
            --i;
        }
        if (i == 1)
        {
            out[n++] = m;
            continue;
        }
        const int2 ch = b.nodes[m].children;
        const int k = (int)(choice >> (3 * (i - 2))) & 7;
        stackRef[sp] = ch.y, stackSlots[sp++] = i - k;
        stackRef[sp] = ch.x, stackSlots[sp++] = k;
    }
    return n;
}

// Smallest biased exponent byte e such that lo + 255 * 2^(e-127) >= hi (evaluated in float, as traversal decodes it)
VKRT_HD uint32_t quantExponent(float lo, float hi)
{
    const float extent = hi - lo;
    if (!(extent > 0.0f))
    {
        return 1u;
    }
    int ex;
    frexpf(extent / 255.0f, &ex); // extent/255 = m * 2^ex, m in [0.5,1) -> 2^ex >= extent/255
    int e = ex + 127;
    // <= 239: traversal adds 15 to the exponent byte (traverse.cuh, intersectNode); extents above 2^120 are out of range
    e = e < 1 ? 1 : (e > 239 ? 239 : e);
    while (e < 239 && !(lo + 255.0f * uintAsFloat((uint32_t)e << 23) >= hi))
    {
        ++e;
    }
    return (uint32_t)e;
}

// Scale exponents of the three axes of one node.  A flat (or nearly flat) axis would get a scale many orders of
// magnitude below the others; its exponent is raised to at most 4 below the largest one.  A perfectly flat axis still
// quantises exactly (q = 0 for both planes), a nearly flat one grows by less than 1/4096 of the node's largest extent,
// and the traversal's packed-half node test (traverse.cuh) sees plane spacings of one node within a factor 16 of
// each other.  Returns the largest exponent.
VKRT_HD uint32_t quantExponents(F3 lo, F3 hi, uint32_t& ex, uint32_t& ey, uint32_t& ez)
{
    ex = quantExponent(lo.x, hi.x), ey = quantExponent(lo.y, hi.y), ez = quantExponent(lo.z, hi.z);
    uint32_t em = ex > ey ? ex : ey;
    em = em > ez ? em : ez;
    const uint32_t floorE = em > 5u ? em - 4u : 1u;
    ex = ex < floorE ? floorE : ex;
    ey = ey < floorE ? floorE : ey;
    ez = ez < floorE ? floorE : ez;
    return em;
}
// The power of two by which a hierarchy's box-test distances are scaled (traverse.cuh::setBoxRay), from the smallest scale
// exponent of any of its nodes.
VKRT_HD float boxScaleFromExponent(uint32_t smallestExponent)
{
    const uint32_t e = smallestExponent < 1u ? 1u : (smallestExponent > 240u ? 240u : smallestExponent);
    return uintAsFloat((241u - e) << 23);
}

VKRT_HD uint32_t quantLo(float v, float origin, float scale)
{
    float q = floorf((v - origin) / scale);
    q = fminf(fmaxf(q, 0.0f), 255.0f);
    uint32_t qi = (uint32_t)q;
    while (qi > 0 && origin + (float)qi * scale > v)
    {
        --qi;
    }
    return qi;
}
VKRT_HD uint32_t quantHi(float v, float origin, float scale)
{
    float q = ceilf((v - origin) / scale);
    q = fminf(fmaxf(q, 0.0f), 255.0f);
    uint32_t qi = (uint32_t)q;
    while (qi < 255 && origin + (float)qi * scale < v)
    {
        ++qi;
    }
    return qi;
}

// Builds wide node `wideIndex` from the binary subtree taskRef[wideIndex].
// `sahLocal` (device kernels): per-thread accumulators {node area, leaf area * count} that the caller reduces per warp;
// when NULL the areas go straight to a.sahAccum.
VKRT_HD void collapseNode(const CollapseArgs& a, uint32_t wideIndex, float* sahLocal = nullptr)
{
    const Bvh2& b = a.bvh2;
    const int rootRef = a.taskRef[wideIndex];
    const CollapseChild self = collapseFetch(b, rootRef);

    CollapseChild c[8];
    int n = 0;
    if (rootRef < 0 || self.count <= b.maxLeaf)
    {
        // whole BLAS has <= 3 triangles: the root holds one leaf child
        c[0] = self;
        n = 1;
    }
    else
    {
        if (b.dpChoice)
        {
            int refs[8];
            n = dpExpand(b, rootRef, refs);
            for (int k = 0; k < n; ++k)
            {
                c[k] = collapseFetch(b, refs[k]);
            }
        }
        else
        {
            const int2 rc = binChildren(b, rootRef);
            c[0] = collapseFetch(b, rc.x);
            c[1] = collapseFetch(b, rc.y);
            n = 2;
        }
        while (n < 8 && !b.dpChoice)
        {
            int pick = -1;
            float bestArea = -1.0f;
            for (int k = 0; k < n; ++k)
            {
                if (c[k].count > b.maxLeaf)
                {
                    const float ar = halfArea(c[k].lo, c[k].hi);
                    if (ar > bestArea)
                    {
                        bestArea = ar;
                        pick = k;
                    }
                }
            }
            if (pick < 0)
            {
                break;
            }
            const int2 ch = binChildren(b, c[pick].ref);
            c[pick] = collapseFetch(b, ch.x);
            c[n++] = collapseFetch(b, ch.y);
        }
        // Spare slots: traversal tests all 8 slot boxes of a node anyway, so multi-triangle leaf slots are opened too
        // (largest box first) until the node is full - tighter boxes, fewer triangle tests, same number of nodes.
        while (n < 8)
        {
            int pick = -1;
            float bestArea = -1.0f;
            for (int k = 0; k < n; ++k)
            {
                if (c[k].count > 1 && c[k].count <= b.maxLeaf)
                {
                    const float ar = halfArea(c[k].lo, c[k].hi);
                    if (ar > bestArea)
                    {
                        bestArea = ar;
                        pick = k;
                    }
                }
            }
            if (pick < 0)
            {
                break;
            }
            const int2 ch = binChildren(b, c[pick].ref);
            c[pick] = collapseFetch(b, ch.x);
            c[n++] = collapseFetch(b, ch.y);
        }
    }

    // octant-ordered slot assignment: slot s suits children lying towards corner (s&1 ? +x : -x, ...)
    const F3 centre = f3(0.5f * (self.lo.x + self.hi.x), 0.5f * (self.lo.y + self.hi.y), 0.5f * (self.lo.z + self.hi.z));
    int slotOf[8];
    bool slotUsed[8];
    bool childDone[8];
    for (int k = 0; k < 8; ++k)
    {
        slotOf[k] = -1;
        slotUsed[k] = false;
        childDone[k] = false;
    }
    for (int iter = 0; iter < n; ++iter)
    {
        float bestCost = -INFINITY;
        int bc = -1, bs = -1;
        for (int k = 0; k < n; ++k)
        {
            if (childDone[k])
            {
                continue;
            }
            const float cx = 0.5f * (c[k].lo.x + c[k].hi.x) - centre.x;
            const float cy = 0.5f * (c[k].lo.y + c[k].hi.y) - centre.y;
            const float cz = 0.5f * (c[k].lo.z + c[k].hi.z) - centre.z;
            for (int s = 0; s < 8; ++s)
            {
                if (slotUsed[s])
                {
                    continue;
                }
                const float cost = ((s & 1) ? cx : -cx) + ((s & 2) ? cy : -cy) + ((s & 4) ? cz : -cz);
                if (cost > bestCost)
                {
                    bestCost = cost, bc = k, bs = s;
                }
            }
        }
        slotOf[bc] = bs;
        slotUsed[bs] = true;
        childDone[bc] = true;
    }
    int childInSlot[8];
    for (int s = 0; s < 8; ++s)
    {
        childInSlot[s] = -1;
    }
    for (int k = 0; k < n; ++k)
    {
        childInSlot[slotOf[k]] = k;
    }

    uint32_t innerCount = 0, primCount = 0;
    for (int k = 0; k < n; ++k)
    {
        if (c[k].count > b.maxLeaf)
        {
            ++innerCount;
        }
        else
        {
            primCount += c[k].count;
        }
    }
    const uint32_t childBase = innerCount ? VKRT_ATOMIC_ADD(a.nodeCounter, innerCount) : 0u;
    const uint32_t primBase = primCount ? VKRT_ATOMIC_ADD(a.primCounter, primCount) : 0u;
    if (childBase + innerCount > a.nodeCapacity)
    {
        *a.overflow = 1u;
        return;
    }

    Node8 node;
    node.px = self.lo.x, node.py = self.lo.y, node.pz = self.lo.z;
    uint32_t ex, ey, ez;
    const uint32_t em = quantExponents(self.lo, self.hi, ex, ey, ez);
    node.sx = packScaleExponent(ex), node.sy = packScaleExponent(ey), node.sz = packScaleExponent(ez);
    const float sx = uintAsFloat(ex << 23), sy = uintAsFloat(ey << 23), sz = uintAsFloat(ez << 23);
    node.imask = 0;
    node.childBase = childBase;
    node.primBase = primBase;
    node.leafBits = 0;
    (void)em;
    node.reserved = 0;
    uint32_t innerSeen = 0, primOffset = 0;
    float leafArea = 0.0f;
    for (int s = 0; s < 8; ++s)
    {
        const int k = childInSlot[s];
        if (k < 0)
        {
            node.qlox[s] = node.qloy[s] = node.qloz[s] = 0;
            node.qhix[s] = node.qhiy[s] = node.qhiz[s] = 0;
            continue;
        }
        node.qlox[s] = (uint8_t)quantLo(c[k].lo.x, self.lo.x, sx);
        node.qloy[s] = (uint8_t)quantLo(c[k].lo.y, self.lo.y, sy);
        node.qloz[s] = (uint8_t)quantLo(c[k].lo.z, self.lo.z, sz);
        node.qhix[s] = (uint8_t)quantHi(c[k].hi.x, self.lo.x, sx);
        node.qhiy[s] = (uint8_t)quantHi(c[k].hi.y, self.lo.y, sy);
        node.qhiz[s] = (uint8_t)quantHi(c[k].hi.z, self.lo.z, sz);
        if (c[k].count > b.maxLeaf)
        {
            node.imask |= (uint8_t)(1u << s);
            a.taskRef[childBase + innerSeen] = c[k].ref;
            ++innerSeen;
        }
        else
        {
            const uint32_t unary = (1u << c[k].count) - 1u; // 1 -> 001, 2 -> 011, 3 -> 111
            node.leafBits |= unary << (3 * s);
            uint32_t leaves[4];
            collectLeaves(b, c[k].ref, leaves);
            for (uint32_t j = 0; j < c[k].count; ++j)
            {
                a.primSource[primBase + primOffset + j] = b.vals[leaves[j]];
            }
            primOffset += c[k].count;
            leafArea += halfArea(c[k].lo, c[k].hi) * (float)c[k].count;
        }
    }
    a.nodes[wideIndex] = node;
    if (sahLocal)
    {
        sahLocal[0] += halfArea(self.lo, self.hi);
        sahLocal[1] += leafArea;
    }
    else if (a.sahAccum)
    {
        VKRT_ATOMIC_ADD(&a.sahAccum[0], halfArea(self.lo, self.hi));
        VKRT_ATOMIC_ADD(&a.sahAccum[1], leafArea);
    }
}

VKRT_HD PackedTri packTriangle(const GeometryTable& g, const uint8_t* vertices, uint32_t stride, const uint32_t* indices,
                               uint32_t buildIndex)
{
    F3 v0, v1, v2;
    PackedTri t;
    fetchTriangle(g, vertices, stride, indices, buildIndex, v0, v1, v2, t.geometry, t.primitive);
    t.v0x = v0.x, t.v0y = v0.y, t.v0z = v0.z;
    t.v1x = v1.x, t.v1y = v1.y, t.v1z = v1.z;
    t.v2x = v2.x, t.v2y = v2.y, t.v2z = v2.z;
    t.buildIndex = buildIndex;
    return t;
}

} // namespace vkrt
