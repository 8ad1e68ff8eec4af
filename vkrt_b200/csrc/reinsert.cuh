// reinsert.cuh — parallel reinsertion on the binary hierarchy (after Meister & Bittner 2018, "Parallel Reinsertion for
// Bounding Volume Hierarchy Optimization"), the optional quality pass of the builder (VKRT_BUILD_PREFER_FAST_TRACE).
// No counterpart in the reference's sources: there the driver behind vkCmdBuildAccelerationStructuresKHR
// (src/Raytracer.cpp:1089-1157) decides how much to spend on hierarchy quality.
//
// One pass = five kernels over the 64-byte binary nodes (builder.cuh), all reading one consistent snapshot:
//   search  one thread per node L (leaves and inner nodes): where else in the tree would L, taken out together with its
//           parent P, cost the least surface area?  Walk from L to the root; at every ancestor the sibling subtree is
//           searched depth-first with branch and bound (the area L adds on the way down only grows).
//   lock    every L with a gain takes the six nodes its move rewrites (L, P, sibling S, grandparent G, target X, X's
//           parent) with a 64-bit atomicMax of (gain, L): the larger gain wins.  The winner per node does not depend on
//           the order the threads run in, so the result is deterministic.  Two moves that hold disjoint locks could still
//           put two subtrees below each other (L1 next to a node inside L2 and L2 next to a node inside L1: a cycle cut off
//           from the root).  The search therefore only accepts targets with a STRICTLY larger surface area than L: a
//           cycle would need area(L1) < area(X1) <= area(L2) < area(X2) <= area(L1).
//   verify  a move goes ahead only if it holds all of its locks.
//   apply   S takes P's place under G; P is put between X and X's parent, with children (X, L).
//   refit   boxes and leaf counts of all inner nodes, bottom-up with arrival counters.
// Functions are host+device: tests/emu runs the same passes sequentially.
#pragma once

#include "builder.cuh"

namespace vkrt
{

constexpr int kReinsertNoMove = (int)0x80000000u; // not a reference: ~leaf would need leaf = 2^31 - 1
constexpr int kReinsertStack = 64;
// Schedule (chosen on the Sponza stand-in with tools/hierarchy_lab.py): sweeps over the area bands, each band - nodes whose
// area is within a factor kReinsertBandRatio - a few times in a row because moves that touch the same nodes take turns.
constexpr int kReinsertSweeps = 2, kReinsertBands = 12, kReinsertIterations = 3;
constexpr float kReinsertBandRatio = 4.0f;

struct ReinsertArgs
{
    Bvh2 bvh; // nodes (64-byte records), vals, triBox, count, root, dpParent (parents, slot layout), flags (arrival counters)
    int* target;               // per slot: reference of the node X that L is put next to, kReinsertNoMove if L stays
    float* gain;               // per slot: surface area saved by the move
    unsigned long long* lock;  // per slot: largest (gain, slot) key that wants this node
    uint32_t* moved;           // number of moves applied in this pass
    float areaLo, areaHi;      // only nodes L with areaLo <= area(L) < areaHi are considered in this pass (large nodes first)
};

// slots: leaves [0, count), inner nodes [count, 2 count - 1) - the layout of Bvh2::dpParent
VKRT_HD uint32_t reinsertSlot(const Bvh2& b, int ref) { return ref < 0 ? (uint32_t)~ref : b.count + (uint32_t)ref; }
VKRT_HD int reinsertRef(const Bvh2& b, uint32_t slot) { return slot < b.count ? ~(int)slot : (int)(slot - b.count); }

VKRT_HD void reinsertBox(const Bvh2& b, int ref, F3& lo, F3& hi)
{
    float4 l, h;
    if (ref < 0)
    {
        const uint32_t t = b.vals[~ref];
        l = b.triBox[2 * (size_t)t], h = b.triBox[2 * (size_t)t + 1];
    }
    else
    {
        l = b.nodes[ref].lo, h = b.nodes[ref].hi;
    }
    lo = f3(l.x, l.y, l.z), hi = f3(h.x, h.y, h.z);
}

VKRT_HD float reinsertMergedArea(F3 alo, F3 ahi, F3 blo, F3 bhi)
{
    return halfArea(f3(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z)),
                    f3(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z)));
}

VKRT_HD int reinsertSibling(const Bvh2& b, int parent, int ref)
{
    const int2 ch = b.nodes[parent].children;
    return ch.x == ref ? ch.y : ch.x;
}

struct ReinsertBest
{
    float gain;
    int target;
};

// Depth-first search of the subtree under `start` for the position that costs L the least.  `induced` = area already
// added to the nodes between `start` and the common ancestor; `base` = area the removal of L has saved up to there.
VKRT_HD void reinsertDescend(const Bvh2& b, int start, float induced, bool startIsCandidate, float base, F3 llo, F3 lhi,
                             float areaL, ReinsertBest& best)
{
    int stackRef[kReinsertStack];
    float stackInduced[kReinsertStack];
    int sp = 0;
    stackRef[sp] = start, stackInduced[sp] = induced, ++sp;
    bool candidate = startIsCandidate; // only the first entry (the start) may be excluded
    while (sp > 0)
    {
        --sp;
        const int y = stackRef[sp];
        const float c = stackInduced[sp];
        const bool isCandidate = candidate;
        candidate = true;
        if (base - c - areaL <= best.gain)
        {
            continue; // pushed before a better position was found
        }
        F3 ylo, yhi;
        reinsertBox(b, y, ylo, yhi);
        const float direct = reinsertMergedArea(ylo, yhi, llo, lhi);
        const float g = base - c - direct;
        const float areaY = halfArea(ylo, yhi);
        if (isCandidate && areaY > areaL && g > best.gain) // only next to a LARGER node (no cycles, see the header)
        {
            best.gain = g, best.target = y;
        }
        if (y >= 0)
        {
            const float below = c + direct - areaY;
            if (base - below - areaL > best.gain && sp + 2 <= kReinsertStack)
            {
                const int2 ch = b.nodes[y].children;
                stackRef[sp] = ch.y, stackInduced[sp] = below, ++sp;
                stackRef[sp] = ch.x, stackInduced[sp] = below, ++sp;
            }
        }
    }
}

// One thread per slot.
VKRT_HD void reinsertSearch(const ReinsertArgs& a, uint32_t slot)
{
    const Bvh2& b = a.bvh;
    const uint32_t n = b.count;
    a.target[slot] = kReinsertNoMove;
    a.lock[slot] = 0ull;
    const int L = reinsertRef(b, slot);
    const int P = b.dpParent[slot];
    if (P < 0)
    {
        return; // the root
    }
    const int G = b.dpParent[n + (uint32_t)P];
    if (G < 0)
    {
        return; // children of the root stay (taking one out would move the root)
    }
    F3 llo, lhi, shrLo, shrHi, plo, phi;
    reinsertBox(b, L, llo, lhi);
    const float areaL = halfArea(llo, lhi);
    if (!(areaL >= a.areaLo && areaL < a.areaHi))
    {
        return;
    }
    reinsertBox(b, P, plo, phi);
    const float areaP = halfArea(plo, phi);
    const int S = reinsertSibling(b, P, L);
    reinsertBox(b, S, shrLo, shrHi);
    ReinsertBest best;
    best.gain = 0.0f, best.target = kReinsertNoMove;
    // below the sibling (the sibling itself is where L is now)
    float base = areaP;
    reinsertDescend(b, S, 0.0f, false, base, llo, lhi, areaL, best);
    // up the path: the sibling subtree of every ancestor, then the ancestor itself, shrunk by the removal
    int prev = P, A = G;
    while (A >= 0)
    {
        const int C = reinsertSibling(b, A, prev);
        reinsertDescend(b, C, 0.0f, true, base, llo, lhi, areaL, best);
        F3 clo, chi, alo, ahi;
        reinsertBox(b, C, clo, chi);
        shrLo = f3(fminf(shrLo.x, clo.x), fminf(shrLo.y, clo.y), fminf(shrLo.z, clo.z));
        shrHi = f3(fmaxf(shrHi.x, chi.x), fmaxf(shrHi.y, chi.y), fmaxf(shrHi.z, chi.z));
        reinsertBox(b, A, alo, ahi);
        const float dec = halfArea(alo, ahi) - halfArea(shrLo, shrHi);
        const int GA = b.dpParent[n + (uint32_t)A];
        if (GA >= 0)
        {
            const float g = base + dec - reinsertMergedArea(shrLo, shrHi, llo, lhi);
            if (halfArea(shrLo, shrHi) > areaL && g > best.gain)
            {
                best.gain = g, best.target = A;
            }
        }
        base += dec;
        prev = A, A = GA;
    }
    // moves that save less than a millionth of L's own area are rounding noise
    if (best.target != kReinsertNoMove && best.gain > 1.0e-6f * areaL)
    {
        a.target[slot] = best.target, a.gain[slot] = best.gain;
    }
}

#if defined(__CUDA_ARCH__)
#define VKRT_ATOMIC_MAX_U64(ptr, v) atomicMax((ptr), (v))
#else
#define VKRT_ATOMIC_MAX_U64(ptr, v) (*(ptr) = (*(ptr) < (v) ? (v) : *(ptr)))
#endif

// Visits every node the move of `slot` rewrites or must keep in place; LOCK: take them, else: check that all are held.
template<bool LOCK>
VKRT_HD bool reinsertWalkLocks(const ReinsertArgs& a, uint32_t slot)
{
    const Bvh2& b = a.bvh;
    const uint32_t n = b.count;
    const unsigned long long key = ((unsigned long long)floatAsUint(a.gain[slot]) << 32) | (unsigned long long)slot;
    const int L = reinsertRef(b, slot);
    const int P = b.dpParent[slot];
    const int G = b.dpParent[n + (uint32_t)P];
    const int S = reinsertSibling(b, P, L);
    const int X = a.target[slot];
    const int GX = b.dpParent[reinsertSlot(b, X)];
    bool ok = true;
    const uint32_t fixedSlots[6] = {slot, n + (uint32_t)P, reinsertSlot(b, S), n + (uint32_t)G, reinsertSlot(b, X), n + (uint32_t)GX};
    for (int k = 0; k < 6; ++k)
    {
        if (LOCK)
        {
            VKRT_ATOMIC_MAX_U64(&a.lock[fixedSlots[k]], key);
        }
        else
        {
            ok = ok && a.lock[fixedSlots[k]] == key;
        }
    }
    return ok;
}

VKRT_HD void reinsertLock(const ReinsertArgs& a, uint32_t slot)
{
    if (a.target[slot] != kReinsertNoMove)
    {
        reinsertWalkLocks<true>(a, slot);
    }
}

VKRT_HD void reinsertVerify(const ReinsertArgs& a, uint32_t slot)
{
    if (a.target[slot] != kReinsertNoMove && !reinsertWalkLocks<false>(a, slot))
    {
        a.target[slot] = kReinsertNoMove;
    }
}

// Only nodes whose lock this move holds are read or written.
VKRT_HD void reinsertApply(const ReinsertArgs& a, uint32_t slot)
{
    const int X = a.target[slot];
    if (X == kReinsertNoMove)
    {
        return;
    }
    const Bvh2& b = a.bvh;
    const uint32_t n = b.count;
    const int L = reinsertRef(b, slot);
    const int P = b.dpParent[slot];
    const int G = b.dpParent[n + (uint32_t)P];
    const int S = reinsertSibling(b, P, L);
    const int GX = b.dpParent[reinsertSlot(b, X)];
    // S takes P's place
    if (b.nodes[G].children.x == P)
    {
        b.nodes[G].children.x = S;
    }
    else
    {
        b.nodes[G].children.y = S;
    }
    b.dpParent[reinsertSlot(b, S)] = G;
    // P goes between GX and X
    if (b.nodes[GX].children.x == X)
    {
        b.nodes[GX].children.x = P;
    }
    else
    {
        b.nodes[GX].children.y = P;
    }
    b.dpParent[n + (uint32_t)P] = GX;
    b.nodes[P].children = make_int2(X, L);
    b.dpParent[reinsertSlot(b, X)] = P;
    VKRT_ATOMIC_ADD(a.moved, 1u);
}

// One thread per sorted leaf: boxes and leaf counts of the inner nodes, bottom-up.  b.flags must be zero on entry.
VKRT_HD void reinsertRefit(const Bvh2& b, uint32_t leaf)
{
    int node = b.dpParent[leaf];
    while (node >= 0)
    {
        if (VKRT_ATOMIC_ADD_ACQ_REL(&b.flags[node], 1) == 0)
        {
            return; // the thread that arrives second finds both children finished
        }
        const int2 ch = b.nodes[node].children;
        float4 lo[2], hi[2];
        uint32_t count = 0;
        for (int k = 0; k < 2; ++k)
        {
            const int c = k ? ch.y : ch.x;
            if (c < 0)
            {
                const uint32_t t = b.vals[~c];
                lo[k] = b.triBox[2 * (size_t)t], hi[k] = b.triBox[2 * (size_t)t + 1];
                count += 1u;
            }
            else
            {
                lo[k] = VKRT_LOAD_CG_F4(&b.nodes[c].lo), hi[k] = VKRT_LOAD_CG_F4(&b.nodes[c].hi);
                count += floatAsUint(lo[k].w);
            }
        }
        b.nodes[node].lo = make_float4(fminf(lo[0].x, lo[1].x), fminf(lo[0].y, lo[1].y), fminf(lo[0].z, lo[1].z), uintAsFloat(count));
        b.nodes[node].hi = make_float4(fmaxf(hi[0].x, hi[1].x), fmaxf(hi[0].y, hi[1].y), fmaxf(hi[0].z, hi[1].z), 0.0f);
        node = b.dpParent[b.count + (uint32_t)node];
    }
}

} // namespace vkrt
