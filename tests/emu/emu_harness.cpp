/*
 * emu_harness.cpp — NO-GPU UNIT-TEST AID.  Not part of the product; nothing in libvkrt_b200.so links or calls it.
 *
 * The kernel LOGIC of libvkrt_b200 lives in `__host__ __device__` headers (vkrt_b200/csrc/*.cuh).  This file compiles
 * those headers with g++ and drives them sequentially — one loop iteration per CUDA thread — so that the builder,
 * the traversal and the shading restatement can be checked against the CPU oracle in the `-m "not gpu"` test tier,
 * before any GPU time is spent.  Warp-level code (radix sort, queue compaction) is not emulated: the sort is replaced
 * by std::stable_sort and the wavefront queues by a per-pixel loop; both are covered by the `-m gpu` tests.
 */
#include "../../vkrt_b200/csrc/builder.cuh"
#include "../../vkrt_b200/csrc/reinsert.cuh"
#include "../../vkrt_b200/csrc/shading.cuh"
#include "../../vkrt_b200/csrc/raster.cuh"
#include "../../vkrt_b200/csrc/traverse.cuh"

#include "hierarchy_lab.hpp"

#include <algorithm>
#include <cfloat>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <numeric>
#include <vector>

using namespace vkrt;

struct EmuBlas
{
    std::vector<uint8_t> vertices;
    std::vector<uint32_t> indices;
    std::vector<uint32_t> firstTri;
    std::vector<uint64_t> indexBase;
    uint32_t stride = 0;
    uint32_t triangleCount = 0;
    std::vector<Node8> nodes;
    std::vector<PackedTri> tris;
    float boundsMin[3], boundsMax[3];
    float sahCost = 0.f;
    float boxScale = 1.f;
};

struct EmuScene
{
    EmuBlas* blas = nullptr;
    std::vector<InstanceRecord> instances;
    std::vector<Node8> topNodes;        // as vkrt_tlas_build makes them (instanceCount > 1)
    std::vector<uint32_t> leafInstance;
    float topBoxScale = 1.f;
    TlasRefs refs() const
    {
        TlasRefs t;
        t.nodes = topNodes.empty() ? nullptr : topNodes.data();
        t.leafInstance = leafInstance.data();
        t.instances = instances.data();
        t.instanceCount = (uint32_t)instances.size();
        t.cullMask = 0xFFu;
        t.boxScale = topBoxScale;
        return t;
    }
    std::vector<VkrtSubmeshInfo> rows;
    std::vector<std::vector<uint8_t>> texels;
    std::vector<TextureRecord> textures;
    std::vector<std::vector<uint8_t>> mipTexels; // levels 1.. packed, as textures.cu lays them out
    std::vector<TextureMips> mips;
    bool lod = false;
};

// k_min_scale_exponent + boxScaleFromExponent
static float emuBoxScale(const std::vector<Node8>& nodes)
{
    uint32_t e = 255u;
    for (const Node8& n : nodes)
    {
        if (n.imask != 0 || (n.leafBits & 0xFFFFFFu) != 0u)
        {
            e = std::min(e, std::min(std::min((uint32_t)n.sx >> 7, (uint32_t)n.sy >> 7), (uint32_t)n.sz >> 7));
        }
    }
    return boxScaleFromExponent(e == 255u ? 127u : e);
}

// The box-only part of the builder (bvh_build.cu::buildHierarchy), sequentially: Morton codes -> stable sort -> LBVH ->
// optional cost-optimal collapse decisions -> collapse into wide nodes.  Returns the node count, 0 on failure.
static uint32_t emuBuildHierarchy(const std::vector<float4>& triBox, uint32_t n, uint32_t maxLeaf, float dpPrimCost, const float clo[3],
                                  const float chi[3], std::vector<Node8>& nodesOut, uint32_t nodeCapacity, std::vector<uint32_t>& primSource,
                                  float sah[2], bool fastTrace = false)
{
    std::vector<uint32_t> keys(n), vals(n);
    const F3 clo3 = f3(clo[0], clo[1], clo[2]);
    const F3 inv = f3(chi[0] > clo[0] ? 1.0f / (chi[0] - clo[0]) : 0.f, chi[1] > clo[1] ? 1.0f / (chi[1] - clo[1]) : 0.f,
                      chi[2] > clo[2] ? 1.0f / (chi[2] - clo[2]) : 0.f);
    for (uint32_t i = 0; i < n; ++i)
    {
        const float4 l = triBox[2 * (size_t)i], h = triBox[2 * (size_t)i + 1];
        keys[i] = morton30(f3(0.5f * (l.x + h.x), 0.5f * (l.y + h.y), 0.5f * (l.z + h.z)), clo3, inv);
        vals[i] = i;
    }
    // radix sort stand-in (stable, like the GPU LSD sort)
    std::vector<uint32_t> perm(n);
    std::iota(perm.begin(), perm.end(), 0u);
    std::stable_sort(perm.begin(), perm.end(), [&](uint32_t a, uint32_t c) { return keys[a] < keys[c]; });
    std::vector<uint32_t> sk(n), sv(n);
    for (uint32_t i = 0; i < n; ++i)
    {
        sk[i] = keys[perm[i]], sv[i] = vals[perm[i]];
    }
    // k_lbvh
    std::vector<BinNode> binNodes(std::max(1u, n));
    std::vector<int> flags(std::max(1u, n), -1);
    int root = 0;
    Bvh2 bvh;
    bvh.nodes32 = nullptr, bvh.sortedBox = nullptr;
    bvh.nodes = binNodes.data(), bvh.flags = flags.data();
    bvh.keys = sk.data(), bvh.vals = sv.data(), bvh.triBox = triBox.data();
    bvh.count = n, bvh.root = &root;
    bvh.dpChoice = nullptr, bvh.dpParent = nullptr, bvh.dpPrimCost = 0.0f, bvh.maxLeaf = maxLeaf;
    if (n > 1)
    {
        for (uint32_t i = 0; i < n; ++i)
        {
            lbvhClimb(bvh, i);
        }
    }
    // experiments on the binary hierarchy (hierarchy_lab.hpp); the default leaves the LBVH as it is
    const char* labMode = getenv("VKRT_EMU_BINARY");
    const char* labReinsert = getenv("VKRT_EMU_REINSERT");
    if (n > 2 && ((labMode && strcmp(labMode, "lbvh") != 0) || labReinsert || getenv("VKRT_LAB_TOP")))
    {
        std::vector<lab::Box> leaves(n);
        for (uint32_t i = 0; i < n; ++i)
        {
            const float4 l = triBox[2 * (size_t)sv[i]], h = triBox[2 * (size_t)sv[i] + 1];
            leaves[i] = lab::Box{{l.x, l.y, l.z}, {h.x, h.y, h.z}};
        }
        lab::Tree t;
        t.init(n, leaves);
        auto fromRef = [&](int ref) { return ref >= 0 ? ref : (int)(n - 1) + ~ref; };
        if (labMode && strcmp(labMode, "ploc") == 0)
        {
            lab::buildPloc(t, 8);
        }
        else if (labMode && strcmp(labMode, "sah") == 0)
        {
            lab::buildSah(t);
        }
        else
        {
            for (uint32_t i = 0; i + 1 < n; ++i)
            {
                t.left[i] = fromRef(binNodes[i].children.x), t.right[i] = fromRef(binNodes[i].children.y);
                t.parent[t.left[i]] = (int)i, t.parent[t.right[i]] = (int)i;
                const float4 l = binNodes[i].lo, h = binNodes[i].hi;
                t.box[i] = lab::Box{{l.x, l.y, l.z}, {h.x, h.y, h.z}};
            }
            t.root = root, t.parent[root] = -1;
        }
        fprintf(stderr, "lab: %s binary SAH %.3f\n", labMode ? labMode : "lbvh", t.sah());
        if (getenv("VKRT_LAB_TOP"))
        {
            lab::rebuildTop(t, (uint32_t)atoi(getenv("VKRT_LAB_TOP")));
            fprintf(stderr, "lab: top rebuilt over %s subtrees, binary SAH %.3f\n", getenv("VKRT_LAB_TOP"), t.sah());
        }
        if (labReinsert)
        {
            lab::reinsert(t, atoi(labReinsert));
        }
        // back to BinNode records: boxes and leaf counts bottom-up
        std::vector<uint32_t> count(2 * (size_t)n - 1, 1u);
        std::vector<std::pair<int, int>> work;
        work.emplace_back(t.root, 0);
        while (!work.empty())
        {
            auto [node, state] = work.back();
            if (t.isLeaf(node))
            {
                work.pop_back();
            }
            else if (state == 0)
            {
                work.back().second = 1;
                work.emplace_back(t.left[node], 0);
                work.emplace_back(t.right[node], 0);
            }
            else
            {
                count[node] = count[t.left[node]] + count[t.right[node]];
                const lab::Box b = lab::merge(t.box[t.left[node]], t.box[t.right[node]]);
                auto toRef = [&](int id) { return t.isLeaf(id) ? ~(id - (int)(n - 1)) : id; };
                binNodes[node].lo = make_float4(b.lo[0], b.lo[1], b.lo[2], uintAsFloat(count[node]));
                binNodes[node].hi = make_float4(b.hi[0], b.hi[1], b.hi[2], 0.0f);
                binNodes[node].children = make_int2(toRef(t.left[node]), toRef(t.right[node]));
                t.box[node] = b;
                work.pop_back();
            }
        }
        root = t.root;
    }
    // the builder's reinsertion passes (reinsert.cuh), phase by phase like the kernels, on the schedule of
    // bvh_build.cu::buildHierarchy (VKRT_BUILD_PREFER_FAST_TRACE); VKRT_EMU_FAST_TRACE=1 turns them on for experiments
    if (n > 3 && (fastTrace || getenv("VKRT_EMU_FAST_TRACE")))
    {
        std::vector<int> parent(2 * (size_t)n - 1, -1), target(2 * (size_t)n - 1);
        std::vector<float> gain(2 * (size_t)n - 1);
        std::vector<unsigned long long> lock(2 * (size_t)n - 1);
        uint32_t moved = 0;
        bvh.dpParent = parent.data();
        for (uint32_t i = 0; i + 1 < n; ++i)
        {
            dpLinkParents(bvh, i);
        }
        parent[n + (uint32_t)root] = -1;
        ReinsertArgs ra;
        ra.bvh = bvh, ra.target = target.data(), ra.gain = gain.data(), ra.lock = lock.data(), ra.moved = &moved;
        const uint32_t slots = 2 * n - 1;
        const BinNode& rootNode = binNodes[root];
        const float rootArea = halfArea(f3(rootNode.lo.x, rootNode.lo.y, rootNode.lo.z), f3(rootNode.hi.x, rootNode.hi.y, rootNode.hi.z));
        int sweeps = kReinsertSweeps, bands = kReinsertBands, iters = kReinsertIterations;
        float ratio = kReinsertBandRatio;
        if (const char* e = getenv("VKRT_BUILD_REINSERT"))
        {
            sscanf(e, "%d,%d,%d,%f", &sweeps, &bands, &iters, &ratio);
        }
        for (int sweep = 0; sweep < sweeps; ++sweep)
        {
            for (int band = 0; band < bands; ++band)
            {
                // the same float expressions as the host loop and k_reinsert_search
                const float hiMul = band == 0 ? FLT_MAX : powf(ratio, -(float)band);
                const float loMul = band == bands - 1 ? 0.0f : powf(ratio, -(float)(band + 1));
                ra.areaHi = hiMul * rootArea, ra.areaLo = loMul * rootArea;
                for (int it = 0; it < iters; ++it)
                {
                    for (uint32_t i = 0; i < slots; ++i)
                    {
                        reinsertSearch(ra, i);
                    }
                    for (uint32_t i = 0; i < slots; ++i)
                    {
                        reinsertLock(ra, i);
                    }
                    for (uint32_t i = 0; i < slots; ++i)
                    {
                        reinsertVerify(ra, i);
                    }
                    for (uint32_t i = 0; i < slots; ++i)
                    {
                        reinsertApply(ra, i);
                    }
                    std::fill(flags.begin(), flags.end(), 0);
                    for (uint32_t i = 0; i < n; ++i)
                    {
                        reinsertRefit(bvh, i);
                    }
                }
            }
        }
        if (getenv("VKRT_EMU_VERBOSE"))
        {
            fprintf(stderr, "emu: reinsertion moved %u nodes in %d passes\n", moved, sweeps * bands * iters);
        }
        bvh.dpParent = nullptr;
    }
    // cost-optimal collapse: children before parents (the GPU climbs from the leaves with arrival counters)
    std::vector<uint32_t> dpChoice;
    if (n > 1 && n > maxLeaf && dpPrimCost > 0.0f)
    {
        dpChoice.assign(n, 0u);
        bvh.dpChoice = dpChoice.data(), bvh.dpPrimCost = dpPrimCost;
        std::vector<std::pair<int, int>> work; // (node, state)
        work.emplace_back(root, 0);
        while (!work.empty())
        {
            auto [node, state] = work.back();
            if (state == 0)
            {
                work.back().second = 1;
                const int2 ch = binNodes[node].children;
                if (ch.x >= 0)
                {
                    work.emplace_back(ch.x, 0);
                }
                if (ch.y >= 0)
                {
                    work.emplace_back(ch.y, 0);
                }
            }
            else
            {
                dpSolve(bvh, node);
                work.pop_back();
            }
        }
    }
    // collapse
    std::vector<int> taskRef(nodeCapacity);
    primSource.assign(n, 0u);
    uint32_t nodeCounter = 1, primCounter = 0, overflow = 0;
    CollapseArgs ca;
    ca.bvh2 = bvh, ca.nodes = nodesOut.data(), ca.taskRef = taskRef.data(), ca.primSource = primSource.data();
    ca.nodeCounter = &nodeCounter, ca.primCounter = &primCounter, ca.sahAccum = sah, ca.nodeCapacity = nodeCapacity;
    ca.overflow = &overflow;
    taskRef[0] = n > 1 ? root : ~0;
    uint32_t first = 0, last = 1;
    while (first < last)
    {
        for (uint32_t i = first; i < last; ++i)
        {
            collapseNode(ca, i);
        }
        first = last;
        last = nodeCounter;
    }
    if (overflow || primCounter != n)
    {
        fprintf(stderr, "emuBuildHierarchy: collapse failed (overflow=%u packed=%u of %u)\n", overflow, primCounter, n);
        return 0;
    }
    return last;
}

extern "C" {

EmuBlas* emu_blas_build_flags(const void* vertices, size_t vertexBytes, uint32_t stride, const uint32_t* indices, size_t indexCount,
                              const VkrtGeometryRange* ranges, uint32_t geometryCount, uint32_t flags);

EmuBlas* emu_blas_build(const void* vertices, size_t vertexBytes, uint32_t stride, const uint32_t* indices, size_t indexCount,
                        const VkrtGeometryRange* ranges, uint32_t geometryCount)
{
    return emu_blas_build_flags(vertices, vertexBytes, stride, indices, indexCount, ranges, geometryCount, 0u);
}

// flags: VkrtBuildFlags; only VKRT_BUILD_PREFER_FAST_TRACE changes anything here (the binary hierarchy is always the LBVH)
EmuBlas* emu_blas_build_flags(const void* vertices, size_t vertexBytes, uint32_t stride, const uint32_t* indices, size_t indexCount,
                              const VkrtGeometryRange* ranges, uint32_t geometryCount, uint32_t flags)
{
    EmuBlas* b = new EmuBlas();
    b->vertices.assign((const uint8_t*)vertices, (const uint8_t*)vertices + vertexBytes);
    b->indices.assign(indices, indices + indexCount);
    b->stride = stride;
    b->firstTri.resize(geometryCount + 1);
    b->indexBase.resize(std::max(geometryCount, 1u));
    uint32_t n = 0;
    for (uint32_t g = 0; g < geometryCount; ++g)
    {
        b->firstTri[g] = n;
        b->indexBase[g] = ranges[g].indexByteOffset / 4;
        n += ranges[g].primitiveCount;
    }
    b->firstTri[geometryCount] = n;
    b->triangleCount = n;

    GeometryTable geo;
    geo.firstTri = b->firstTri.data();
    geo.indexBase = b->indexBase.data();
    geo.geometryCount = geometryCount;

    const uint32_t nodeCapacity = getenv("VKRT_EMU_MAX_LEAF") ? n + 8u : std::max(1u, n / 2u + 8u);
    b->nodes.assign(nodeCapacity, Node8{});
    b->tris.assign(std::max(1u, n), PackedTri{});
    if (n == 0)
    {
        Node8 root;
        memset(&root, 0, sizeof(root));
        root.sx = root.sy = root.sz = packScaleExponent(1);
        b->nodes[0] = root;
        b->nodes.resize(1);
        for (int k = 0; k < 3; ++k)
        {
            b->boundsMin[k] = b->boundsMax[k] = 0.f;
        }
        return b;
    }

    // k_tri_bounds
    std::vector<float4> triBox(2 * (size_t)n);
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = 0; i < n; ++i)
    {
        F3 v0, v1, v2;
        uint32_t g, p;
        fetchTriangle(geo, b->vertices.data(), stride, b->indices.data(), i, v0, v1, v2, g, p);
        const float4 l = make_float4(fminf(fminf(v0.x, v1.x), v2.x), fminf(fminf(v0.y, v1.y), v2.y), fminf(fminf(v0.z, v1.z), v2.z), 0.f);
        const float4 h = make_float4(fmaxf(fmaxf(v0.x, v1.x), v2.x), fmaxf(fmaxf(v0.y, v1.y), v2.y), fmaxf(fmaxf(v0.z, v1.z), v2.z), 0.f);
        triBox[2 * (size_t)i] = l, triBox[2 * (size_t)i + 1] = h;
        const float lv[3] = {l.x, l.y, l.z}, hv[3] = {h.x, h.y, h.z};
        for (int k = 0; k < 3; ++k)
        {
            const float c = 0.5f * (lv[k] + hv[k]);
            lo[k] = fminf(lo[k], lv[k]), hi[k] = fmaxf(hi[k], hv[k]);
            clo[k] = fminf(clo[k], c), chi[k] = fmaxf(chi[k], c);
        }
    }
    for (int k = 0; k < 3; ++k)
    {
        b->boundsMin[k] = lo[k], b->boundsMax[k] = hi[k];
    }
    const char* dpEnv = getenv("VKRT_EMU_DP_PRIM_COST");
    const float dpPrimCost = dpEnv ? (float)atof(dpEnv) : 0.5f;
    std::vector<uint32_t> primSource;
    float sah[2] = {0.f, 0.f};
    const uint32_t maxLeaf = getenv("VKRT_EMU_MAX_LEAF") ? (uint32_t)atoi(getenv("VKRT_EMU_MAX_LEAF")) : kMaxLeafTris; // experiments
    const uint32_t last = emuBuildHierarchy(triBox, n, maxLeaf, dpPrimCost, clo, chi, b->nodes, nodeCapacity, primSource, sah,
                                            (flags & VKRT_BUILD_PREFER_FAST_TRACE) != 0);
    if (last == 0)
    {
        delete b;
        return nullptr;
    }
    b->nodes.resize(last);
    b->boxScale = emuBoxScale(b->nodes);
    for (uint32_t i = 0; i < n; ++i)
    {
        b->tris[i] = packTriangle(geo, b->vertices.data(), stride, b->indices.data(), primSource[i]);
    }
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    const float ra = dx * dy + dy * dz + dz * dx;
    b->sahCost = ra > 0.f ? (sah[0] + sah[1]) / ra : 0.f;
    return b;
}

// pathToPixel of shading.cuh for one (rank, size): x, y, pixel-in-tile per path; returns the number of valid paths
uint32_t emu_path_to_pixel(uint32_t width, uint32_t height, uint32_t rank, uint32_t size, uint32_t paths, uint32_t* x, uint32_t* y,
                           uint32_t* pit)
{
    TileMap m;
    m.width = width, m.height = height;
    m.tilesX = (width + kTile - 1) / kTile, m.tilesY = (height + kTile - 1) / kTile;
    m.rank = rank, m.size = size;
    const uint32_t tiles = m.tilesX * m.tilesY;
    m.ownedTiles = tiles > rank ? (tiles - rank + size - 1) / size : 0u;
    uint32_t valid = 0;
    for (uint32_t p = 0; p < paths; ++p)
    {
        uint32_t tl;
        const bool ok = pathToPixel(m, p, x[p], y[p], tl, pit[p]);
        if (!ok)
        {
            x[p] = y[p] = 0xFFFFFFFFu;
        }
        valid += ok ? 1u : 0u;
    }
    return valid;
}

void emu_blas_destroy(EmuBlas* b) { delete b; }
uint32_t emu_blas_node_count(const EmuBlas* b) { return (uint32_t)b->nodes.size(); }
uint32_t emu_blas_triangle_count(const EmuBlas* b) { return b->triangleCount; }
const void* emu_blas_nodes(const EmuBlas* b) { return b->nodes.data(); }
const void* emu_blas_triangles(const EmuBlas* b) { return b->tris.data(); }
float emu_blas_sah(const EmuBlas* b) { return b->sahCost; }

EmuScene* emu_scene_create(EmuBlas* blas, const float* instances, uint32_t instanceCount, const VkrtSubmeshInfo* rows, uint32_t rowCount,
                           const VkrtImage* images, uint32_t imageCount)
{
    EmuScene* s = new EmuScene();
    s->blas = blas;
    s->instances.resize(instanceCount);
    for (uint32_t i = 0; i < instanceCount; ++i)
    {
        InstanceRecord& r = s->instances[i];
        memset(&r, 0, sizeof(r));
        memcpy(r.o2w, instances + 12 * (size_t)i, 48);
        affineInverseHost(r.o2w, r.w2o);
        r.nodes = blas->nodes.data(), r.tris = blas->tris.data();
        r.vertices = blas->vertices.data(), r.indices = blas->indices.data();
        r.vertexStride = blas->stride;
        r.mask = 0xFFu;
        r.boxScale = blas->boxScale;
        for (int k = 0; k < 3; ++k)
        {
            r.blasLo[k] = blas->boundsMin[k], r.blasHi[k] = blas->boundsMax[k];
        }
    }
    if (instanceCount > 1)
    {
        // buildTlasOnDevice: k_instance_bounds, the box-only builder with one instance per leaf slot, k_tlas_finalize
        const uint32_t n = instanceCount;
        std::vector<float4> box(2 * (size_t)n);
        float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (uint32_t i = 0; i < n; ++i)
        {
            float l[3], h[3];
            instanceWorldBounds(s->instances[i], l, h);
            box[2 * (size_t)i] = make_float4(l[0], l[1], l[2], 0.f), box[2 * (size_t)i + 1] = make_float4(h[0], h[1], h[2], 0.f);
            for (int k = 0; k < 3; ++k)
            {
                const float c = 0.5f * (l[k] + h[k]);
                clo[k] = fminf(clo[k], c), chi[k] = fmaxf(chi[k], c);
            }
        }
        s->topNodes.assign(n + 8u, Node8{});
        float sah[2] = {0.f, 0.f};
        const uint32_t last = emuBuildHierarchy(box, n, 1u, 0.5f, clo, chi, s->topNodes, n + 8u, s->leafInstance, sah);
        s->topNodes.resize(last);
        s->topBoxScale = emuBoxScale(s->topNodes);
        for (Node8& node : s->topNodes)
        {
            uint32_t m = 0;
            for (int sl = 0; sl < 8; ++sl)
            {
                m |= ((node.leafBits >> (3 * sl)) & 1u) << sl;
            }
            node.reserved = (uint8_t)m;
        }
    }
    if (rows)
    {
        s->rows.assign(rows, rows + rowCount);
    }
    s->texels.resize(imageCount);
    s->textures.resize(imageCount);
    for (uint32_t i = 0; i < imageCount; ++i)
    {
        s->texels[i].assign(images[i].rgba8, images[i].rgba8 + (size_t)images[i].width * images[i].height * 4);
        s->textures[i].texels = s->texels[i].data();
        s->textures[i].width = images[i].width, s->textures[i].height = images[i].height;
    }
    return s;
}

void emu_scene_destroy(EmuScene* s) { delete s; }

// Sequential form of generateMips (textures.cu): blitTexel per destination texel, level after level.
void emu_scene_generate_mips(EmuScene* s)
{
    const size_t n = s->textures.size();
    s->mipTexels.assign(n, {});
    s->mips.assign(n, TextureMips{nullptr, 1u, 0u});
    for (size_t i = 0; i < n; ++i)
    {
        const TextureRecord& t = s->textures[i];
        const uint32_t levels = mipLevelCount(t.width, t.height);
        size_t bytes = 0;
        for (uint32_t k = 1; k < levels; ++k)
        {
            bytes += (size_t)mipExtent(t.width, k) * mipExtent(t.height, k) * 4;
        }
        s->mipTexels[i].resize(std::max<size_t>(bytes, 4));
        const uint8_t* src = t.texels;
        uint8_t* dst = s->mipTexels[i].data();
        for (uint32_t k = 1; k < levels; ++k)
        {
            const uint32_t sw = mipExtent(t.width, k - 1), sh = mipExtent(t.height, k - 1), dw = mipExtent(t.width, k), dh = mipExtent(t.height, k);
            for (uint32_t j = 0; j < dh; ++j)
            {
                for (uint32_t x = 0; x < dw; ++x)
                {
                    const uint32_t px = blitTexel(src, sw, sh, dw, dh, x, j);
                    memcpy(dst + ((size_t)j * dw + x) * 4, &px, 4);
                }
            }
            src = dst;
            dst += (size_t)dw * dh * 4;
        }
        s->mips[i].texels = s->mipTexels[i].data();
        s->mips[i].levels = levels;
    }
}

// Sequential form of rasterizeFrame (raster.cu): set-up of every triangle, the small / large split of k_raster_setup,
// whole bounding boxes for small triangles, 16x16 tiles with the conservative edge rejection for large ones, resolve.
void emu_rasterize(const EmuScene* s, const float* wvp, uint32_t width, uint32_t height, uint8_t* bgra8, uint32_t* triangleId)
{
    const EmuBlas* b = s->blas;
    GeometryTable geo;
    geo.firstTri = b->firstTri.data(), geo.indexBase = b->indexBase.data();
    geo.geometryCount = (uint32_t)b->firstTri.size() - 1u;
    const uint32_t n = b->triangleCount;
    std::vector<RasterSetup> setups(std::max(n, 1u));
    std::vector<uint64_t> visibility((size_t)width * height * kRasterSamples, kRasterEmpty);
    auto minKey = [](uint64_t* key, uint64_t candidate) { *key = std::min(*key, candidate); };
    // triangles are visited in REVERSE draw order on purpose: the keys make the result independent of the order
    for (uint32_t k = 0; k < n; ++k)
    {
        const uint32_t t = n - 1u - k;
        const uint32_t g = findGeometry(geo, t);
        const uint64_t base = geo.indexBase[g] + 3ull * (t - geo.firstTri[g]);
        const uint32_t i0 = b->indices[base], i1 = b->indices[base + 1], i2 = b->indices[base + 2];
        const uint8_t* vb = b->vertices.data();
        const float4 p0 = loadF4(vb, (size_t)i0 * b->stride), p1 = loadF4(vb, (size_t)i1 * b->stride), p2 = loadF4(vb, (size_t)i2 * b->stride);
        const float4 t0 = loadF4(vb, (size_t)i0 * b->stride + 32), t1 = loadF4(vb, (size_t)i1 * b->stride + 32), t2 = loadF4(vb, (size_t)i2 * b->stride + 32);
        const int32_t baseColor = s->rows[g].baseColorTextureIndex;
        const uint32_t texture = (baseColor < 0 || (size_t)baseColor >= s->textures.size()) ? 0u : (uint32_t)baseColor;
        RasterSetup& rs = setups[t];
        if (!rasterSetup(wvp, f3(p0.x, p0.y, p0.z), f3(p1.x, p1.y, p1.z), f3(p2.x, p2.y, p2.z), t0.x, t0.y, t1.x, t1.y, t2.x, t2.y, texture, width,
                         height, rs))
        {
            continue;
        }
        const uint32_t x0 = rs.bboxX & 0xFFFFu, x1 = rs.bboxX >> 16, y0 = rs.bboxY & 0xFFFFu, y1 = rs.bboxY >> 16;
        if (x1 - x0 < 32u && y1 - y0 < 32u) // k_raster_small: the whole bounding box
        {
            for (uint32_t py = y0; py <= y1; ++py)
            {
                for (uint32_t px = x0; px <= x1; ++px)
                {
                    rasterPixel(rs, t, s->textures.data(), s->mips.data(), width, px, py, visibility.data(), minKey);
                }
            }
            continue;
        }
        // k_raster_large: 16x16-pixel tiles of the bounding box, each rejected conservatively against the three edges
        for (uint32_t ty = y0 / kRasterTile; ty <= y1 / kRasterTile; ++ty)
        {
            for (uint32_t tx = x0 / kRasterTile; tx <= x1 / kRasterTile; ++tx)
            {
                const uint32_t px0 = std::max(tx * kRasterTile, x0), px1 = std::min(tx * kRasterTile + kRasterTile - 1u, x1);
                const uint32_t py0 = std::max(ty * kRasterTile, y0), py1 = std::min(ty * kRasterTile + kRasterTile - 1u, y1);
                if (rasterTileOutside(rs, px0, py0, px1, py1))
                {
                    continue;
                }
                for (uint32_t py = py0; py <= py1; ++py)
                {
                    for (uint32_t px = px0; px <= px1; ++px)
                    {
                        rasterPixel(rs, t, s->textures.data(), s->mips.data(), width, px, py, visibility.data(), minKey);
                    }
                }
            }
        }
    }
    for (uint32_t py = 0; py < height; ++py)
    {
        for (uint32_t px = 0; px < width; ++px)
        {
            uint32_t first;
            const uint32_t c = rasterResolve(setups.data(), s->textures.data(), s->mips.data(), width, px, py, visibility.data(), first);
            const size_t pix = (size_t)py * width + px;
            memcpy(bgra8 + pix * 4, &c, 4);
            if (triangleId)
            {
                triangleId[pix] = first;
            }
        }
    }
}

void emu_scene_set_texture_lod(EmuScene* s, int mode) { s->lod = mode == 1 && s->mips.size() == s->textures.size(); }

const uint8_t* emu_texture_level(const EmuScene* s, uint32_t image, uint32_t level, uint32_t* width, uint32_t* height)
{
    const TextureRecord& t = s->textures[image];
    *width = mipExtent(t.width, level), *height = mipExtent(t.height, level);
    if (level == 0)
    {
        return t.texels;
    }
    const uint8_t* p = s->mips[image].texels;
    for (uint32_t k = 1; k < level; ++k)
    {
        p += (size_t)mipExtent(t.width, k) * mipExtent(t.height, k) * 4;
    }
    return p;
}

void emu_sample_texture_lod(const EmuScene* s, uint32_t image, float u, float v, float footprint2, float rgba[4])
{
    sampleTextureLod(s->textures[image], s->mips[image], u, v, footprint2, rgba);
}

void emu_trace_rays(const EmuScene* s, const VkrtRay* rays, uint64_t count, int anyHit, VkrtHit* hits, uint64_t* nodesVisited,
                    uint64_t* trisTested)
{
    TraceCounters tc;
    uint64_t nodes = 0, tris = 0;
    for (uint64_t i = 0; i < count; ++i)
    {
        BestHit best;
        best.t = rays[i].tmax, best.b1 = best.b2 = 0.f;
        best.tri = kInvalid, best.instance = kInvalid, best.buildIndex = kInvalid;
        const F3 o = f3(rays[i].origin[0], rays[i].origin[1], rays[i].origin[2]);
        const F3 d = f3(rays[i].direction[0], rays[i].direction[1], rays[i].direction[2]);
        tc.nodes = tc.tris = 0;
        bool found;
        if (anyHit)
        {
            found = traceScene<true>(s->refs(), o, d, rays[i].tmin, best, &tc);
        }
        else
        {
            found = traceScene<false>(s->refs(), o, d, rays[i].tmin, best, &tc);
        }
        nodes += tc.nodes, tris += tc.tris;
        VkrtHit& h = hits[i];
        if (!found)
        {
            h.t = h.b1 = h.b2 = 0.f;
            h.instance = h.geometry = h.primitive = kInvalid;
        }
        else
        {
            const PackedTri& t = s->instances[best.instance].tris[best.tri];
            h.t = best.t, h.b1 = anyHit ? 0.f : best.b1, h.b2 = anyHit ? 0.f : best.b2;
            h.instance = best.instance;
            h.geometry = anyHit ? 0u : t.geometry, h.primitive = anyHit ? 0u : t.primitive;
        }
        if (tc.overflow)
        {
            fprintf(stderr, "emu_trace_rays: traversal stack overflow\n");
        }
    }
    if (nodesVisited)
    {
        *nodesVisited = nodes;
    }
    if (trisTested)
    {
        *trisTested = tris;
    }
}

// Per-pixel emulation of vkrt_trace (the wavefront stages run back to back for one path).
void emu_render(const EmuScene* s, const VkrtUniforms* u, uint32_t width, uint32_t height, uint32_t spp, uint32_t x0, uint32_t y0,
                uint32_t x1, uint32_t y1, uint32_t strideX, uint32_t strideY, float* rgb32f, uint8_t* bgra8, uint32_t* triangleId,
                uint8_t* shadowMask)
{
    const InstanceRecord* inst = s->instances.data();
    const TlasRefs top = s->refs();
    spp = std::max(spp, 1u);
    for (uint32_t y = y0; y < std::min(y1, height); y += std::max(strideY, 1u))
    {
        for (uint32_t x = x0; x < std::min(x1, width); x += std::max(strideX, 1u))
        {
            F3 sampleSum = f3(0, 0, 0);
            uint32_t aovTri = kInvalid;
            uint8_t aovMask = 0;
            for (uint32_t sample = 0; sample < spp; ++sample)
            {
                float jx, jy;
                sampleOffset(sample, spp, jx, jy);
                F3 o, d;
                primaryRay(*u, width, height, x, y, jx, jy, o, d);
                F3 acc = f3(0, 0, 0);
                float attenuation = 1.0f;
                float travelled = 0.0f; // ray cones: k_shade_surface keeps the primary hit distance in coneT
                for (int depth = 0; depth < 2; ++depth)
                {
                    BestHit best;
                    best.t = kRayTMax, best.b1 = best.b2 = 0.f;
                    best.tri = kInvalid, best.instance = kInvalid, best.buildIndex = kInvalid;
                    traceScene<false>(top, o, d, kRayTMin, best, nullptr);
                    if (best.tri == kInvalid)
                    {
                        acc.x += 0.8f, acc.y += 0.8f, acc.z += 1.0f;
                        break;
                    }
                    const PackedTri& t = inst[best.instance].tris[best.tri];
                    SurfacePoint sp;
                    if (s->lod)
                    {
                        travelled = depth == 0 ? best.t : travelled + best.t;
                        LodInput lod;
                        lod.mips = s->mips.data(), lod.worldRayDir = d;
                        lod.coneWidth = pixelSpreadAngle(*u, height) * travelled;
                        sp = shadeSurface<true>(inst[best.instance], s->rows.data(), s->textures.data(), t.geometry, t.primitive, best.b1,
                                                best.b2, &lod);
                    }
                    else
                    {
                        sp = shadeSurface<false>(inst[best.instance], s->rows.data(), s->textures.data(), t.geometry, t.primitive, best.b1,
                                                 best.b2);
                    }
                    uint32_t bits = 0, cast = 0;
                    for (int i = 0; i < 4; ++i)
                    {
                        const LightTerm lt = lightTerm(*u, i, sp.worldPos, sp.normal);
                        if (lt.castShadow)
                        {
                            cast |= 1u << i;
                            BestHit sh;
                            sh.t = lt.distance, sh.b1 = sh.b2 = 0.f;
                            sh.tri = kInvalid, sh.instance = kInvalid, sh.buildIndex = kInvalid;
                            if (traceScene<true>(top, sp.worldPos, lt.dir, kRayTMin, sh, nullptr))
                            {
                                bits |= 1u << i;
                            }
                        }
                    }
                    if (depth == 0 && sample == 0)
                    {
                        aovTri = (best.instance << 24) + (uint32_t)s->rows[t.geometry].indexBufferOffset + t.primitive;
                        aovMask = (uint8_t)(bits | (cast << 4));
                    }
                    const ResolveOut r = s->lod ? shadeResolve<true>(*u, s->rows.data(), s->textures.data(), sp, bits, attenuation, d, s->mips.data())
                                                : shadeResolve<false>(*u, s->rows.data(), s->textures.data(), sp, bits, attenuation, d);
                    acc.x += r.hitValue.x, acc.y += r.hitValue.y, acc.z += r.hitValue.z;
                    attenuation = r.attenuation;
                    if (!r.reflect)
                    {
                        break;
                    }
                    o = sp.worldPos, d = r.rayDir;
                }
                sampleSum = sample == 0 ? acc : sampleSum + acc;
            }
            const F3 out = spp > 1 ? sampleSum / (float)spp : sampleSum;
            const size_t pix = (size_t)y * width + x;
            if (rgb32f)
            {
                rgb32f[pix * 3] = out.x, rgb32f[pix * 3 + 1] = out.y, rgb32f[pix * 3 + 2] = out.z;
            }
            if (bgra8)
            {
                const uint32_t px = packBgra8(out);
                memcpy(bgra8 + pix * 4, &px, 4);
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
}

} // extern "C"
