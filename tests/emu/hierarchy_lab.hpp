// hierarchy_lab.hpp — host-only experiments on the BINARY hierarchy that feeds the collapse (test infrastructure, like
// the rest of tests/emu): sequential PLOC, a top-down binned-SAH build and Bittner-style reinsertion.  They answer
// "how many node visits per ray would a better binary hierarchy save" on the CPU, before any GPU time is spent on a
// device version.  Selected with VKRT_EMU_BINARY = lbvh (default) | ploc | sah and VKRT_EMU_REINSERT = <passes>.
#pragma once

#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <queue>
#include <vector>

namespace lab
{
struct Box
{
    float lo[3], hi[3];
};
inline Box emptyBox() { return Box{{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}}; }
inline Box merge(const Box& a, const Box& b)
{
    Box r;
    for (int k = 0; k < 3; ++k)
    {
        r.lo[k] = std::min(a.lo[k], b.lo[k]), r.hi[k] = std::max(a.hi[k], b.hi[k]);
    }
    return r;
}
inline float area(const Box& b)
{
    const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
    return x * y + y * z + z * x;
}

// Unified ids: [0, n-1) internal nodes, [n-1, 2n-1) leaves (leaf i = id n-1+i).
struct Tree
{
    uint32_t n = 0;
    int root = 0;
    std::vector<int> left, right, parent;
    std::vector<Box> box; // 2n-1
    bool isLeaf(int id) const { return id >= (int)n - 1; }
    void init(uint32_t count, const std::vector<Box>& leaves)
    {
        n = count;
        left.assign(n - 1, -1), right.assign(n - 1, -1), parent.assign(2 * n - 1, -1);
        box.assign(2 * n - 1, emptyBox());
        for (uint32_t i = 0; i < n; ++i)
        {
            box[n - 1 + i] = leaves[i];
        }
    }
    void refitUp(int id)
    {
        while (id >= 0)
        {
            box[id] = merge(box[left[id]], box[right[id]]);
            id = parent[id];
        }
    }
    double sah() const // sum of internal areas / root area (binary-tree SAH, node cost 1)
    {
        double s = 0;
        for (uint32_t i = 0; i + 1 < n; ++i)
        {
            s += area(box[i]);
        }
        return s / area(box[root]);
    }
};

// Sequential PLOC over leaves given in Morton order (the GPU's k_ploc_round, radius 8, mutual nearest neighbours).
inline void buildPloc(Tree& t, int radius)
{
    std::vector<int> cur(t.n), next;
    for (uint32_t i = 0; i < t.n; ++i)
    {
        cur[i] = (int)(t.n - 1 + i);
    }
    int nextInternal = 0;
    std::vector<int> nn;
    while (cur.size() > 1)
    {
        const int m = (int)cur.size();
        static const int exactBelow = getenv("VKRT_LAB_PLOC_EXACT") ? atoi(getenv("VKRT_LAB_PLOC_EXACT")) : 0;
        if (m <= exactBelow)
        {
            radius = m; // exhaustive nearest-neighbour search for the top of the tree
        }
        nn.assign(m, -1);
        for (int i = 0; i < m; ++i)
        {
            float best = FLT_MAX;
            for (int j = std::max(0, i - radius); j <= std::min(m - 1, i + radius); ++j)
            {
                if (j == i)
                {
                    continue;
                }
                const float a = area(merge(t.box[cur[i]], t.box[cur[j]]));
                if (a < best || (a == best && (uint32_t)(i ^ j) < (uint32_t)(i ^ nn[i])))
                {
                    best = a, nn[i] = j;
                }
            }
        }
        next.clear();
        for (int i = 0; i < m; ++i)
        {
            const int j = nn[i];
            if (nn[j] == i)
            {
                if (i < j)
                {
                    const int id = nextInternal++;
                    t.left[id] = cur[i], t.right[id] = cur[j];
                    t.parent[cur[i]] = id, t.parent[cur[j]] = id;
                    t.box[id] = merge(t.box[cur[i]], t.box[cur[j]]);
                    next.push_back(id);
                }
            }
            else
            {
                next.push_back(cur[i]);
            }
        }
        cur.swap(next);
    }
    t.root = cur[0];
    t.parent[t.root] = -1;
}

// Top-down binned SAH (32 bins on the centroids, all three axes), down to single leaves.
inline void buildSah(Tree& t)
{
    constexpr int kBins = 32;
    std::vector<uint32_t> order(t.n);
    for (uint32_t i = 0; i < t.n; ++i)
    {
        order[i] = i;
    }
    struct Task
    {
        uint32_t first, count;
        int parent;
        bool asLeft;
    };
    std::vector<Task> stack;
    stack.push_back({0u, t.n, -1, false});
    int nextInternal = 0;
    auto leafBox = [&](uint32_t leaf) -> const Box& { return t.box[t.n - 1 + leaf]; };
    while (!stack.empty())
    {
        const Task task = stack.back();
        stack.pop_back();
        int id;
        if (task.count == 1)
        {
            id = (int)(t.n - 1 + order[task.first]);
        }
        else
        {
            id = nextInternal++;
            Box cb = emptyBox(), bb = emptyBox();
            for (uint32_t i = 0; i < task.count; ++i)
            {
                const Box& b = leafBox(order[task.first + i]);
                bb = merge(bb, b);
                for (int k = 0; k < 3; ++k)
                {
                    const float c = 0.5f * (b.lo[k] + b.hi[k]);
                    cb.lo[k] = std::min(cb.lo[k], c), cb.hi[k] = std::max(cb.hi[k], c);
                }
            }
            t.box[id] = bb;
            int bestAxis = -1, bestSplit = 0;
            float bestCost = FLT_MAX;
            for (int axis = 0; axis < 3; ++axis)
            {
                const float ext = cb.hi[axis] - cb.lo[axis];
                if (!(ext > 0.f))
                {
                    continue;
                }
                Box bins[kBins];
                uint32_t cnt[kBins] = {};
                for (auto& b : bins)
                {
                    b = emptyBox();
                }
                const float scale = kBins / ext;
                for (uint32_t i = 0; i < task.count; ++i)
                {
                    const Box& b = leafBox(order[task.first + i]);
                    const int bin = std::min(kBins - 1, (int)((0.5f * (b.lo[axis] + b.hi[axis]) - cb.lo[axis]) * scale));
                    bins[bin] = merge(bins[bin], b), cnt[bin]++;
                }
                float rightArea[kBins];
                uint32_t rightCnt[kBins];
                Box acc = emptyBox();
                uint32_t c = 0;
                for (int b = kBins - 1; b > 0; --b)
                {
                    acc = merge(acc, bins[b]), c += cnt[b];
                    rightArea[b] = c ? area(acc) : 0.f, rightCnt[b] = c;
                }
                acc = emptyBox(), c = 0;
                for (int b = 0; b + 1 < kBins; ++b)
                {
                    acc = merge(acc, bins[b]), c += cnt[b];
                    if (c == 0 || rightCnt[b + 1] == 0)
                    {
                        continue;
                    }
                    const float cost = area(acc) * c + rightArea[b + 1] * rightCnt[b + 1];
                    if (cost < bestCost)
                    {
                        bestCost = cost, bestAxis = axis, bestSplit = b;
                    }
                }
            }
            uint32_t mid;
            if (bestAxis < 0)
            {
                mid = task.first + task.count / 2; // identical centroids
            }
            else
            {
                const float ext = cb.hi[bestAxis] - cb.lo[bestAxis], scale = kBins / ext;
                auto it = std::partition(order.begin() + task.first, order.begin() + task.first + task.count, [&](uint32_t leaf) {
                    const Box& b = leafBox(leaf);
                    return std::min(kBins - 1, (int)((0.5f * (b.lo[bestAxis] + b.hi[bestAxis]) - cb.lo[bestAxis]) * scale)) <= bestSplit;
                });
                mid = (uint32_t)(it - order.begin());
            }
            stack.push_back({mid, task.first + task.count - mid, id, false});
            stack.push_back({task.first, mid - task.first, id, true});
        }
        if (task.parent < 0)
        {
            t.root = id, t.parent[id] = -1;
        }
        else
        {
            (task.asLeft ? t.left : t.right)[task.parent] = id;
            t.parent[id] = task.parent;
        }
    }
}

// Rebuilds the TOP of a tree: the K largest subtrees (cut found by opening the largest node again and again) are kept,
// the inner nodes above them are thrown away and rebuilt top-down with the binned SAH over the K boxes, weighted by
// the number of leaves below each.
inline void rebuildTop(Tree& t, uint32_t K)
{
    std::vector<uint32_t> leaves(2 * (size_t)t.n - 1, 1u);
    {
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
                work.emplace_back(t.left[node], 0), work.emplace_back(t.right[node], 0);
            }
            else
            {
                leaves[node] = leaves[t.left[node]] + leaves[t.right[node]];
                work.pop_back();
            }
        }
    }
    auto cmp = [&](int a, int b) { return area(t.box[a]) < area(t.box[b]); };
    std::priority_queue<int, std::vector<int>, decltype(cmp)> open(cmp);
    std::vector<int> cut, freed;
    open.push(t.root);
    while (!open.empty() && open.size() + cut.size() < K)
    {
        const int node = open.top();
        open.pop();
        if (t.isLeaf(node))
        {
            cut.push_back(node);
            continue;
        }
        freed.push_back(node);
        open.push(t.left[node]), open.push(t.right[node]);
    }
    while (!open.empty())
    {
        cut.push_back(open.top()), open.pop();
    }
    if (cut.size() < 2)
    {
        return;
    }
    constexpr int kBins = 32;
    struct Task
    {
        uint32_t first, count;
        int parent;
        bool asLeft;
    };
    std::vector<Task> stack;
    stack.push_back({0u, (uint32_t)cut.size(), -1, false});
    size_t nextFreed = 0;
    while (!stack.empty())
    {
        const Task task = stack.back();
        stack.pop_back();
        int id;
        if (task.count == 1)
        {
            id = cut[task.first];
        }
        else
        {
            id = freed[nextFreed++];
            Box cb = emptyBox(), bb = emptyBox();
            for (uint32_t i = 0; i < task.count; ++i)
            {
                const Box& b = t.box[cut[task.first + i]];
                bb = merge(bb, b);
                for (int k = 0; k < 3; ++k)
                {
                    const float c = 0.5f * (b.lo[k] + b.hi[k]);
                    cb.lo[k] = std::min(cb.lo[k], c), cb.hi[k] = std::max(cb.hi[k], c);
                }
            }
            t.box[id] = bb;
            int bestAxis = -1, bestSplit = 0;
            float bestCost = FLT_MAX;
            for (int axis = 0; axis < 3; ++axis)
            {
                const float ext = cb.hi[axis] - cb.lo[axis];
                if (!(ext > 0.f))
                {
                    continue;
                }
                Box bins[kBins];
                uint32_t cnt[kBins] = {};
                for (auto& b : bins)
                {
                    b = emptyBox();
                }
                const float scale = kBins / ext;
                for (uint32_t i = 0; i < task.count; ++i)
                {
                    const Box& b = t.box[cut[task.first + i]];
                    const int bin = std::min(kBins - 1, (int)((0.5f * (b.lo[axis] + b.hi[axis]) - cb.lo[axis]) * scale));
                    bins[bin] = merge(bins[bin], b), cnt[bin] += leaves[cut[task.first + i]];
                }
                float rightArea[kBins];
                uint32_t rightCnt[kBins];
                Box acc = emptyBox();
                uint32_t c = 0;
                for (int b = kBins - 1; b > 0; --b)
                {
                    acc = merge(acc, bins[b]), c += cnt[b];
                    rightArea[b] = c ? area(acc) : 0.f, rightCnt[b] = c;
                }
                acc = emptyBox(), c = 0;
                for (int b = 0; b + 1 < kBins; ++b)
                {
                    acc = merge(acc, bins[b]), c += cnt[b];
                    if (c == 0 || rightCnt[b + 1] == 0)
                    {
                        continue;
                    }
                    const float cost = area(acc) * c + rightArea[b + 1] * rightCnt[b + 1];
                    if (cost < bestCost)
                    {
                        bestCost = cost, bestAxis = axis, bestSplit = b;
                    }
                }
            }
            uint32_t mid;
            if (bestAxis < 0)
            {
                mid = task.first + task.count / 2;
            }
            else
            {
                const float ext = cb.hi[bestAxis] - cb.lo[bestAxis], scale = kBins / ext;
                auto it = std::partition(cut.begin() + task.first, cut.begin() + task.first + task.count, [&](int node) {
                    const Box& b = t.box[node];
                    return std::min(kBins - 1, (int)((0.5f * (b.lo[bestAxis] + b.hi[bestAxis]) - cb.lo[bestAxis]) * scale)) <= bestSplit;
                });
                mid = (uint32_t)(it - cut.begin());
            }
            stack.push_back({mid, task.first + task.count - mid, id, false});
            stack.push_back({task.first, mid - task.first, id, true});
        }
        if (task.parent < 0)
        {
            t.root = id, t.parent[id] = -1;
        }
        else
        {
            (task.asLeft ? t.left : t.right)[task.parent] = id;
            t.parent[id] = task.parent;
        }
    }
}

// Insertion-based optimisation (Bittner, Hapala, Havran 2013), sequential: every node in order of decreasing area is
// taken out with its parent and put back where the total area grows least (branch and bound from the root).
inline void reinsert(Tree& t, int passes)
{
    std::vector<int> cand;
    struct Item
    {
        float induced;
        int node;
        bool operator<(const Item& o) const { return induced > o.induced; }
    };
    for (int pass = 0; pass < passes; ++pass)
    {
        cand.clear();
        for (int id = 0; id < (int)(2 * t.n - 1); ++id)
        {
            if (id != t.root && t.parent[id] != t.root && t.parent[id] >= 0)
            {
                cand.push_back(id);
            }
        }
        const char* order = getenv("VKRT_LAB_ORDER");
        if (order && order[0] == 'a')
        {
            std::sort(cand.begin(), cand.end(), [&](int a, int b) { return area(t.box[a]) < area(t.box[b]); });
        }
        else if (order && order[0] == 'r')
        {
            uint64_t state = 88172645463325252ull;
            for (size_t i = cand.size(); i > 1; --i)
            {
                state ^= state << 13, state ^= state >> 7, state ^= state << 17;
                std::swap(cand[i - 1], cand[state % i]);
            }
        }
        else
        {
            std::sort(cand.begin(), cand.end(), [&](int a, int b) { return area(t.box[a]) > area(t.box[b]); });
        }
        uint64_t moved = 0;
        for (int L : cand)
        {
            const int P = t.parent[L];
            if (P < 0 || P == t.root)
            {
                continue;
            }
            const int G = t.parent[P];
            const int S = t.left[P] == L ? t.right[P] : t.left[P];
            // take L and P out: S replaces P under G
            (t.left[G] == P ? t.left[G] : t.right[G]) = S;
            t.parent[S] = G;
            t.refitUp(G);
            const float areaL = area(t.box[L]);
            int best = S;
            float bestCost = FLT_MAX;
            std::priority_queue<Item> queue;
            queue.push({0.f, t.root});
            while (!queue.empty())
            {
                const Item it = queue.top();
                queue.pop();
                if (it.induced + areaL >= bestCost)
                {
                    break;
                }
                const float direct = area(merge(t.box[it.node], t.box[L]));
                const float total = it.induced + direct;
                static const bool rule = getenv("VKRT_LAB_AREA_RULE") != nullptr;
                if (total < bestCost && (!rule || area(t.box[it.node]) > areaL))
                {
                    bestCost = total, best = it.node;
                }
                const float childInduced = total - area(t.box[it.node]);
                if (!t.isLeaf(it.node) && childInduced + areaL < bestCost)
                {
                    queue.push({childInduced, t.left[it.node]});
                    queue.push({childInduced, t.right[it.node]});
                }
            }
            // put P back as the parent of (best, L)
            const int X = best, GX = t.parent[X];
            if (GX < 0)
            {
                t.root = P;
            }
            else
            {
                (t.left[GX] == X ? t.left[GX] : t.right[GX]) = P;
            }
            t.parent[P] = GX;
            t.left[P] = X, t.right[P] = L;
            t.parent[X] = P, t.parent[L] = P;
            t.refitUp(P);
            moved += X != S;
        }
        fprintf(stderr, "lab: reinsertion pass %d moved %llu of %zu nodes, binary SAH %.3f\n", pass, (unsigned long long)moved, cand.size(), t.sah());
    }
}
} // namespace lab
