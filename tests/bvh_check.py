"""Structural invariants of a compressed 8-wide BVH read back from the builder (GPU or emulated)."""
import numpy as np


def decode_child_boxes(node):
    p = node["p"].astype(np.float32)
    # sx, sy, sz hold the upper half of the float `scale`
    half = np.array([node["sx"], node["sy"], node["sz"]], dtype=np.uint32)
    scale = (half << 16).view(np.float32)
    lo = np.stack([p[0] + node["qlox"].astype(np.float32) * scale[0], p[1] + node["qloy"].astype(np.float32) * scale[1],
                   p[2] + node["qloz"].astype(np.float32) * scale[2]], axis=1)
    hi = np.stack([p[0] + node["qhix"].astype(np.float32) * scale[0], p[1] + node["qhiy"].astype(np.float32) * scale[1],
                   p[2] + node["qhiz"].astype(np.float32) * scale[2]], axis=1)
    return lo, hi


def check_bvh8(nodes, tris, expected_triangles):
    """Walks the tree from the root.  Verifies: every packed triangle slot is referenced exactly once, every build
    index appears exactly once, internal children are contiguous (childBase + rank in imask), quantised child boxes
    contain the child's exact contents.  Returns (max_depth, leaf_slots, inner_nodes)."""
    n_tris = len(tris)
    assert n_tris == expected_triangles
    seen_slot = np.zeros(max(n_tris, 1), dtype=np.int32)
    seen_node = np.zeros(len(nodes), dtype=np.int32)
    tri_lo = np.minimum(np.minimum(tris["v0"], tris["v1"]), tris["v2"]) if n_tris else np.zeros((0, 3), np.float32)
    tri_hi = np.maximum(np.maximum(tris["v0"], tris["v1"]), tris["v2"]) if n_tris else np.zeros((0, 3), np.float32)
    node_lo = np.full((len(nodes), 3), np.inf, dtype=np.float32)
    node_hi = np.full((len(nodes), 3), -np.inf, dtype=np.float32)
    max_depth = 0
    order = []
    stack = [(0, 0)]
    while stack:
        ni, depth = stack.pop()
        seen_node[ni] += 1
        order.append(ni)
        max_depth = max(max_depth, depth)
        node = nodes[ni]
        imask = int(node["imask"])
        leaf_bits = int(node["leafBits"])
        assert leaf_bits < (1 << 24)
        half = [int(node["sx"]), int(node["sy"]), int(node["sz"])]
        exps = [h >> 7 for h in half]
        assert max(exps) - min(exps) <= 4, (ni, exps)  # builder.cuh::quantExponents
        for s in range(8):
            unary = (leaf_bits >> (3 * s)) & 7
            if (imask >> s) & 1:
                assert unary == 0, "a slot is internal or leaf, not both"
                child = int(node["childBase"]) + bin(imask & ((1 << s) - 1)).count("1")
                assert 0 < child < len(nodes)
                stack.append((child, depth + 1))
            elif unary:
                count = {1: 1, 3: 2, 7: 3}[unary]
                off = bin(leaf_bits & ((1 << (3 * s)) - 1)).count("1")  # leaves are stored in slot order
                for k in range(count):
                    seen_slot[int(node["primBase"]) + off + k] += 1
    assert (seen_node == 1).all(), "every node reachable exactly once"
    if n_tris:
        assert (seen_slot == 1).all(), "every packed triangle referenced exactly once"
        assert np.array_equal(np.sort(tris["buildIndex"]), np.arange(n_tris, dtype=np.uint32))
    # bottom-up exact bounds, then containment in the quantised boxes
    for ni in reversed(order):
        node = nodes[ni]
        lo, hi = decode_child_boxes(node)
        imask = int(node["imask"])
        leaf_bits = int(node["leafBits"]) & 0xFFFFFF
        for s in range(8):
            unary = (leaf_bits >> (3 * s)) & 7
            if (imask >> s) & 1:
                child = int(node["childBase"]) + bin(imask & ((1 << s) - 1)).count("1")
                clo, chi = node_lo[child], node_hi[child]
            elif unary:
                count = {1: 1, 3: 2, 7: 3}[unary]
                first = int(node["primBase"]) + bin(leaf_bits & ((1 << (3 * s)) - 1)).count("1")
                clo, chi = tri_lo[first:first + count].min(axis=0), tri_hi[first:first + count].max(axis=0)
            else:
                continue
            assert (lo[s] <= clo).all() and (hi[s] >= chi).all(), (ni, s, lo[s], clo, hi[s], chi)
            node_lo[ni] = np.minimum(node_lo[ni], clo)
            node_hi[ni] = np.maximum(node_hi[ni], chi)
        assert (node["p"] <= node_lo[ni]).all() or not np.isfinite(node_lo[ni]).all()
    return max_depth, int(seen_slot.sum()), len(nodes)


def instance_world_boxes(blas_lo, blas_hi, transforms):
    """Exact (float64) world boxes of instances: the 8 corners of the BLAS box through each row-major 3x4 transform."""
    t = np.asarray(transforms, dtype=np.float64).reshape(-1, 3, 4)
    corners = np.array([[(blas_hi if (c >> k) & 1 else blas_lo)[k] for k in range(3)] for c in range(8)], dtype=np.float64)
    w = np.einsum("nij,cj->nci", t[:, :, :3], corners) + t[:, None, :, 3]
    return w.min(axis=1), w.max(axis=1)


def check_tlas(nodes, leaf_instance, inst_lo, inst_hi):
    """Top-level hierarchy (vkrt_tlas_read): every instance sits in exactly one leaf slot (Node8.reserved = mask of leaf
    slots, one instance each), internal children are contiguous, quantised child boxes contain the instances' world
    boxes below them.  Returns the depth."""
    n = len(leaf_instance)
    assert np.array_equal(np.sort(leaf_instance), np.arange(n, dtype=np.uint32)), "every instance exactly once"
    seen_node = np.zeros(len(nodes), dtype=np.int32)
    seen_leaf = np.zeros(n, dtype=np.int32)
    node_lo = np.full((len(nodes), 3), np.inf)
    node_hi = np.full((len(nodes), 3), -np.inf)
    order, stack, depth_max = [], [(0, 1)], 0
    while stack:
        ni, depth = stack.pop()
        seen_node[ni] += 1
        order.append(ni)
        depth_max = max(depth_max, depth)
        node = nodes[ni]
        imask, leaves, bits = int(node["imask"]), int(node["reserved"]), int(node["leafBits"]) & 0xFFFFFF
        assert imask & leaves == 0
        assert bits == sum(1 << (3 * s) for s in range(8) if (leaves >> s) & 1), "one instance per leaf slot"
        for s in range(8):
            if (imask >> s) & 1:
                child = int(node["childBase"]) + bin(imask & ((1 << s) - 1)).count("1")
                assert 0 < child < len(nodes)
                stack.append((child, depth + 1))
            elif (leaves >> s) & 1:
                seen_leaf[int(node["primBase"]) + bin(leaves & ((1 << s) - 1)).count("1")] += 1
    assert (seen_node == 1).all() and (seen_leaf == 1).all()
    for ni in reversed(order):
        node = nodes[ni]
        lo, hi = decode_child_boxes(node)
        imask, leaves = int(node["imask"]), int(node["reserved"])
        for s in range(8):
            if (imask >> s) & 1:
                child = int(node["childBase"]) + bin(imask & ((1 << s) - 1)).count("1")
                clo, chi = node_lo[child], node_hi[child]
            elif (leaves >> s) & 1:
                i = int(leaf_instance[int(node["primBase"]) + bin(leaves & ((1 << s) - 1)).count("1")])
                clo, chi = inst_lo[i], inst_hi[i]
            else:
                continue
            assert (lo[s] <= clo).all() and (hi[s] >= chi).all(), (ni, s)
            node_lo[ni] = np.minimum(node_lo[ni], clo)
            node_hi[ni] = np.maximum(node_hi[ni], chi)
    return depth_max
