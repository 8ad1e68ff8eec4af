"""Structural invariants of a compressed 8-wide BVH read back from the builder (GPU or emulated)."""
import numpy as np


def decode_child_boxes(node):
    p = node["p"].astype(np.float32)
    # sx, sy, sz hold the upper half of the float 2^15 * scale
    half = np.array([node["sx"], node["sy"], node["sz"]], dtype=np.uint32)
    scale = ((half << 16).view(np.float32) * np.float32(2.0 ** -15)).astype(np.float32)
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
        leaf_bits = int(node["leafBits"])
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
