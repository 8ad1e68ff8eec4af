"""CPU tier: the kernel LOGIC of libvkrt_b200 (the `__host__ __device__` headers compiled by g++ into the emulation
harness, tests/emu) against the oracle — builder invariants, ray-level parity, frame-level parity.  The same
comparisons run against the real CUDA library in test_gpu_parity.py."""
import numpy as np
import pytest

import helpers as H
from bvh_check import check_bvh8
from vkrt_b200 import camera, scenes
from vkrt_b200.raytracer import flatten_model


def test_builder_invariants_small_counts():
    for n in (0, 1, 2, 3, 4, 5, 9, 33):
        flat = flatten_model(scenes.random_soup(max(n, 1), seed=n + 1, geometries=1))
        if n == 0:
            flat["ranges"][0, 0] = 0
            flat["indices"] = flat["indices"][:0]
        e = H.EmuScene(flat, scenes.identity_instance(1.0))
        nodes, tris = e.read_blas()
        check_bvh8(nodes, tris, n)


def test_builder_invariants_soup_and_duplicates():
    flat = flatten_model(scenes.random_soup(5000, seed=7, geometries=4))
    e = H.EmuScene(flat, scenes.identity_instance(1.0))
    nodes, tris = e.read_blas()
    depth, slots, inner = check_bvh8(nodes, tris, 5000)
    assert depth < 40 and inner < 5000 // 2 + 8
    # all centroids identical: Morton codes tie, the index tie-break must still give a valid tree
    m = scenes.random_soup(300, seed=9, geometries=1)
    m["submeshes"][0]["vertices"][:, :3] = np.tile(m["submeshes"][0]["vertices"][:3, :3], (300, 1))
    e = H.EmuScene(flatten_model(m), scenes.identity_instance(1.0))
    nodes, tris = e.read_blas()
    check_bvh8(nodes, tris, 300)


def test_reinsertion_passes_fast_trace():
    """VKRT_BUILD_PREFER_FAST_TRACE (reinsert.cuh), the same search / lock / verify / apply / refit passes the kernels run:
    the tree stays valid (every triangle once, boxes contain), costs less, answers every ray like the oracle, and
    degenerate inputs (tiny counts, identical boxes: no strictly larger target, so nothing may move) survive."""
    from vkrt_b200 import binding as B

    flat = flatten_model(scenes.random_soup(4000, seed=21, geometries=3))
    inst = scenes.identity_instance(1.0)
    plain, fast = H.EmuScene(flat, inst), H.EmuScene(flat, inst, flags=B.BUILD_PREFER_FAST_TRACE)
    nodes, tris = fast.read_blas()
    depth, slots, inner = check_bvh8(nodes, tris, 4000)
    assert depth < 40
    assert fast.sah() < 0.97 * plain.sah()
    rays = H.random_rays(20000, 5, (-12, -8, -10), (12, 8, 10))
    rays["tmax"][::5] = 4.0
    hp, np_, tp = plain.trace_rays(rays)
    hf, nf, tf = fast.trace_rays(rays)
    assert np.array_equal(hp, hf)
    ho, _ = H.OracleScene(flat, inst).trace_rays(rays)
    assert np.array_equal(ho, hf)
    assert nf < np_  # fewer node visits for the same answers
    ap, _, _ = plain.trace_rays(rays, any_hit=True)
    af, _, _ = fast.trace_rays(rays, any_hit=True)
    assert np.array_equal(ap["instance"] != 0xFFFFFFFF, af["instance"] != 0xFFFFFFFF)
    for n in (1, 2, 3, 4, 5, 9, 33):
        small = flatten_model(scenes.random_soup(n, seed=n + 40, geometries=1))
        nodes, tris = H.EmuScene(small, inst, flags=B.BUILD_PREFER_FAST_TRACE).read_blas()
        check_bvh8(nodes, tris, n)
    m = scenes.random_soup(300, seed=9, geometries=1)
    m["submeshes"][0]["vertices"][:, :3] = np.tile(m["submeshes"][0]["vertices"][:3, :3], (300, 1))
    nodes, tris = H.EmuScene(flatten_model(m), inst, flags=B.BUILD_PREFER_FAST_TRACE).read_blas()
    check_bvh8(nodes, tris, 300)


def test_path_to_pixel_partition():
    """pathToPixel (shading.cuh): over the ranks every pixel of a ragged frame is produced exactly once, a warp of 32
    consecutive paths is one 8x4 block and 4 consecutive paths a 2x2 quad (the layout the pixel-major waves rely on)."""
    import ctypes as C

    lib = H.emu_lib()
    lib.emu_path_to_pixel.restype = C.c_uint32
    lib.emu_path_to_pixel.argtypes = [C.c_uint32] * 5 + [C.c_void_p] * 3
    W, Hh, size = 333, 190, 3
    seen = np.zeros((Hh, W), dtype=np.int32)
    for rank in range(size):
        tiles = ((W + 63) // 64) * ((Hh + 63) // 64)
        owned = len(range(rank, tiles, size))
        n = owned * 4096
        x, y, pit = (np.zeros(n, dtype=np.uint32) for _ in range(3))
        valid = lib.emu_path_to_pixel(W, Hh, rank, size, n, x.ctypes.data, y.ctypes.data, pit.ctypes.data)
        ok = x != 0xFFFFFFFF
        assert valid == ok.sum()
        np.add.at(seen, (y[ok], x[ok]), 1)
        # geometry of a warp / quad, checked on the paths of the first (full) tile of rank 0
        if rank == 0:
            bx, by = x[:4096].reshape(128, 32), y[:4096].reshape(128, 32)
            assert ((bx.max(1) - bx.min(1)) == 7).all() and ((by.max(1) - by.min(1)) == 3).all()
            qx, qy = x[:4096].reshape(1024, 4), y[:4096].reshape(1024, 4)
            assert ((qx.max(1) - qx.min(1)) == 1).all() and ((qy.max(1) - qy.min(1)) == 1).all()
            assert np.array_equal(pit[:4096], (y[:4096] % 64) * 64 + x[:4096] % 64)
    assert (seen == 1).all()


def test_cost_optimal_collapse(monkeypatch):
    """The cost-optimal collapse (dpSolve / dpExpand in builder.cuh) against the greedy one on the same binary tree:
    both give valid trees and identical closest hits; the optimal one has the lower SAH cost and tests fewer triangles."""
    flat = flatten_model(scenes.random_soup(20000, seed=11, geometries=3))
    inst = scenes.identity_instance(1.0)
    rays = H.random_rays(20000, 5, (-12, -8, -10), (12, 8, 10))
    out = {}
    for name, cost in (("greedy", "0"), ("optimal", "0.5")):
        monkeypatch.setenv("VKRT_EMU_DP_PRIM_COST", cost)
        e = H.EmuScene(flat, inst)
        nodes, tris = e.read_blas()
        check_bvh8(nodes, tris, 20000)
        hits, n_nodes, n_tris = e.trace_rays(rays)
        out[name] = (e.sah(), len(nodes), hits, n_nodes, n_tris)
    assert np.array_equal(out["greedy"][2], out["optimal"][2])
    assert out["optimal"][0] < out["greedy"][0]
    assert out["optimal"][3] * 0.5 + out["optimal"][4] < out["greedy"][3] * 0.5 + out["greedy"][4]


def test_builder_invariants_sponza(sponza_flat):
    e = H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    nodes, tris = e.read_blas()
    depth, slots, inner = check_bvh8(nodes, tris, 262267)
    assert slots == 262267 and depth < 48
    # geometry / primitive tags follow the reference's (geometryIndex, primitiveID) numbering
    first = np.concatenate([[0], np.cumsum(sponza_flat["ranges"][:, 0].astype(np.int64))])
    assert np.array_equal(first[tris["geometry"]] + tris["primitive"], tris["buildIndex"])


@pytest.mark.parametrize("seed", [3, 4])
def test_ray_parity_soup(seed):
    flat = flatten_model(scenes.random_soup(6000, seed=seed))
    inst = scenes.identity_instance(1.0)
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    rays = H.random_rays(30000, seed, (-12, -8, -10), (12, 8, 10))
    rays["tmax"][::7] = 3.0  # short rays exercise the exclusive tmax
    ho, _ = o.trace_rays(rays)
    he, nodes, tris = e.trace_rays(rays)
    assert np.array_equal(ho, he)
    ao, _ = o.trace_rays(rays, any_hit=True)
    ae, _, _ = e.trace_rays(rays, any_hit=True)
    assert np.array_equal(ao["instance"] != 0xFFFFFFFF, ae["instance"] != 0xFFFFFFFF)


def test_ray_parity_degenerate_rays():
    flat = flatten_model(scenes.analytic_scene())
    inst = scenes.identity_instance(1.0)
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    rays = np.zeros(8, dtype=H.B.RAY_DTYPE)
    rays["origin"] = (0, 3, 0)
    rays["tmin"], rays["tmax"] = 0.001, 1000
    rays["direction"] = [(0, -1, 0), (0, 0, 0), (np.nan, -1, 0), (np.inf, -1, 0), (0, -1, 0), (1, 0, 0), (0, -1e-30, 0), (0, 1, 0)]
    rays["tmax"][4] = 0.0005  # tmax < tmin
    rays["origin"][5] = (-20, 0, 0)  # ray lying exactly in the floor plane
    ho, _ = o.trace_rays(rays, mode=H.MODE_BRUTE)
    he, _, _ = e.trace_rays(rays)
    assert np.array_equal(ho, he)
    assert ho["instance"][0] == 0 and (ho["instance"][1:5] == 0xFFFFFFFF).all()


def test_ray_parity_shared_edges_and_duplicates():
    """Rays aimed exactly at shared edges/vertices of a regular grid and at exactly duplicated triangles: watertight
    test + equal-t tie-break must give the same winner as the oracle."""
    n = 12
    xs = np.linspace(-3, 3, n + 1, dtype=np.float32)
    gx, gz = np.meshgrid(xs, xs)
    pos = np.stack([gx.ravel(), np.zeros_like(gx).ravel(), gz.ravel()], axis=1)
    quads = [(r * (n + 1) + c, r * (n + 1) + c + 1, (r + 1) * (n + 1) + c + 1, (r + 1) * (n + 1) + c) for r in range(n) for c in range(n)]
    idx = np.array([[a, b, c, a, c, d] for a, b, c, d in quads], dtype=np.uint32).ravel()
    idx = np.concatenate([idx, np.roll(idx.reshape(-1, 3), 1, axis=1).ravel()])  # every triangle twice
    v = scenes.make_vertex_array(pos, np.tile([0, 1, 0], (len(pos), 1)), pos[:, [0, 2]], np.tile([1, 0, 0, 1], (len(pos), 1)))
    model = {"submeshes": [{"vertices": v, "indices": idx, "material": 0}], "materials": [(1, 2, 0)],
             "images": scenes.analytic_scene()["images"]}
    flat = flatten_model(model)
    inst = scenes.identity_instance(1.0)
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    rays = np.zeros(len(pos) * 2, dtype=H.B.RAY_DTYPE)
    rays["origin"][: len(pos)] = pos + np.array([0, 5, 0], dtype=np.float32)
    rays["direction"][: len(pos)] = (0, -1, 0)
    rays["origin"][len(pos):] = (0.1, 4, 0.2)
    rays["direction"][len(pos):] = pos - np.array([0.1, 4, 0.2], dtype=np.float32)
    rays["tmin"], rays["tmax"] = 0.001, 1000
    ho, _ = o.trace_rays(rays, mode=H.MODE_BRUTE)
    hb, _ = o.trace_rays(rays, mode=H.MODE_BVH)
    he, _, _ = e.trace_rays(rays)
    assert np.array_equal(ho, hb) and np.array_equal(ho, he)
    interior = ho["instance"][: len(pos)] != 0xFFFFFFFF
    assert interior.mean() > 0.8  # watertight: grid vertices and edges are hit, not leaked through
    assert (ho["primitive"][ho["instance"] != 0xFFFFFFFF] < 2 * n * n).all()  # ties go to the first copy


def test_ray_parity_top_level_many_instances():
    """Two-level traversal over a built top level: 300 rotated, scaled, overlapping instances of one small
    BLAS; closest and any hit equal to the oracle, which simply loops over every instance."""
    flat = flatten_model(scenes.random_soup(400, seed=9))
    inst = scenes.random_instance_transforms(300, seed=2, extent=(80.0, 30.0, 80.0), scale=(0.3, 1.5))
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    rays = H.random_rays(12000, 5, (-45, -20, -45), (45, 20, 45))
    rays["tmax"][::5] = 12.0
    ho, _ = o.trace_rays(rays)
    he, nodes, tris = e.trace_rays(rays)
    assert np.array_equal(ho, he)
    assert len(np.unique(ho["instance"][ho["instance"] != 0xFFFFFFFF])) > 100
    ao, _ = o.trace_rays(rays, any_hit=True)
    ae, _, _ = e.trace_rays(rays, any_hit=True)
    assert np.array_equal(ao["instance"] != 0xFFFFFFFF, ae["instance"] != 0xFFFFFFFF)


def test_frame_parity_analytic_and_instances():
    flat = flatten_model(scenes.analytic_scene(metallic_floor=True))
    inst = np.stack([scenes.identity_instance(1.0), scenes.identity_instance(0.5, (10, 0, -4)), scenes.identity_instance(1.0, (-18, 1, 2))])
    o, e = H.OracleScene(flat, inst), H.EmuScene(flat, inst)
    W, Hh = 128, 72
    u = camera.make_uniforms(W, Hh, position=(0, 5, 16), rotation=(-0.25, 0.1, 0))
    a, b = o.render(u, W, Hh), e.render(u, W, Hh)
    c = H.compare_frames(a, b)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"] and c["bgra_max_diff"] == 0
    assert len(np.unique(a["tri"] >> 24)) >= 3  # several instances (and the sky) are visible


def test_frame_parity_sponza_config1_subset(sponza_flat):
    o, e = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE), H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    W, Hh = 1280, 720
    u = camera.make_uniforms(W, Hh)
    a, b = o.render(u, W, Hh, stride=(5, 5)), e.render(u, W, Hh, stride=(5, 5))
    sel = np.zeros((Hh, W), dtype=bool)
    sel[::5, ::5] = True
    c = H.compare_frames(a, b, sel)
    assert c["tri_agree"] >= 0.999 and c["mask_agree"] >= 0.999 and c["bgra_max_diff"] <= 1
    assert c["tri_agree"] == 1.0 and c["rgb_bit_equal"]  # the design goal is bit equality, the bar above is the contract


def test_frame_parity_spp8_jitter(sponza_flat):
    o, e = H.OracleScene(sponza_flat, scenes.REFERENCE_INSTANCE), H.EmuScene(sponza_flat, scenes.REFERENCE_INSTANCE)
    W, Hh = 384, 216
    u = camera.make_uniforms(W, Hh)
    a, b = o.render(u, W, Hh, spp=8, stride=(3, 3)), e.render(u, W, Hh, spp=8, stride=(3, 3))
    sel = np.zeros((Hh, W), dtype=bool)
    sel[::3, ::3] = True
    c = H.compare_frames(a, b, sel)
    assert c["tri_agree"] == 1.0 and c["mask_agree"] == 1.0 and c["rgb_bit_equal"]
