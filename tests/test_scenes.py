"""CPU tier: the Sponza stand-in honours the real glTF's structure; flatten_model restates the reference's
createVertexAndIndexBuffer / SubmeshInfo table (src/Raytracer.cpp:642-742, :1412-1427)."""
import numpy as np

from vkrt_b200 import scenes
from vkrt_b200.raytracer import flatten_model


def test_standin_counts_and_bounds():
    layout = scenes.load_layout()
    subs = scenes.sponza_standin_submeshes(layout)
    assert len(subs) == 103
    assert sum(s["indices"].size for s in subs) // 3 == 262267
    assert sum(s["vertices"].shape[0] for s in subs) == 192496
    shared = {}
    attained = 0
    for s, p in zip(subs, layout["primitives"]):
        assert s["vertices"].shape[0] == p["vertexCount"] and s["indices"].size == p["indexCount"]
        assert s["indices"].max() < min(65536, p["vertexCount"])
        lo, hi = s["vertices"][:, :3].min(0), s["vertices"][:, :3].max(0)
        assert (lo >= np.float32(p["min"])).all() and (hi <= np.float32(p["max"])).all()
        attained += int(np.array_equal(lo, np.float32(p["min"])) and np.array_equal(hi, np.float32(p["max"])))
        key = p["indexAccessor"]
        if key in shared:
            assert np.array_equal(shared[key], s["indices"])  # primitives sharing one index accessor
        shared[key] = s["indices"]
        n = np.linalg.norm(s["vertices"][:, 4:7], axis=1)
        assert np.allclose(n, 1.0, atol=1e-3)
    assert attained >= 100
    assert not subs[50]["vertices"][:, 12:16].any()  # primitive 50 has no TANGENT accessor


def test_standin_is_deterministic():
    a, b = scenes.sponza_standin_submeshes(), scenes.sponza_standin_submeshes()
    assert all(np.array_equal(x["vertices"], y["vertices"]) and np.array_equal(x["indices"], y["indices"]) for x, y in zip(a, b))


def test_flatten_model_matches_reference_rules():
    m = scenes.analytic_scene()
    m["materials"][1] = (1, -1, -1)  # missing normal / metallic-roughness clamp to image 0, base colour is kept
    f = flatten_model(m)
    assert f["vertices"].shape == (12, 16) and f["indices"].size == 18
    assert f["indices"][6:12].tolist() == [4, 5, 6, 4, 6, 7]  # rebased by the running vertex count
    assert f["ranges"].tolist() == [[2, 3, 0], [2, 3, 24], [2, 3, 48]]
    assert f["rows"].tolist() == [[1, 2, 0, 0], [1, 0, 0, 2], [1, 2, 0, 4]]
