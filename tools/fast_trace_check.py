"""Node visits and triangle tests per ray with and without VKRT_BUILD_PREFER_FAST_TRACE on scenes other than Sponza
(the schedule of the reinsertion passes was chosen on the Sponza stand-in): the config-5 triangle soup and a city-like
grid of boxes.  Prints one line per scene and flavour."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import helpers as H  # noqa: E402
from vkrt_b200 import binding as B  # noqa: E402
from vkrt_b200 import scenes  # noqa: E402
from vkrt_b200.raytracer import flatten_model  # noqa: E402


def boxes_scene(n_side=40, seed=3):
    """n_side^2 axis-aligned boxes of random height on a ground quad: 12 triangles each, large and small triangles mixed."""
    rng = np.random.default_rng(seed)
    pos, idx = [], []
    corners = np.array([[x, y, z] for x in (0, 1) for y in (0, 1) for z in (0, 1)], dtype=np.float32)
    faces = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    for i in range(n_side):
        for j in range(n_side):
            size = np.array([0.6, rng.uniform(0.5, 6.0), 0.6], dtype=np.float32)
            base = np.array([i - n_side / 2, 0.0, j - n_side / 2], dtype=np.float32)
            o = len(pos)
            pos.extend(base + corners * size)
            for a, b, c, d in faces:
                idx.extend([o + a, o + b, o + c, o + a, o + c, o + d])
    o = len(pos)
    g = n_side / 2 + 2
    pos.extend(np.array([[-g, 0, -g], [g, 0, -g], [g, 0, g], [-g, 0, g]], dtype=np.float32))
    idx.extend([o, o + 1, o + 2, o, o + 2, o + 3])
    pos = np.array(pos, dtype=np.float32)
    v = scenes.make_vertex_array(pos, np.tile([0, 1, 0], (len(pos), 1)), pos[:, [0, 2]], np.tile([1, 0, 0, 1], (len(pos), 1)))
    return {"submeshes": [{"vertices": v, "indices": np.array(idx, dtype=np.uint32), "material": 0}], "materials": [(1, 2, 0)],
            "images": scenes.analytic_scene()["images"]}, (-n_side / 2, 0.1, -n_side / 2), (n_side / 2, 7.0, n_side / 2)


def main():
    ctx = B.Context(0)
    cases = [("soup 500k", scenes.random_soup(500000, seed=5), (-12, -8, -10), (12, 8, 10))]
    m, lo, hi = boxes_scene()
    cases.append(("boxes 19k", m, lo, hi))
    for name, model, lo, hi in cases:
        flat = flatten_model(model)
        rays = H.random_rays(1000000, 4, lo, hi)
        rays["tmax"][::2] = 6.0
        base = None
        for flags in (B.BUILD_DEFAULT, B.BUILD_PREFER_FAST_TRACE):
            blas = B.Blas(ctx, flat["vertices"], flat["indices"], flat["ranges"], flags=flags)
            tlas = B.Tlas(ctx, [blas], scenes.identity_instance(1.0))
            out = []
            for any_hit in (False, True):
                ctx.stats_enable(True, False)
                ctx.stats_reset()
                hits = tlas.trace_rays(rays, any_hit=any_hit)
                st = ctx.stats()
                out.append((st.nodesVisited / len(rays), st.trianglesTested / len(rays), hits))
                ctx.stats_enable(False, False)
            info = blas.info()
            if base is None:
                base = out
            else:
                assert np.array_equal(base[0][2], out[0][2])
                assert np.array_equal(base[1][2]["instance"] != 0xFFFFFFFF, out[1][2]["instance"] != 0xFFFFFFFF)
            print(f"{name:10s} flags {flags}: build {info.buildMilliseconds:7.2f} ms  SAH {info.sahCost:7.2f}  closest {out[0][0]:6.2f} nodes {out[0][1]:6.2f} tris"
                  f"  any {out[1][0]:6.2f} nodes {out[1][1]:6.2f} tris", flush=True)
            tlas.close(), blas.close()
    ctx.close()


if __name__ == "__main__":
    main()
