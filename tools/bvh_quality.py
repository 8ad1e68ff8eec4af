#!/usr/bin/env python
"""CPU-side BVH quality probe (test tier: uses the host emulation of the builder and traversal in tests/emu).
Builds the Sponza stand-in with the emulated builder, traces the config-1 primary rays of a reduced frame plus the
shadow rays towards the four lights, and prints nodes visited / triangles tested per ray.  Used to choose builder
heuristics before spending GPU time; set VKRT_EMU_DP_PRIM_COST (0 = greedy collapse)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import helpers as H
    from vkrt_b200 import binding as B
    from vkrt_b200 import camera, scenes
    from vkrt_b200.raytracer import flatten_model

    w, h = 320, 180
    flat = flatten_model(scenes.sponza_standin(max_texture_size=8))
    inst = np.array([[1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0]], dtype=np.float32)
    sc = H.EmuScene(flat, inst)
    nodes, tris = sc.read_blas()
    u = camera.make_uniforms(w, h)
    vi = np.array(u.viewInverse, dtype=np.float32).reshape(4, 4).T  # column-major glm
    pi = np.array(u.projInverse, dtype=np.float32).reshape(4, 4).T
    ys, xs = np.mgrid[0:h, 0:w]
    d = np.stack([(xs + 0.5) / w * 2 - 1, (ys + 0.5) / h * 2 - 1], axis=-1).reshape(-1, 2).astype(np.float32)
    target = (pi @ np.concatenate([d, np.ones((len(d), 2), np.float32)], axis=1).T).T
    t3 = target[:, :3] / np.linalg.norm(target[:, :3], axis=1, keepdims=True)
    dirs = (vi[:3, :3] @ t3.T).T
    origin = vi[:3, 3]
    rays = np.zeros(len(d), dtype=B.RAY_DTYPE)
    rays["origin"] = origin
    rays["direction"] = dirs
    rays["tmin"] = 0.001
    rays["tmax"] = 1000.0
    hits, n_nodes, n_tris = sc.trace_rays(rays)
    hit = hits["t"] < 1000.0
    print(f"nodes {len(nodes)} sah {sc.sah():.2f}")
    print(f"closest: {n_nodes / len(rays):.2f} nodes/ray {n_tris / len(rays):.2f} tris/ray  (hit {hit.mean():.3f})")
    pos = rays["origin"][hit] + rays["direction"][hit] * hits["t"][hit, None]
    lights = np.array(u.lightPositions, dtype=np.float32).reshape(4, 4)[:, :3]
    tn = tt = tr = 0
    for l in lights:
        v = l[None, :] - pos
        dist = np.linalg.norm(v, axis=1)
        sr = np.zeros(len(pos), dtype=B.RAY_DTYPE)
        sr["origin"] = pos
        sr["direction"] = v / dist[:, None]
        sr["tmin"] = 0.001
        sr["tmax"] = dist
        _, a, b = sc.trace_rays(sr, any_hit=True)
        tn += a
        tt += b
        tr += len(sr)
    print(f"any:     {tn / tr:.2f} nodes/ray {tt / tr:.2f} tris/ray")
    # instruction-weighted cost: node test ~205 instr at 27 lanes, triangle test ~144 at 13 lanes
    wn, wt = 205 / 27.0, 144 / 13.0
    cost = (n_nodes * wn + n_tris * wt + (tn * wn + tt * wt)) / (len(rays) + tr)
    print(f"weighted cost per ray {cost:.1f}")


if __name__ == "__main__":
    main()
