"""CPU-only experiment (tests/emu/hierarchy_lab.hpp): node visits and triangle tests per ray on the Sponza stand-in for
different BINARY hierarchies under the same collapse and the same traversal code as the GPU (host emulation).

    python tools/hierarchy_lab.py [--width 160 --height 90] lbvh ploc sah lbvh+2 ploc+2 ...

"<mode>+<k>" runs k reinsertion passes on the binary tree.  Primary rays follow shader.rgen at the default pose; shadow
rays go from every hit point to the four lights (tmin 0.001, tmax = distance), like shader.rchit:132-143."""
import argparse
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]


def run_one(mode, width, height, npose=6):
    import helpers as H
    from vkrt_b200 import binding as B
    from vkrt_b200 import camera, scenes
    from vkrt_b200.raytracer import flatten_model

    flat = flatten_model(scenes.sponza_standin(with_images=False) | {"images": scenes.analytic_scene()["images"]})
    t0 = time.time()
    emu = H.EmuScene(flat, scenes.identity_instance(1.0))
    build_s = time.time() - t0
    poses = [((6.3, 4.5, -0.7), (0.0, 1.57, 0.0)), ((0.0, 2.0, 0.0), (0.0, 0.0, 0.0)), ((-6.0, 3.0, 0.5), (0.0, -1.57, 0.0)),
             ((3.0, 8.0, -2.0), (0.3, 3.0, 0.0)), ((-3.0, 1.5, 2.0), (0.0, 0.8, 0.0)), ((8.0, 2.0, 3.0), (-0.2, 2.2, 0.0))]
    tot = np.zeros(6)
    per_pose = []
    for position, rotation in poses[:npose]:
        u = camera.make_uniforms(width, height, position=position, rotation=rotation)
        vi = np.array(u.viewInverse, dtype=np.float32).reshape(4, 4).T  # column-major
        pi = np.array(u.projInverse, dtype=np.float32).reshape(4, 4).T
        ys, xs = np.mgrid[0:height, 0:width]
        px = (xs.ravel() + 0.5) / width * 2 - 1
        py = (ys.ravel() + 0.5) / height * 2 - 1
        target = (pi @ np.stack([px, py, np.ones_like(px), np.ones_like(px)]).astype(np.float32)).T
        d = target[:, :3] / np.linalg.norm(target[:, :3], axis=1, keepdims=True)
        d = (vi[:3, :3] @ d.T).T
        rays = np.zeros(len(d), dtype=B.RAY_DTYPE)
        rays["origin"] = vi[:3, 3]
        rays["direction"] = d.astype(np.float32)
        rays["tmin"], rays["tmax"] = 0.001, 10000.0
        hits, n_c, t_c = emu.trace_rays(rays)
        ok = hits["instance"] != 0xFFFFFFFF
        p = rays["origin"][ok] + rays["direction"][ok] * hits["t"][ok, None]
        lights = np.array(u.lightPositions, dtype=np.float32).reshape(4, 4)[:, :3]
        sh = np.zeros(len(p) * 4, dtype=B.RAY_DTYPE)
        for i in range(4):
            v = lights[i] - p
            dist = np.linalg.norm(v, axis=1)
            sh["origin"][i::4] = p
            sh["direction"][i::4] = v / dist[:, None]
            sh["tmax"][i::4] = dist
        sh["tmin"] = 0.001
        sh["origin"] += sh["direction"] * np.float32(0.01)  # off the surface, as the shading normal offset would
        ah, n_a, t_a = emu.trace_rays(sh, any_hit=True)
        tot += [n_c, t_c, len(rays), n_a, t_a, len(sh)]
        per_pose.append(f"{n_c / len(rays):.1f}/{n_a / max(len(sh), 1):.1f}")
    n_c, t_c, nr, n_a, t_a, ns = tot
    cost = (n_c * 180 + t_c * 137) / nr, (n_a * 180 + t_a * 137) / ns
    # probe rays that know nothing about the cameras: segments between random points on random triangles (area-weighted),
    # any-hit - what a builder could trace by itself to compare candidate hierarchies.  It does NOT predict the cost of the
    # camera and light rays: schedules that do badly there (band factor 2, 3.5) score better here than the shipped one.
    rng = np.random.default_rng(7)
    vtx = flat["vertices"][:, :3][flat["indices"].reshape(-1, 3)]
    tri_area = 0.5 * np.linalg.norm(np.cross(vtx[:, 1] - vtx[:, 0], vtx[:, 2] - vtx[:, 0]), axis=1)
    pick = rng.choice(len(vtx), size=(20000, 2), p=tri_area / tri_area.sum())
    uvw = rng.dirichlet((1, 1, 1), size=(20000, 2)).astype(np.float32)
    pts = (vtx[pick] * uvw[..., None]).sum(2)
    seg = pts[:, 1] - pts[:, 0]
    length = np.linalg.norm(seg, axis=1)
    okp = length > 1e-3
    pr = np.zeros(int(okp.sum()), dtype=B.RAY_DTYPE)
    pr["origin"] = pts[okp, 0]
    pr["direction"] = seg[okp] / length[okp, None]
    pr["tmin"], pr["tmax"] = length[okp] * 0.01, length[okp] * 0.99
    _, n_p, t_p = emu.trace_rays(pr, any_hit=True)
    probe = (n_p * 180 + t_p * 137) / len(pr)
    nodes, _ = emu.read_blas()
    # point-stabbing diagnostics: child boxes (inner children and leaf slots) that contain each camera position, and the
    # expected number of child boxes around a random point of the scene box (sum of volumes / root volume)
    scale = np.stack([(nodes[k].astype(np.uint32) << 16).view(np.float32) for k in ("sx", "sy", "sz")], axis=1)
    lo = np.stack([nodes["p"][:, k, None] + nodes[q].astype(np.float32) * scale[:, k, None] for k, q in enumerate(("qlox", "qloy", "qloz"))], axis=2)
    hi = np.stack([nodes["p"][:, k, None] + nodes[q].astype(np.float32) * scale[:, k, None] for k, q in enumerate(("qhix", "qhiy", "qhiz"))], axis=2)
    slot = np.arange(8)
    leaf_count = (nodes["leafBits"][:, None] >> (3 * slot)[None, :]) & 7
    valid = (((nodes["imask"][:, None] >> slot[None, :]) & 1) != 0) | (leaf_count != 0)
    stab = []
    for position, _ in poses[:npose]:
        q = np.array(position, dtype=np.float32)
        stab.append(int((valid & np.all((lo <= q) & (q <= hi), axis=2)).sum()))
    vol = np.prod(np.maximum(hi - lo, 0), axis=2)
    root_vol = float(np.prod(hi[0][valid[0]].max(0) - lo[0][valid[0]].min(0)))
    print(f"{mode:10s} boxes around the camera {stab}  expected boxes around a random point {float((vol * valid).sum()) / root_vol:.2f}")
    print(f"{mode:10s} wide nodes {len(nodes):6d}  wide SAH {emu.sah():7.2f}  closest {n_c / nr:6.2f} nodes {t_c / nr:6.2f} tris"
          f"  any {n_a / ns:6.2f} nodes {t_a / ns:6.2f} tris  est. instr closest {cost[0]:5.0f} any {cost[1]:5.0f}  per pose nodes c/a {' '.join(per_pose)}"
          f"  probe {probe:5.0f}  (build {build_s:.1f} s)", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=160)
    ap.add_argument("--height", type=int, default=90)
    ap.add_argument("--one", default=None)
    ap.add_argument("modes", nargs="*", default=["lbvh", "ploc", "sah"])
    a = ap.parse_args()
    if a.one:
        return run_one(a.one, a.width, a.height)
    for m in a.modes:
        env = dict(os.environ)
        base, _, par = m.partition("*")  # "<mode>*1": the builder's own parallel reinsertion (reinsert.cuh) on its default schedule, "<mode>*s,b,i,r": on another
        base, _, passes = base.partition("+")
        env["VKRT_EMU_BINARY"] = base
        env["VKRT_EMU_VERBOSE"] = "1"
        if passes:
            env["VKRT_EMU_REINSERT"] = passes
        if par:
            env["VKRT_EMU_FAST_TRACE"] = "1"
            if par != "1":
                env["VKRT_BUILD_REINSERT"] = par  # "<sweeps>,<bands>,<iterations>,<ratio>"
        subprocess.run([sys.executable, __file__, "--one", m, "--width", str(a.width), "--height", str(a.height)], env=env, check=True)


if __name__ == "__main__":
    main()
