#!/usr/bin/env python
"""BVH build throughput (BASELINE.json "BVH build Mtri/s"): builds and rebuilds the BLAS of the Sponza stand-in and of
the config-5 procedural soup at several sizes with device-resident input, and prints one JSON line per case.
Times are the library's own CUDA-event bracket around the whole builder (VkrtBlasInfo::buildMilliseconds)."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="262144,2000000,10000000,50000000")
    ap.add_argument("--repeat", type=int, default=5)
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--no-sponza", action="store_true")
    args = ap.parse_args()

    import torch

    from vkrt_b200 import binding as B
    from vkrt_b200 import scenes
    from vkrt_b200.raytracer import flatten_model
    from vkrt_b200.scenes_device import SoupDevice

    ctx = B.Context(0)
    peak = 6546.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass

    def report(name, blas, n, times):
        info = blas.info()
        best = min(times)
        # ~430 algorithmic bytes per triangle for the whole builder chain
        print(json.dumps({"case": name, "triangles": n, "build_ms_best": best, "build_ms_median": float(np.median(times)),
                          "mtri_s": n / best / 1e3, "nodes": int(info.nodeCount), "sah_cost": float(info.sahCost), "flags": args.flags,
                          "algorithmic_GBs": n * 430 / (best * 1e-3) / 1e9, "frac_of_measured_hbm": n * 430 / (best * 1e-3) / 1e9 / peak}),
              flush=True)

    if not args.no_sponza:
        flat = flatten_model(scenes.sponza_standin(max_texture_size=8))
        v = torch.from_numpy(flat["vertices"]).cuda()
        i = torch.from_numpy(flat["indices"].astype(np.int32)).cuda()
        torch.cuda.synchronize()
        blas = B.Blas(ctx, v, i, flat["ranges"], flags=args.flags)
        times = [blas.info().buildMilliseconds]
        for _ in range(args.repeat):
            blas.rebuild(v, flags=args.flags)
            times.append(blas.info().buildMilliseconds)
        report("sponza_standin", blas, int(blas.info().triangleCount), times[1:])
        blas.close()
        del v, i

    for n in [int(x) for x in args.sizes.split(",") if x]:
        soup = SoupDevice(n, 0)
        blas = B.Blas(ctx, soup.vertices, soup.indices, soup.ranges(), flags=args.flags)
        times = [blas.info().buildMilliseconds]
        for f in range(1, args.repeat + 1):
            blas.rebuild(soup.set_frame(f), flags=args.flags)
            times.append(blas.info().buildMilliseconds)
        report("soup", blas, n, times[1:])
        blas.close()
        del soup
        torch.cuda.empty_cache()
    ctx.close()


if __name__ == "__main__":
    main()
