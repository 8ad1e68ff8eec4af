"""Times vkrt_rasterize (the reference's Rasterizer path) on the Sponza stand-in with CUDA events on the
library stream: ms per frame, Mtri/s submitted, Mpixel/s, and the byte model of DESIGN.md 5.4 against the measured HBM peak."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    import torch

    from vkrt_b200 import binding as B
    from vkrt_b200 import scenes
    from vkrt_b200.raytracer import Rasterizer

    rz = Rasterizer(scenes.sponza_standin(), args.width, args.height)
    stream = torch.cuda.ExternalStream(rz.ctx.stream, device=0)
    for _ in range(args.warmup):
        rz.render(B.TRACE_NO_AOV)
    rz.ctx.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            rz.render(B.TRACE_NO_AOV)
        e1.record(stream)
    rz.ctx.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    tris = int(rz.blas.info().triangleCount)
    px = args.width * args.height
    # bytes: keys cleared (64 B/pixel) + keys read by the resolve (64 B/pixel) + image (4 B/pixel) + set-up (76 B read, 96 B write per triangle)
    algo = px * (64 + 64 + 4) + tris * (76 + 96)
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    print(json.dumps({"metric": "rasterizer frame (Sponza stand-in, 8x MSAA)", "width": args.width, "height": args.height, "ms_per_frame": ms,
                      "mtri_s": tris / ms / 1e3, "mpixel_s": px / ms / 1e3, "algorithmic_bytes": algo, "achieved_gb_s": algo / ms / 1e6,
                      "measured_peaks": peak, "steps": args.steps, "warmup": args.warmup}))
    rz.close()


if __name__ == "__main__":
    main()
