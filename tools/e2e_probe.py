#!/usr/bin/env python
"""Per-frame host-side timing of the config-2 end-to-end loop (pose update, vkrt_trace call, frame read-back)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
from vkrt_b200 import binding as B
from vkrt_b200.raytracer import Raytracer

model, inst = bench.build_scene("config2")
rt = Raytracer(model, 1920, 1080, device=0)
host = torch.empty((1080, 1920, 4), dtype=torch.uint8, pin_memory=True).numpy()
rows = []
for step in range(0, 320):
    t0 = time.perf_counter()
    bench.set_pose(rt, "config2", step)
    t1 = time.perf_counter()
    rt.render(spp=1, flags=B.TRACE_NO_AOV, gather=True)
    t2 = time.perf_counter()
    rt.ctx.synchronize()
    t3 = time.perf_counter()
    rt.readback(B.AOV_BGRA8, host)
    t4 = time.perf_counter()
    rows.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
a = np.array(rows[20:]) * 1e3
print("mean ms: pose %.3f  trace call %.3f  gpu wait %.3f  readback %.3f" % tuple(a.mean(0)))
print("max  ms: pose %.3f  trace call %.3f  gpu wait %.3f  readback %.3f" % tuple(a.max(0)))
for lo in range(0, 300, 50):
    print(lo, "gpu wait mean %.3f" % a[lo:lo + 50, 2].mean(), "trace call %.3f" % a[lo:lo + 50, 1].mean())
