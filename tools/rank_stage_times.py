"""Stage times of ONE rank of an N-rank split, measured on one GPU (vkrt_comm_emulate): what a rank of the 8-GPU run
spends in each of the ten stages of a 4K 8 spp frame, without paying for 8 GPUs.
    python tools/rank_stage_times.py [--size 8] [--rank 0] [--frames 6]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vkrt_b200 import binding as B  # noqa: E402
from vkrt_b200 import scenes  # noqa: E402
from vkrt_b200.raytracer import Raytracer  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--spp", type=int, default=8)
    a = ap.parse_args()
    model = scenes.sponza_standin()
    rt = Raytracer(model, 3840, 2160, device=0, build_flags=B.BUILD_PREFER_FAST_TRACE, rank=a.rank, world_size=a.size,
                   emulate_ranks=a.size > 1)
    rt.ctx.stats_enable(False, True)
    rows = []
    for _ in range(a.frames):
        rt.render(spp=a.spp, flags=B.TRACE_NO_AOV, gather=False)
        rt.ctx.synchronize()
        st = rt.ctx.stats()
        rows.append(list(st.msStage) + [st.msTotal])
    m = np.median(np.array(rows[2:]), axis=0)
    names = ["raygen", "closest0", "surface0", "any0", "resolve0", "closest1", "surface1", "any1", "resolve1", "sample", "total"]
    print(f"rank {a.rank}/{a.size}: " + "  ".join(f"{n} {v:.3f}" for n, v in zip(names, m)), flush=True)
    # the same frames as a user runs them (no per-stage events: the bounce overlaps the depth-0 shadow rays), wall clock
    import time
    rt.ctx.stats_enable(False, False)
    for _ in range(3):
        rt.render(spp=a.spp, flags=B.TRACE_NO_AOV, gather=False)
    rt.ctx.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        rt.render(spp=a.spp, flags=B.TRACE_NO_AOV, gather=False)
    rt.ctx.synchronize()
    print(f"rank {a.rank}/{a.size}: untimed frame {(time.perf_counter() - t0) / 20 * 1e3:.3f} ms (stages serialised: {m[-1]:.3f})", flush=True)


if __name__ == "__main__":
    main()
