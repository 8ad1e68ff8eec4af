#!/usr/bin/env python
"""Per-rank cost of the interleaved tile partition, measured on ONE GPU with emulated ranks (vkrt_comm_emulate):
renders the config-3 frame once per rank and prints device ms and ray counts, i.e. the load imbalance the real
multi-GPU run would see, separated from launch/tail effects (sum of the rank times vs the single-rank time)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--spp", type=int, default=8)
    args = ap.parse_args()
    from vkrt_b200 import binding as B
    from vkrt_b200 import scenes
    from vkrt_b200.raytracer import Raytracer

    model = scenes.sponza_standin()
    out = []
    for world in (1, args.ranks):
        for r in range(world):
            rt = Raytracer(model, 3840, 2160, device=0, rank=r, world_size=world, emulate_ranks=world > 1)
            for _ in range(2):
                rt.render(spp=args.spp, flags=B.TRACE_NO_AOV, gather=False)
            rt.ctx.synchronize()
            rt.ctx.stats_reset()
            t0 = time.perf_counter()
            for _ in range(3):
                rt.render(spp=args.spp, flags=B.TRACE_NO_AOV, gather=False)
            rt.ctx.synchronize()
            ms = (time.perf_counter() - t0) / 3 * 1e3
            st = rt.ctx.stats()
            rays = (st.raysPrimary + st.raysShadow + st.raysReflection) / 3
            out.append({"world": world, "rank": r, "ms": ms, "rays": rays, "mrays_s": rays / ms / 1e3})
            print(json.dumps(out[-1]), flush=True)
            rt.close()
    one = out[0]["ms"]
    per = [o["ms"] for o in out[1:]]
    print(json.dumps({"single_ms": one, "sum_rank_ms": sum(per), "max_rank_ms": max(per), "ideal_rank_ms": one / args.ranks,
                      "imbalance_max_over_mean": max(per) / (sum(per) / len(per)), "overhead_sum_over_single": sum(per) / one}))


if __name__ == "__main__":
    main()
