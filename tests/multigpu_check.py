#!/usr/bin/env python
"""Run under torchrun on N GPUs: renders the Sponza stand-in split into interleaved tiles over N ranks, gathers the
BGRA8 frame on rank 0 over NCCL (vkrt_tiles_gather) and checks that it, and the copy that vkrt_frame_present makes while
the next frame is being traced, are byte-identical to a single-GPU render.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/multigpu_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vkrt_b200 import binding as B  # noqa: E402
from vkrt_b200 import scenes  # noqa: E402
from vkrt_b200.raytracer import Rasterizer, Raytracer  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # rank 0 decodes the textures; the other ranks receive them through vkrt_textures_broadcast (scene ingest over N ranks)
    model = scenes.sponza_standin(max_texture_size=256, with_images=(rank == 0))
    ok = True
    for (w, h, spp) in ((1000, 600, 1), (640, 360, 4)):
        ids = [B.Context.comm_unique_id() if rank == 0 else None]  # one NCCL communicator per Raytracer
        dist.broadcast_object_list(ids, src=0)
        rt = Raytracer(model, w, h, device=local, rank=rank, world_size=world, unique_id=ids[0], textures_from_root=0)
        rt.render(spp=spp)
        rt.ctx.synchronize()
        presented = None
        if rank == 0:
            got = rt.readback(B.AOV_BGRA8)
            # presentation ring on the root: the copy of this frame overlaps the next frame's trace + gather
            presented = rt.ctx.host_array((h, w, 4))
            ticket = rt.present(presented)
        rt.render(spp=spp)
        if rank == 0:
            rt.present_wait(ticket)
            presented = presented.copy()
        rt.ctx.synchronize()
        rt.close()
        if rank == 0:
            single = Raytracer(model, w, h, device=local)
            single.render(spp=spp)
            ref = single.readback(B.AOV_BGRA8)
            single.close()
            same = np.array_equal(got, ref) and np.array_equal(presented, ref)
            print(f"multigpu_check {w}x{h} spp={spp} ranks={world}: gathered == presented == single-GPU: {same}", flush=True)
            ok = ok and same
        dist.barrier()
    # the second renderer over the same tile split and gather
    ids = [B.Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    w, h = 1000, 600
    rz = Rasterizer(model, w, h, device=local, rank=rank, world_size=world, unique_id=ids[0], textures_from_root=0)
    rz.render()
    rz.ctx.synchronize()
    got = rz.readback(B.AOV_BGRA8) if rank == 0 else None
    rz.close()
    if rank == 0:
        single = Rasterizer(model, w, h, device=local)
        single.render()
        same = np.array_equal(got, single.readback(B.AOV_BGRA8))
        single.close()
        print(f"multigpu_check rasteriser {w}x{h} ranks={world}: gathered == single-GPU: {same}", flush=True)
        ok = ok and same
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
