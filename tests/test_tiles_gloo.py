"""CPU tier, world_size 2 over gloo: the host-side tile partition / gather logic of the N>1 path (each rank owns
tile i % size, sends its tile-major buffer to rank 0, rank 0 reassembles the frame)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vkrt_b200 import binding as B
from vkrt_b200 import tiles

W, H = 200, 150  # partial tiles on both axes


def _frame():
    y, x = np.mgrid[0:H, 0:W]
    return np.stack([x % 256, y % 256, (x * 7 + y * 13) % 256, np.full_like(x, 255)], axis=-1).astype(np.uint8)


def _worker(rank, size, port, out_path):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    mine = tiles.to_tile_major(_frame(), rank, size)  # what this rank's k_resolve_sample would have written
    counts = [len(tiles.owned_tiles(W, H, r, size)) for r in range(size)]
    if rank == 0:
        bufs = [mine]
        for r in range(1, size):
            t = torch.empty((counts[r], tiles.TILE, tiles.TILE, 4), dtype=torch.uint8)
            dist.recv(t, src=r)
            bufs.append(t.numpy())
        np.save(out_path, tiles.from_tile_major(bufs, W, H))
    else:
        dist.send(torch.from_numpy(mine), dst=0)
    dist.barrier()
    dist.destroy_process_group()


def test_tile_gather_world_size_2(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert np.array_equal(np.load(out), _frame())


def test_partition_matches_library():
    lib = B.load_library()
    for size in (1, 2, 3, 8):
        seen = np.zeros((H, W), dtype=np.int32)
        for r in range(size):
            tm = tiles.to_tile_major(np.ones((H, W, 1), dtype=np.int32), r, size)
            seen += tiles.from_tile_major([tm if k == r else np.zeros_like(tiles.to_tile_major(np.ones((H, W, 1), dtype=np.int32), k, size))
                                           for k in range(size)], W, H)[..., 0]
        assert (seen == 1).all()  # every pixel owned by exactly one rank
        for x, y in ((0, 0), (64, 0), (199, 149), (130, 70)):
            assert lib.vkrt_tile_owner(W, x, y, size) == tiles.owner(W, x, y, size)
