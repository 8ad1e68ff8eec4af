"""Tile partition of a frame over ranks — numpy mirror of the layout libvkrt_b200 uses (64x64 tiles, tile i owned by
rank i % size, tile-major send buffers; csrc/shading.cuh pathToPixel, csrc/comm.cu k_untile).  Used by the host-side
multi-process tests and by callers that gather frames with torch.distributed instead of vkrt_tiles_gather."""
from __future__ import annotations

import numpy as np

TILE = 64


def tiles_xy(width: int, height: int):
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


def owner(width: int, x: int, y: int, size: int) -> int:
    tx, _ = tiles_xy(width, 1)
    return ((y // TILE) * tx + (x // TILE)) % size


def owned_tiles(width: int, height: int, rank: int, size: int) -> np.ndarray:
    tx, ty = tiles_xy(width, height)
    return np.arange(rank, tx * ty, size)


def to_tile_major(image: np.ndarray, rank: int, size: int) -> np.ndarray:
    """(H,W,C) -> (owned, 64, 64, C) send buffer of `rank` (pixels outside the image are zero)."""
    h, w = image.shape[:2]
    tx, _ = tiles_xy(w, h)
    tiles = owned_tiles(w, h, rank, size)
    out = np.zeros((len(tiles), TILE, TILE) + image.shape[2:], dtype=image.dtype)
    for k, t in enumerate(tiles):
        y0, x0 = (t // tx) * TILE, (t % tx) * TILE
        part = image[y0:y0 + TILE, x0:x0 + TILE]
        out[k, : part.shape[0], : part.shape[1]] = part
    return out


def from_tile_major(buffers, width: int, height: int) -> np.ndarray:
    """Inverse of to_tile_major over the buffers of all ranks (list indexed by rank)."""
    size = len(buffers)
    tx, _ = tiles_xy(width, height)
    first = buffers[0]
    image = np.zeros((height, width) + first.shape[3:], dtype=first.dtype)
    for rank, buf in enumerate(buffers):
        for k, t in enumerate(owned_tiles(width, height, rank, size)):
            y0, x0 = (t // tx) * TILE, (t % tx) * TILE
            hh, ww = min(TILE, height - y0), min(TILE, width - x0)
            image[y0:y0 + hh, x0:x0 + ww] = buf[k, :hh, :ww]
    return image
