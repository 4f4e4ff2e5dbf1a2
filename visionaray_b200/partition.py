"""Image-space tile partition for multi-GPU frames: 64x64 tiles, tile k of
the frame belongs to rank k % nranks; every rank renders its tiles into a compact buffer
[tiles_of_rank, 64*64, 4]; one all-gather + de-interleave rebuilds the row-major image.  The numpy
functions here mirror `vsb_tiles_of_rank` / `deinterleave_kernel` for host-side tests and tools."""
from __future__ import annotations

import numpy as np

TILE = 64


def tiles_xy(width: int, height: int):
    return (width + TILE - 1) // TILE, (height + TILE - 1) // TILE


def tiles_of_rank(width: int, height: int, rank: int, nranks: int) -> int:
    tx, ty = tiles_xy(width, height)
    return (tx * ty - rank + nranks - 1) // nranks


def padded_tiles_per_rank(width: int, height: int, nranks: int) -> int:
    return max(tiles_of_rank(width, height, r, nranks) for r in range(nranks))


def pack_tiles(image: np.ndarray, rank: int, nranks: int) -> np.ndarray:
    """The compact tile buffer rank `rank` would render for `image` (h, w, 4), padded to the
    per-rank maximum (pixels outside the image stay zero)."""
    h, w = image.shape[:2]
    tx, _ = tiles_xy(w, h)
    per = padded_tiles_per_rank(w, h, nranks)
    out = np.zeros((per, TILE * TILE, 4), dtype=image.dtype)
    for local in range(tiles_of_rank(w, h, rank, nranks)):
        tile = local * nranks + rank
        x0, y0 = (tile % tx) * TILE, (tile // tx) * TILE
        block = np.zeros((TILE, TILE, 4), dtype=image.dtype)
        sub = image[y0:y0 + TILE, x0:x0 + TILE]
        block[: sub.shape[0], : sub.shape[1]] = sub
        out[local] = block.reshape(TILE * TILE, 4)
    return out


def deinterleave(gathered: np.ndarray, width: int, height: int, nranks: int) -> np.ndarray:
    """gathered: (nranks, tiles_per_rank_padded, 64*64, 4) -> (height, width, 4)."""
    tx, ty = tiles_xy(width, height)
    image = np.zeros((height, width, 4), dtype=gathered.dtype)
    for tile in range(tx * ty):
        rank, local = tile % nranks, tile // nranks
        x0, y0 = (tile % tx) * TILE, (tile // tx) * TILE
        block = gathered[rank, local].reshape(TILE, TILE, 4)
        image[y0:y0 + TILE, x0:x0 + TILE] = block[: min(TILE, height - y0), : min(TILE, width - x0)]
    return image
