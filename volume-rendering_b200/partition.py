"""Image-space partition of one frame over the ranks of a multi-GPU job (the reference is single-GPU).

Rule (the same one the kernel and vr5_unpack_strips implement, vr5_cast.cuh / vr5_render.cu): the frame is cut into
strips of 4 rows (= the height of the 8x4 warp tile); strip s belongs to rank s % world. A rank's packed buffer holds
its strips back to back (strip j of the rank at rows 4j..4j+3), padded to the row count of the rank owning the most.
Interleaving balances the uneven ray lengths (centre rows carry far more samples than edge rows). Pure index helpers:
host-side plumbing for bench.py and the gloo test; no pixel is computed here.
"""
from __future__ import annotations

import numpy as np

from . import lib

STRIP = 4


def strip_rows(height: int, rank: int, world: int) -> int:
    """Rows of the packed buffer of `rank` (C ABI: vr5_strip_rows)."""
    return int(lib().vr5_strip_rows(height, rank, world))


def rows_padded(height: int, world: int) -> int:
    return max(strip_rows(height, r, world) for r in range(world))


def owned_rows(height: int, rank: int, world: int) -> np.ndarray:
    """Image rows rendered by `rank`, in packed order."""
    rows = np.arange(height)
    return rows[(rows // STRIP) % world == rank]


def pack(frame: np.ndarray, rank: int, world: int, padded_rows: int | None = None) -> np.ndarray:
    h = frame.shape[0]
    mine = owned_rows(h, rank, world)
    out = np.zeros((padded_rows if padded_rows is not None else len(mine),) + frame.shape[1:], dtype=frame.dtype)
    out[: len(mine)] = frame[mine]
    return out


def unpack(packed_all: np.ndarray, height: int) -> np.ndarray:
    """[world, rows_padded, W, C] -> [height, W, C]; the numpy statement of unpack_strips_kernel."""
    world = packed_all.shape[0]
    frame = np.zeros((height,) + packed_all.shape[2:], dtype=packed_all.dtype)
    for r in range(world):
        mine = owned_rows(height, r, world)
        frame[mine] = packed_all[r, : len(mine)]
    return frame
