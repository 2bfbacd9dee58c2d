"""Multi-GPU partition of a frame: image rows in bands of `band_rows`, band b owned by
rank b mod world; pixel indices and RNG streams stay global so the union of all ranks' rows is the
1-GPU image bit for bit.  The scene, BVHs and weights are replicated; the only exchange per frame is
the gather of the framebuffer to rank 0.  (The reference is single-GPU: this is an addition.)
"""
from __future__ import annotations

import numpy as np


def owned_rows(height: int, band_rows: int, rank: int, world: int) -> np.ndarray:
    """Global row indices owned by `rank`, ascending (= nrb_set_tile's rule in libnrb200)."""
    y = np.arange(height)
    return y[(y // band_rows) % world == rank]


def owned_bands(height: int, band_rows: int, rank: int, world: int):
    """[(y0, y1), ...] contiguous row ranges owned by `rank`."""
    out = []
    for y0 in range(rank * band_rows, height, world * band_rows):
        out.append((y0, min(height, y0 + band_rows)))
    return out


def gather_frame(frame, dst: int = 0, group=None):
    """The per-frame exchange: `frame` is this rank's full-size uint8 framebuffer in which only the
    owned rows are non-zero (libnrb200 zero-fills the buffer at allocation and writes owned rows
    only).  Rows are disjoint across ranks, so a SUM-reduce to `dst` is exactly a gather.  Works with
    any torch.distributed backend (NCCL on GPUs over NVLink; gloo in the CPU tests).  Returns the
    reduced tensor (valid on `dst`)."""
    import torch.distributed as dist
    out = frame.clone()
    dist.reduce(out, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return out
