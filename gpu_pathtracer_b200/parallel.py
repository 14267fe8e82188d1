"""Multi-GPU plumbing: one process per GPU, scene replicated, image tile-sharded.

Pixels are independent in path tracing and each pixel's RNG stream is keyed by its GLOBAL
pixel index (src/util/general.h:15, src/render/pathtracing.h:272-273), so any partition gives
bit-identical pixels.  The frame is cut into 32x32 tiles dealt round-robin to ranks
(gpt_render_pathtrace_device does the same dealing on the device side); the only
communication is the end-of-frame tile gather, implemented as a SUM reduction of frame
buffers whose non-owned pixels are zero.  SPPM adds one all-reduce of per-hit-point flux and
counts per pass (``allreduce_flux``).
"""
from __future__ import annotations

import numpy as np

TILE = 32


def tile_of_pixel(width: int, height: int) -> np.ndarray:
    """(H, W) int array: tile index of every pixel (row-major tiles of 32x32)."""
    tx = (width + TILE - 1) // TILE
    yy, xx = np.meshgrid(np.arange(height) // TILE, np.arange(width) // TILE, indexing="ij")
    return yy * tx + xx


def owner_mask(width: int, height: int, rank: int, nranks: int) -> np.ndarray:
    """(H, W) bool: pixels rendered by `rank` (tile % nranks == rank)."""
    return tile_of_pixel(width, height) % nranks == rank


def tile_pixels(width: int, height: int, rank: int, nranks: int) -> np.ndarray:
    """Slot -> global pixel index in the order the device uses: tile-major, rows inside a
    tile, so one warp covers one 32-pixel tile row."""
    tx, ty = (width + TILE - 1) // TILE, (height + TILE - 1) // TILE
    out = []
    for t in range(rank, tx * ty, nranks):
        x0, y0 = (t % tx) * TILE, (t // tx) * TILE
        ys = np.arange(y0, min(y0 + TILE, height))
        xs = np.arange(x0, min(x0 + TILE, width))
        out.append((ys[:, None] * width + xs[None, :]).reshape(-1))
    return np.concatenate(out) if out else np.zeros(0, np.int64)


def gather_frame(frame, dst: int = 0, group=None):
    """End-of-frame tile gather: every rank holds a full-size frame tensor that is zero
    outside its own tiles; a SUM reduce to `dst` assembles the image (exact: each pixel has
    one non-zero contributor)."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(frame, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return frame


def allreduce_flux(flux, group=None):
    """SPPM: per-hit-point (flux.rgb, photon count) of this pass summed over ranks."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flux, op=dist.ReduceOp.SUM, group=group)
    return flux


def photon_range(num_photons: int, rank: int, nranks: int):
    """Photons [begin, end) traced by `rank`; RNG sequence = global photon index."""
    per = (num_photons + nranks - 1) // nranks
    b = min(num_photons, rank * per)
    return b, min(num_photons, b + per)
