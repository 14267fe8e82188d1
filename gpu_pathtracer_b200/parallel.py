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


def init_comm(pt, rank: int, nranks: int, group=None):
    """Bring up the LIBRARY's own NCCL communicator (gpt_comm_init, include/gpt.h "multi-GPU") on
    PathTracer `pt`.  torch.distributed is used for one thing only: to hand rank 0's 128-byte
    ncclUniqueId to the other ranks (any backend, gloo included); the collectives of the path then
    run inside libgpt.so on the context's stream."""
    import torch.distributed as dist
    if nranks <= 1:
        return
    box = [pt.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    pt.comm_init(box[0], rank, nranks)


def _lib_comm_ranks(pt) -> int:
    info = getattr(pt, "comm_info", None)
    return info()["nranks"] if info is not None else 1


class _DevicePtr:
    """A raw device pointer dressed up with __cuda_array_interface__ so torch can alias it."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = dict(shape=(n_floats,), typestr="<f4", data=(ptr, False),
                                             version=2)


def device_tensor(ptr: int, n_floats: int, device):
    """torch.float32 view (no copy) of library-owned device memory, e.g. the SPPM flux buffer
    of gpt_sppm_flux_device, so NCCL can reduce it in place."""
    import torch
    return torch.as_tensor(_DevicePtr(ptr, n_floats), device=device)


def sppm_frame(pt, rank: int = 0, nranks: int = 1, passes: int | None = None, flux_tensor=None,
               group=None, on_pass=None):
    """One render_mode-2 frame (src/main.cu:1282-1870) with the photons of every pass split by
    index range over the ranks: each rank traces its photons, builds its photon trees, runs the
    eye pass and gathers into ALL hit points; the per-hit-point accumulators are summed over
    ranks and every rank applies the same radius / flux update, so hit-point state stays
    replicated.  With the library's communicator up (``init_comm``) gpt_sppm_pass also
    all-gathers the volume photons and shards the eye pass by tiles (one more all-reduce hands
    every rank every hit point), and the sum is gpt_sppm_allreduce inside
    libgpt.so; otherwise the eye pass is replicated and (CPU tests over gloo,
    a caller that owns its own process group) it is torch.distributed on the aliased buffer.  `pt` is a PathTracer (or
    anything with its sppm_* methods); `flux_tensor()` returns the tensor aliasing the
    accumulator buffer (defaults to the device buffer of pt.sppm_flux_device())."""
    n_photons = int(pt.scene.params["ppm_num_photon_per_pass"])
    passes = passes or int(pt.scene.params["ppm_num_pass"])
    b, e = photon_range(n_photons, rank, nranks)
    pt.sppm_init()
    flux = None
    lib_comm = nranks > 1 and _lib_comm_ranks(pt) == nranks
    for p in range(passes):
        pt.sppm_pass(p, b, e)
        if lib_comm:
            # gpt_sppm_allreduce: ncclAllReduce on the context's stream, between the traversal and
            # the update half of the pass -- stream-ordered, no host synchronisation
            pt.sppm_allreduce()
        elif nranks > 1:
            if flux is None:
                if flux_tensor is not None:
                    flux = flux_tensor()
                else:
                    import torch
                    ptr, n = pt.sppm_flux_device()
                    flux = device_tensor(ptr, n, torch.device("cuda", torch.cuda.current_device()))
            pt.synchronize()
            allreduce_flux(flux, group=group)
            _sync_current_stream(flux)
        pt.sppm_update(p)
        if on_pass is not None:
            on_pass(p)
    return passes


def _sync_current_stream(t):
    if getattr(t, "is_cuda", False):
        import torch
        torch.cuda.current_stream(t.device).synchronize()


def allreduce_flux(flux, group=None):
    """SPPM: per-hit-point (flux.rgb, photon count, volume term) of this pass summed over
    ranks."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flux, op=dist.ReduceOp.SUM, group=group)
    return flux


def photon_range(num_photons: int, rank: int, nranks: int):
    """Photons [begin, end) traced by `rank`; RNG sequence = global photon index."""
    per = (num_photons + nranks - 1) // nranks
    b = min(num_photons, rank * per)
    return b, min(num_photons, b + per)
