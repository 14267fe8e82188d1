"""CPU, world_size 2 over gloo: the multi-GPU host logic (tile dealing, end-of-frame gather,
SPPM flux all-reduce, photon split) without any GPU."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpu_pathtracer_b200 import parallel as P


def test_tiles_partition_the_image():
    for (w, h, n) in [(128, 128, 2), (1920, 1080, 8), (100, 70, 3), (33, 31, 4)]:
        cover = np.zeros(w * h, np.int32)
        for r in range(n):
            px = P.tile_pixels(w, h, r, n)
            cover[px] += 1
            assert np.array_equal(np.sort(px), np.flatnonzero(P.owner_mask(w, h, r, n).reshape(-1)))
        assert (cover == 1).all()


def test_photon_ranges_partition():
    for total, n in [(10_000_000, 8), (7, 3), (5, 8)]:
        spans = [P.photon_range(total, r, n) for r in range(n)]
        assert spans[0][0] == 0 and spans[-1][1] == total
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        W, H = 100, 70
        g = torch.Generator().manual_seed(1)
        truth = torch.rand((H, W, 3), generator=g)
        mine = torch.from_numpy(P.owner_mask(W, H, rank, world))
        frame = torch.where(mine[..., None], truth, torch.zeros(()))
        P.gather_frame(frame, dst=0)
        ok_frame = bool(torch.equal(frame, truth)) if rank == 0 else True
        flux = torch.full((H * W, 4), float(rank + 1))
        P.allreduce_flux(flux)
        ok_flux = bool((flux == sum(range(1, world + 1))).all())
        ret[rank] = ok_frame and ok_flux
    finally:
        dist.destroy_process_group()


class _FakeTracer:
    """Stands in for PathTracer in the SPPM frame loop: `flux` is what this rank's photons
    contribute, the update accumulates whatever the buffer holds after the all-reduce."""

    def __init__(self, n_photons, P=16):
        from types import SimpleNamespace
        self.scene = SimpleNamespace(params=dict(ppm_num_photon_per_pass=n_photons, ppm_num_pass=3))
        self.flux = torch.zeros(8 * P)
        self.acc = torch.zeros(8 * P)
        self.calls = []

    def sppm_init(self):
        self.calls.append("init")

    def sppm_pass(self, p, b, e):
        self.calls.append(("pass", p, b, e))
        self.flux[:] = float(e - b) * (p + 1)  # linear in the photon set

    def synchronize(self):
        pass

    def sppm_update(self, p):
        self.acc += self.flux
        self.calls.append(("update", p))


def _sppm_worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 1001
        pt = _FakeTracer(n)
        P.sppm_frame(pt, rank, world, flux_tensor=lambda: pt.flux)
        want = float(n) * (1 + 2 + 3)  # what a single rank tracing all photons accumulates
        order_ok = pt.calls[0] == "init" and [c[0] for c in pt.calls[1:]] == ["pass", "update"] * 3
        ret[rank] = bool((pt.acc == want).all()) and order_ok
    finally:
        dist.destroy_process_group()


def test_sppm_frame_photon_split_world2():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + (os.getpid() + 517) % 1000
    mp.spawn(_sppm_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret[r] for r in range(world)), dict(ret)


def test_sppm_frame_single_rank_needs_no_collective():
    pt = _FakeTracer(10)
    P.sppm_frame(pt)
    assert pt.acc[0] == 10 * 6


def test_gather_and_allreduce_world2():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    port = 29500 + os.getpid() % 1000
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    assert all(ret[r] for r in range(world)), dict(ret)
