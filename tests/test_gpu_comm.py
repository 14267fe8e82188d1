"""The library's own NCCL communicator (include/gpt.h "multi-GPU", csrc/comm.cu) through the
C-ABI.  One GPU here, so one rank: NCCL is loaded and a communicator is really created, and the
collectives must then be exact no-ops; the 2/4/8-rank data path is exercised by bench.py under
torchrun (SCALE runs) and its host logic by tests/test_parallel_gloo.py."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import scene as S

pytestmark = pytest.mark.gpu


def test_one_rank_communicator_and_noop_collectives():
    import torch
    from gpu_pathtracer_b200.api import GptError, PathTracer
    sc = S.make_scene("C4", cells=8, width=64, height=48, photons=20000, sphere_subdiv=2)
    pt = PathTracer(sc)
    assert pt.comm_info()["nranks"] == 1
    uid = PathTracer.comm_unique_id()
    assert len(uid) == 128 and any(uid)
    pt.comm_init(uid, 0, 1)
    info = pt.comm_info()
    assert info["rank"] == 0 and info["nranks"] == 1 and info["nccl_version"] >= 22000
    with pytest.raises(GptError):
        pt.comm_init(uid, 0, 1)  # already initialised
    pt.build()
    pt.sppm_init()
    pt.sppm_pass(0, 0, 20000)
    before = pt.sppm_export_hit_points()["flux_lm"].copy()
    pt.sppm_allreduce()
    pt.synchronize()
    assert np.array_equal(pt.sppm_export_hit_points()["flux_lm"], before)
    fb = torch.rand((48, 64, 3), device="cuda")
    keep = fb.clone()
    pt.gather_frame(fb.data_ptr(), 0)
    pt.synchronize()
    assert torch.equal(fb, keep)
    pt.comm_destroy()
    assert pt.comm_info()["nranks"] == 1
