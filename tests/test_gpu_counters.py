"""The visit counters behind the roofline numerator (algorithmic bytes of the
trace stage = rays * 48 + V_int * bytes per node + V_leaf * 48) against a host count with the
same culling rule (oracle.c orc_trace_count).  Exact for the ordered walk of the binary LBVH,
whose visit order is a function of the ray alone; the wide walk batches its triangle tests per
warp, so its counts depend on the schedule and are only bounded here."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import rays, scene as S

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scene_kind", ["small", "C1"])
def test_visit_counters_equal_a_host_count(scene_kind):
    import torch
    from gpu_pathtracer_b200.api import PathTracer
    from tests.oracle_lib import Oracle
    sc = S.make_scene("C1") if scene_kind == "C1" else S.make_scene("small", n=40, textured=True)
    pt = PathTracer(sc)
    pt.build()
    r = np.concatenate([rays.camera_rays(sc)[::11], rays.random_rays(20000), rays.random_rays(33, seed=3)])
    d_rays = torch.from_numpy(np.ascontiguousarray(r)).cuda()
    want = Oracle(sc).trace_count(r)
    # the plain ordered walk: one ray per lane from start to end, leaves tested where found
    pt.set_trace_bvh(0)
    pt.set_trace_tail(0)
    pt.set_trace_leaf_batching(0)
    got = pt.trace_count_device(d_rays.data_ptr(), r.shape[0])
    assert got["rays_traced"] == r.shape[0]
    assert got["internal_visits"] == want["internal_visits"], (got["internal_visits"], want["internal_visits"])
    assert got["leaf_tests"] == want["leaf_tests"], (got["leaf_tests"], want["leaf_tests"])
    # batched triangle tests cull with a staler bound: never fewer visits, and not many more
    pt.set_trace_leaf_batching(1)
    lazy = pt.trace_count_device(d_rays.data_ptr(), r.shape[0])
    assert want["internal_visits"] <= lazy["internal_visits"] <= 1.15 * want["internal_visits"]
    # the 4-wide tree: a visit tests four boxes, so far fewer visits; every ray still finds
    # every triangle the host walk found (the hits themselves are compared in test_gpu_wide.py)
    pt.set_trace_bvh(1)
    wide = pt.trace_count_device(d_rays.data_ptr(), r.shape[0])
    assert 0.2 * want["internal_visits"] <= wide["internal_visits"] <= 0.65 * want["internal_visits"]
    assert 0.8 * want["leaf_tests"] <= wide["leaf_tests"] <= 2.0 * want["leaf_tests"]
    pt.set_trace_bvh(-1)
    pt.set_trace_tail(-1)
    pt.set_trace_leaf_batching(-1)
