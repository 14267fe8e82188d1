"""gpt_trace_static (include/gpt.h): the allocation-free variant of gpt_trace
(no counterpart in the reference).  Its staging and kernel
scratch live in a `static __device__` array, so the file is collected LAST: a failure here must
not hide any other test under `pytest -x`.  Bar: bit-exact against gpt_trace, on two calls in a
row (the second one shorter and on other rays: stale scratch of the first must not leak)."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import cabi, rays, scene as S
from tests.parity import assert_bits_equal

pytestmark = pytest.mark.gpu


def test_trace_static_equals_trace_on_two_calls_in_a_row():
    import torch
    from gpu_pathtracer_b200.api import GptError, PathTracer
    sc = S.make_scene("small", n=40, textured=True)
    pt = PathTracer(sc)
    pt.build()
    first = np.concatenate([rays.camera_rays(sc), rays.random_rays(80000, seed=5),
                            rays.edge_case_rays(sc, 8192)])
    second = np.concatenate([rays.random_rays(60000, seed=9), rays.camera_rays(sc)[::3]])
    assert second.shape[0] < first.shape[0] <= PathTracer.TRACE_STATIC_MAX_RAYS
    hits = 0
    for k, r in enumerate((first, second)):
        for which in (cabi.GPT_BVH_MAIN, cabi.GPT_BVH_TRANSPARENT):
            want = pt.trace(r, which)
            if k == 1 and which == cabi.GPT_BVH_MAIN:
                torch.cuda.synchronize()
                free_before = torch.cuda.mem_get_info()[0]
            got = pt.trace_static(r, which)
            if k == 1 and which == cabi.GPT_BVH_MAIN:
                # rays + hit records of this call would be 64 B x 76 K = 4.9 MB if they were allocated
                assert torch.cuda.mem_get_info()[0] >= free_before - (1 << 20)
            for key in ("leaf_id", "tri_id", "t", "bary"):
                assert_bits_equal(got[key], want[key], f"call {k} tree {which}: {key}")
            hits += int((got["tri_id"] >= 0).sum())
    assert hits > 10000
    # both walks of the tree, and the empty call
    for mode in (0, 1):
        pt.set_trace_bvh(mode)
        got, want = pt.trace_static(second), pt.trace(second)
        for key in ("leaf_id", "tri_id", "t", "bary"):
            assert_bits_equal(got[key], want[key], f"walk {mode}: {key}")
    pt.set_trace_bvh(-1)
    assert pt.trace_static(np.zeros((0, 8), np.float32))["t"].shape == (0,)
    # beyond the static scratch: a clear error, and the context stays usable
    too_many = rays.random_rays(PathTracer.TRACE_STATIC_MAX_RAYS + 1, seed=2)
    with pytest.raises(GptError, match="GPT_TRACE_STATIC_MAX_RAYS"):
        pt.trace_static(too_many)
    again = pt.trace_static(second)
    assert_bits_equal(again["t"], pt.trace(second)["t"], "after the rejected call")
