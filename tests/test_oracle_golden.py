"""CPU: the plain-C oracle against the committed golden vectors (tests/golden/small_scene.npz,
generated from the reference's own headers by tests/golden/make_golden.py).  Needs neither the
reference sources nor a GPU.  Bar: bit-exact."""
import os

import numpy as np
import pytest

from tests.golden.make_golden import golden_rays, golden_scene
from tests.oracle_lib import Oracle
from tests.parity import BVH_KEYS, TRI_KEYS, assert_bits_equal, point_from_bary

GOLD = os.path.join(os.path.dirname(__file__), "golden", "small_scene.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def orc():
    sc = golden_scene()
    return sc, Oracle(sc)


def test_fixture_inputs_are_reproducible(gold, orc):
    sc, _ = orc
    assert_bits_equal(golden_rays(sc), gold["rays"], "fixture rays")


def test_triangles(gold, orc):
    a = orc[1].export_triangles()
    for k in TRI_KEYS:
        assert_bits_equal(a[k], gold["tri_" + k], k)


@pytest.mark.parametrize("which", [0, 1, 2])
def test_bvh(gold, orc, which):
    a = orc[1].export_bvh(which)
    for k in BVH_KEYS:
        assert_bits_equal(a[k], gold[f"bvh{which}_{k}"], k)


@pytest.mark.parametrize("which", [0, 1])
def test_trace(gold, orc, which):
    sc, o = orc
    a = o.trace(gold["rays"], which)
    for k in ("leaf_id", "tri_id", "t"):
        assert_bits_equal(a[k], gold[f"trace{which}_{k}"], k)
    assert_bits_equal(point_from_bary(sc, a["tri_id"], a["bary"]), gold[f"trace{which}_point"],
                      "hit point")


def test_lbvh_invariants(orc):
    """Size-independent properties of any Karras tree, checked on the oracle's output."""
    b = orc[1].export_bvh(0)
    n = b["morton"].shape[0]
    assert (np.diff(b["morton"].astype(np.int64)) >= 0).all()            # sorted
    assert sorted(b["sorted_to_orig"].tolist()) == list(range(n))        # a permutation
    kids = np.concatenate([b["left"], b["right"]])
    assert sorted(kids[kids >= 0].tolist()) == list(range(1, n - 1))     # every node once
    assert sorted((~kids[kids < 0]).tolist()) == list(range(n))          # every leaf once
    assert b["parent_internal"][0] == -1
    # every node box contains its children's boxes
    for side in ("left", "right"):
        c = b[side]
        cb = np.where((c >= 0)[:, None], b["node_aabb"][np.maximum(c, 0)],
                      b["leaf_aabb"][np.maximum(~c, 0)])
        assert (b["node_aabb"][:, 0::2] <= cb[:, 0::2]).all()
        assert (b["node_aabb"][:, 1::2] >= cb[:, 1::2]).all()
