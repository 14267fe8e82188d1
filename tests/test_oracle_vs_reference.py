"""CPU: pins the plain-C oracle (oracle/oracle.c) bit-for-bit against the host replay of the
reference's own headers (oracle/_ref), on scenes with duplicate Morton codes, transparent and
emissive subsets, degenerate uvs and textured materials.  Skipped only where neither the
prebuilt replay nor the reference sources exist."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import cabi, rays, scene as S
from tests import refreplay
from tests.oracle_lib import Oracle
from tests.parity import BVH_KEYS, TRI_KEYS, assert_bits_equal, point_from_bary

pytestmark = pytest.mark.skipif(not refreplay.available(), reason="reference replay unavailable")

SCENES = {
    "small": lambda: S.make_scene("small", n=20),
    "textured": lambda: S.make_scene("small", n=16, textured=True),
    "dense_duplicates": lambda: S.make_scene("small", n=72, glass=False, metal=False),
    "bottle": lambda: S.make_scene("C3", width=128, height=128),
}


@pytest.fixture(scope="module", params=sorted(SCENES))
def trio(request):
    sc = SCENES[request.param]()
    return sc, Oracle(sc), refreplay.RefReplay(sc)


def test_triangle_derived_data(trio):
    sc, orc, ref = trio
    a, b = orc.export_triangles(), ref.export_triangles()
    for k in TRI_KEYS:
        assert_bits_equal(a[k], b[k], k)


@pytest.mark.parametrize("which", [0, 1, 2])
def test_bvh(trio, which):
    sc, orc, ref = trio
    assert orc.num_leaves(which) == ref.num_leaves(which)
    if ref.num_leaves(which) < 2:
        pytest.skip("subset has fewer than 2 triangles")
    a, b = orc.export_bvh(which), ref.export_bvh(which)
    for k in BVH_KEYS:
        assert_bits_equal(a[k], b[k], k)


def test_trace(trio):
    sc, orc, ref = trio
    r = np.concatenate([rays.camera_rays(sc)[::5], rays.random_rays(3000),
                        rays.edge_case_rays(sc, 2048)])
    for which in (0, 1):
        if ref.num_leaves(which) < 2:
            continue
        a, b = orc.trace(r, which), ref.trace(r, which)
        for k in ("leaf_id", "tri_id", "t"):
            assert_bits_equal(a[k], b[k], k)
        assert_bits_equal(point_from_bary(sc, a["tri_id"], a["bary"]), b["point"], "point")
        assert (a["tri_id"] >= 0).any()


def test_duplicate_morton_codes_exist():
    """The tie-break path of delta() must actually be exercised by the test scenes."""
    sc = SCENES["dense_duplicates"]()
    m = Oracle(sc).export_bvh(0)["morton"]
    assert (np.diff(m.astype(np.int64)) == 0).sum() > 10
