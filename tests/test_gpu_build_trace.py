"""GPU parity: the CUDA path (through the C-ABI) against the host replay of the reference's
own headers (oracle/_ref) on identical inputs.  Bar: BIT-EXACT for everything here
(per-triangle derived data, Morton codes, permutation, topology, boxes, hit ids, t)."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import cabi, rays, scene as S

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pair():
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay
    sc = S.make_scene("small", n=40, textured=True)
    pt = PathTracer(sc)
    pt.build()
    return sc, pt, RefReplay(sc)


def _eq(a, b, name):
    a, b = np.asarray(a), np.asarray(b)
    same = (a.view(np.uint32) == b.view(np.uint32)) if a.dtype == np.float32 else (a == b)
    # NaN payloads may differ between host and device; NaN == NaN counts as equal
    if a.dtype == np.float32:
        same = same | (np.isnan(a) & np.isnan(b))
    assert same.all(), f"{name}: {np.count_nonzero(~same)} of {same.size} differ"


def test_triangle_derived_data_bit_exact(pair):
    sc, pt, ref = pair
    a, b = pt.export_triangles(), ref.export_triangles()
    for k in ("tolerance", "aabb", "normal", "area", "flags", "energy", "tangents"):
        _eq(a[k], b[k], k)


@pytest.mark.parametrize("which", [cabi.GPT_BVH_MAIN, cabi.GPT_BVH_TRANSPARENT,
                                   cabi.GPT_BVH_LIGHT])
def test_bvh_bit_exact(pair, which):
    sc, pt, ref = pair
    assert pt.num_leaves(which) == ref.num_leaves(which)
    a, b = pt.export_bvh(which), ref.export_bvh(which)
    for k in ("world_aabb", "morton", "sorted_to_orig", "left", "right", "parent_internal",
              "parent_leaf", "leaf_aabb", "node_aabb"):
        _eq(a[k], b[k], k)


def _point_from_bary(sc, tri_id, bary):
    m = sc.mesh
    ok = tri_id >= 0
    t = np.where(ok, tri_id, 0)
    P = [np.stack([m["x"][m[k][t]], m["y"][m[k][t]], m["z"][m[k][t]]], 1)
         for k in ("point_1_idx", "point_2_idx", "point_3_idx")]
    # rec.point = b1 * p1 + b2 * p2 + b3 * p3 in fp32, left to right (triangle.h:402-403)
    b = bary.astype(np.float32)
    pt = (b[:, 0:1] * P[0] + b[:, 1:2] * P[1]) + b[:, 2:3] * P[2]
    return np.where(ok[:, None], pt, 0).astype(np.float32)


@pytest.mark.parametrize("which", [cabi.GPT_BVH_MAIN, cabi.GPT_BVH_TRANSPARENT])
@pytest.mark.parametrize("kind", ["camera", "random", "edge"])
def test_trace_bit_exact(pair, which, kind):
    sc, pt, ref = pair
    r = {"camera": lambda: rays.camera_rays(sc),
         "random": lambda: rays.random_rays(20000),
         "edge": lambda: rays.edge_case_rays(sc, 8192)}[kind]()
    a, b = pt.trace(r, which), ref.trace(r, which)
    _eq(a["leaf_id"], b["leaf_id"], "leaf_id")
    _eq(a["tri_id"], b["tri_id"], "tri_id")
    _eq(a["t"], b["t"], "t")
    _eq(_point_from_bary(sc, a["tri_id"], a["bary"]), b["point"], "hit point")
    assert (a["tri_id"] >= 0).any()


def test_trace_tmax_and_empty(pair):
    sc, pt, ref = pair
    r = rays.camera_rays(sc)[::7].copy()
    full = pt.trace(r)
    hit = full["tri_id"] >= 0
    r2 = r.copy()
    r2[:, 3] = np.where(hit, full["t"], 1.0)          # tmax == t: still a hit (t > t_max rejects)
    again = pt.trace(r2)
    _eq(again["tri_id"][hit], full["tri_id"][hit], "tmax == t")
    r2[:, 3] = np.where(hit, full["t"] * np.float32(0.5), 1.0)
    none = pt.trace(r2)
    assert (none["tri_id"][hit] != full["tri_id"][hit]).all()
    assert pt.trace(np.zeros((0, 8), np.float32))["t"].shape == (0,)


def _doubled(sc):
    """Every triangle twice (same vertices): each hit is an equal-t tie that only the
    reference's visit rank decides."""
    import copy
    d = copy.deepcopy(sc)
    m = d.mesh
    for k in ("point_1_idx", "point_2_idx", "point_3_idx", "norm_1_idx", "norm_2_idx",
              "norm_3_idx", "tex_1_idx", "tex_2_idx", "tex_3_idx", "material_idx"):
        m[k] = np.concatenate([m[k], m[k]])
    n2 = m["point_1_idx"].shape[0]
    m["triangle_object_idx"] = np.zeros(n2, np.int32)
    m["object_num_primitives"] = np.array([n2], np.int32)
    m["object_primitive_offset_idx"] = np.array([0], np.int32)
    return d


@pytest.mark.parametrize("scene_kind", ["small", "doubled", "C1"])
def test_trace_tail_modes_agree(scene_kind):
    """The cooperative tail of the persistent trace (gpt_set_trace_tail) visits nodes in another
    order than the per-lane walk, and the batched triangle tests (gpt_set_trace_leaf_batching) test
    leaves later than they are found; hits must be bit-identical, ties included, and equal to
    the reference replay where that is affordable."""
    from gpu_pathtracer_b200.api import PathTracer
    sc = S.make_scene("C1") if scene_kind == "C1" else S.make_scene("small", n=40, textured=True)
    if scene_kind == "doubled":
        sc = _doubled(sc)
    pt = PathTracer(sc)
    pt.build()
    # few rays per warp at the end of the queue AND long incoherent walks: 3 batch sizes
    sets = [rays.random_rays(50000), rays.edge_case_rays(sc, 8192), rays.camera_rays(sc)[::3].copy(),
            rays.random_rays(37)]
    out = {}
    # ... and the same for the batched triangle tests (gpt_set_trace_leaf_batching): every
    # (tail, batching) combination against (0, 0), the plain test-where-found walk
    for mode in (0, 1, 2, 3):
        pt.set_trace_tail(mode & 1)
        pt.set_trace_leaf_batching(mode >> 1)
        trees = (cabi.GPT_BVH_MAIN,) if scene_kind == "C1" else (cabi.GPT_BVH_MAIN, cabi.GPT_BVH_TRANSPARENT)
        out[mode] = [pt.trace(r, w) for r in sets for w in trees]
    pt.set_trace_tail(-1)
    pt.set_trace_leaf_batching(-1)
    for mode in (1, 2, 3):
        for a, b in zip(out[0], out[mode]):
            for k in ("leaf_id", "tri_id", "t", "bary"):
                _eq(a[k], b[k], f"{scene_kind} mode {mode} {k}")
    assert any((a["tri_id"] >= 0).any() for a in out[1])
    if scene_kind == "doubled":
        from tests.refreplay import RefReplay
        ref = RefReplay(sc)
        want = ref.trace(sets[0], cabi.GPT_BVH_MAIN)
        _eq(out[3][0]["leaf_id"], want["leaf_id"], "doubled vs replay leaf_id")
        _eq(out[3][0]["t"], want["t"], "doubled vs replay t")


def test_against_committed_golden_vectors():
    """Same checks against tests/golden/small_scene.npz (made from the reference's own
    headers by tests/golden/make_golden.py): needs no reference code at run time."""
    import os
    from gpu_pathtracer_b200.api import PathTracer
    from tests.golden.make_golden import golden_scene
    from tests.parity import BVH_KEYS, TRI_KEYS, assert_bits_equal, point_from_bary
    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "small_scene.npz"))
    sc = golden_scene()
    pt = PathTracer(sc)
    pt.build()
    tri = pt.export_triangles()
    for k in TRI_KEYS:
        assert_bits_equal(tri[k], gold["tri_" + k], k)
    for which in (0, 1, 2):
        b = pt.export_bvh(which)
        for k in BVH_KEYS:
            assert_bits_equal(b[k], gold[f"bvh{which}_{k}"], f"bvh{which} {k}")
    for which in (0, 1):
        a = pt.trace(gold["rays"], which)
        for k in ("leaf_id", "tri_id", "t"):
            assert_bits_equal(a[k], gold[f"trace{which}_{k}"], k)
        assert_bits_equal(point_from_bary(sc, a["tri_id"], a["bary"]),
                          gold[f"trace{which}_point"], "hit point")


def test_large_scene_properties():
    """BASELINE config-1 size (100K triangles): size-independent properties instead of a
    replay -- sorted codes, permutation, every node/leaf referenced once, boxes nested,
    rebuild idempotent, and closest hits agree with the plain-C oracle on a ray sample."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.oracle_lib import Oracle
    from tests.parity import assert_bits_equal
    sc = S.make_scene("C1")
    pt = PathTracer(sc)
    pt.build()
    b = pt.export_bvh(0)
    n = b["morton"].shape[0]
    assert n == sc.num_triangles
    assert (np.diff(b["morton"].astype(np.int64)) >= 0).all()
    assert np.array_equal(np.sort(b["sorted_to_orig"]), np.arange(n))
    kids = np.concatenate([b["left"], b["right"]])
    assert np.array_equal(np.sort(kids[kids >= 0]), np.arange(1, n - 1))
    assert np.array_equal(np.sort(~kids[kids < 0]), np.arange(n))
    for side in ("left", "right"):
        c = b[side]
        cb = np.where((c >= 0)[:, None], b["node_aabb"][np.maximum(c, 0)],
                      b["leaf_aabb"][np.maximum(~c, 0)])
        assert (b["node_aabb"][:, 0::2] <= cb[:, 0::2]).all()
        assert (b["node_aabb"][:, 1::2] >= cb[:, 1::2]).all()
    pt.rebuild_main_bvh()
    b2 = pt.export_bvh(0)
    for k in b:
        assert_bits_equal(b[k], b2[k], "rebuild " + k)
    orc = Oracle(sc)
    ob = orc.export_bvh(0)
    for k in ("morton", "sorted_to_orig", "left", "right", "node_aabb"):
        assert_bits_equal(b[k], ob[k], "oracle " + k)
    r = np.concatenate([rays.camera_rays(sc)[::37], rays.random_rays(4000)])
    a, o = pt.trace(r), orc.trace(r)
    for k in ("leaf_id", "tri_id", "t"):
        assert_bits_equal(a[k], o[k], k)
