"""GPU parity of the 4-wide walk (trace_wide.cuh, gpt_set_trace_bvh) at the sizes BASELINE.json
names: against the ordered walk of the binary LBVH, and against the host replay of the
reference's own traverse_bvh + Triangle::hit (oracle/_ref).  Bar: BIT-EXACT hit ids, t and
barycentrics, ties and degenerate rays included."""
import os

import numpy as np
import pytest

from gpu_pathtracer_b200 import cabi, rays, scene as S
from tests.parity import assert_bits_equal, point_from_bary

pytestmark = pytest.mark.gpu


def _bounce_set(sc, pt, stride, seed=7):
    """cosine-hemisphere bounce rays from the primary hits of every stride-th pixel."""
    r1 = rays.camera_rays(sc)[::stride].copy()
    h = pt.trace(r1)
    ok = h["tri_id"] >= 0
    d = r1[ok, 4:7].astype(np.float64)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    hp = r1[ok, 0:3] + h["t"][ok, None].astype(np.float64) * d
    nrm = pt.export_triangles()["normal"][h["tri_id"][ok]].astype(np.float64)
    flip = (nrm * d).sum(1) > 0
    nrm[flip] *= -1
    return rays.bounce_rays(hp, nrm, seed=seed)


def _both_walks(pt, r, which=cabi.GPT_BVH_MAIN):
    out = []
    for mode in (0, 1):
        pt.set_trace_bvh(mode)
        out.append(pt.trace(r, which))
    pt.set_trace_bvh(-1)
    for k in ("leaf_id", "tri_id", "t", "bary"):
        assert_bits_equal(out[0][k], out[1][k], f"binary vs wide walk: {k}")
    return out[1]


@pytest.mark.parametrize("scene_kind", ["small", "doubled", "C1", "C3"])
def test_wide_walk_equals_binary_walk(scene_kind):
    from gpu_pathtracer_b200.api import PathTracer
    from tests.test_gpu_build_trace import _doubled
    if scene_kind in ("C1", "C3"):
        sc = S.make_scene(scene_kind)
    else:
        sc = S.make_scene("small", n=40, textured=True)
    if scene_kind == "doubled":
        sc = _doubled(sc)
    pt = PathTracer(sc)
    pt.build()
    sets = [rays.random_rays(100000), rays.edge_case_rays(sc, 16384), rays.camera_rays(sc)[::3].copy(),
            rays.random_rays(37), _bounce_set(sc, pt, 5)]
    # tmax below / at / above the hit, and denormal / huge direction components
    weird = rays.random_rays(4096, seed=11)
    weird[0::4, 4] = 1e-30
    weird[1::4, 5] = -1e-42
    weird[2::4, 6] = 0.0
    weird[3::8, 3] = 2.5
    sets.append(weird)
    trees = (cabi.GPT_BVH_MAIN, cabi.GPT_BVH_TRANSPARENT)
    hits = 0
    for r in sets:
        for w in trees:
            if pt.num_leaves(w) < 2:
                continue
            hits += int((_both_walks(pt, r, w)["tri_id"] >= 0).sum())
    assert hits > 0
    if scene_kind == "doubled":  # every hit is an equal-t tie: the reference's visit rank decides
        from tests.refreplay import RefReplay
        ref = RefReplay(sc)
        pt.set_trace_bvh(1)
        a, b = pt.trace(sets[0]), ref.trace(sets[0])
        pt.set_trace_bvh(-1)
        assert_bits_equal(a["leaf_id"], b["leaf_id"], "doubled vs replay leaf_id")
        assert_bits_equal(a["t"], b["t"], "doubled vs replay t")


def test_r2_bounce_rays_c1_vs_replay():
    """SURVEY 8(d) R2 at BASELINE config 1 (100 K triangles): 1 M cosine-hemisphere bounce rays
    from the primary hits, against the reference replay (all host threads)."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay
    sc = S.make_scene("C1")
    pt = PathTracer(sc)
    pt.build()
    parts = [_bounce_set(sc, pt, 1, seed=s) for s in range(8)]
    r = np.concatenate(parts)[:1_000_000]
    assert r.shape[0] >= 500_000
    ref = RefReplay(sc)
    want = ref.trace(r, nthreads=os.cpu_count() or 1)
    got = _both_walks(pt, r)
    for k in ("leaf_id", "tri_id", "t"):
        assert_bits_equal(got[k], want[k], f"R2 on C1: {k}")
    assert_bits_equal(point_from_bary(sc, got["tri_id"], got["bary"]), want["point"], "R2 hit point")
    assert (got["tri_id"] >= 0).mean() > 0.05  # open terrain: most bounce rays leave the scene


def test_c2_size_build_and_trace_vs_replay():
    """BASELINE config 2 size (1.02 M triangles): LBVH export and 200 K primary + 200 K bounce
    rays, both walks, bit-exact against the reference replay."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay
    sc = S.make_scene("C2")
    pt = PathTracer(sc)
    pt.build()
    ref = RefReplay(sc)
    a, b = pt.export_bvh(0), ref.export_bvh(0)
    for k in ("world_aabb", "morton", "sorted_to_orig", "left", "right", "parent_internal",
              "parent_leaf", "leaf_aabb", "node_aabb"):
        assert_bits_equal(a[k], b[k], "C2 " + k)
    nt = os.cpu_count() or 1
    r1 = rays.camera_rays(sc)[::10].copy()
    r2 = _bounce_set(sc, pt, 6)
    for name, r in (("primary", r1), ("bounce", r2)):
        want = ref.trace(r, nthreads=nt)
        got = _both_walks(pt, r)
        for k in ("leaf_id", "tri_id", "t"):
            assert_bits_equal(got[k], want[k], f"C2 {name}: {k}")
        assert_bits_equal(point_from_bary(sc, got["tri_id"], got["bary"]), want["point"],
                          f"C2 {name}: hit point")
