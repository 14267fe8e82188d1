"""Shared bit-exact comparison helpers for the parity tests."""
import numpy as np


def assert_bits_equal(a, b, name):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, f"{name}: shape {a.shape} vs {b.shape}"
    if a.dtype == np.float32:
        same = (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
    else:
        same = a == b
    assert same.all(), f"{name}: {np.count_nonzero(~same)} of {same.size} differ"


def point_from_bary(scene, tri_id, bary):
    """rec.point = b1*p1 + b2*p2 + b3*p3 in fp32, left to right (triangle.h:402-403)."""
    m = scene.mesh
    ok = tri_id >= 0
    t = np.where(ok, tri_id, 0)
    P = [np.stack([m["x"][m[k][t]], m["y"][m[k][t]], m["z"][m[k][t]]], 1)
         for k in ("point_1_idx", "point_2_idx", "point_3_idx")]
    b = bary.astype(np.float32)
    pt = (b[:, 0:1] * P[0] + b[:, 1:2] * P[1]) + b[:, 2:3] * P[2]
    return np.where(ok[:, None], pt, 0).astype(np.float32)


TRI_KEYS = ("tolerance", "aabb", "normal", "area", "flags", "energy", "tangents")
BVH_KEYS = ("world_aabb", "morton", "sorted_to_orig", "left", "right", "parent_internal",
            "parent_leaf", "leaf_aabb", "node_aabb")
