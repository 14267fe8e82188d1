"""Test-side ctypes wrapper of oracle/_build/liboracle.so (the plain-C restatement).
TEST INFRASTRUCTURE: never imported by gpu_pathtracer_b200/."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from gpu_pathtracer_b200 import cabi
from gpu_pathtracer_b200.scene import Scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_SO = os.path.join(ROOT, "oracle", "_build", "liboracle.so")


def _check(rc):
    if rc != 0:
        raise RuntimeError(f"oracle returned {rc}")


class Oracle:
    def __init__(self, scene: Scene):
        src = os.path.join(ROOT, "oracle", "oracle.c")
        if not os.path.exists(ORC_SO) or os.path.getmtime(ORC_SO) < os.path.getmtime(src):
            subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True)
        self.lib = C.CDLL(ORC_SO)
        self.lib.orc_destroy.restype = None
        self.ctx = C.c_void_p()
        _check(self.lib.orc_create(C.byref(self.ctx)))
        self.scene = scene
        mesh, mats = cabi.pack_mesh(scene.mesh), cabi.pack_materials(scene.materials)
        _check(self.lib.orc_set_mesh(self.ctx, C.byref(mesh)))
        _check(self.lib.orc_set_materials(self.ctx, C.byref(mats)))
        m = scene.materials
        _check(self.lib.orc_set_textures(self.ctx, cabi.fptr(m["tex_r"]), cabi.fptr(m["tex_g"]),
                                         cabi.fptr(m["tex_b"]), C.c_int64(m["tex_r"].shape[0])))
        _check(self.lib.orc_build(self.ctx))
        self.n = scene.num_triangles

    def __del__(self):
        try:
            self.lib.orc_destroy(self.ctx)
        except Exception:
            pass

    def export_triangles(self):
        n = self.n
        out = dict(tolerance=np.zeros(n, np.float32), aabb=np.zeros((n, 6), np.float32),
                   normal=np.zeros((n, 3), np.float32), area=np.zeros(n, np.float32),
                   flags=np.zeros(n, np.int32), energy=np.zeros((n, 3), np.float32),
                   tangents=np.zeros((n, 9), np.float32))
        _check(self.lib.orc_export_triangles(
            self.ctx, cabi.fptr(out["tolerance"]), cabi.fptr(out["aabb"]),
            cabi.fptr(out["normal"]), cabi.fptr(out["area"]), cabi.iptr(out["flags"]),
            cabi.fptr(out["energy"]), cabi.fptr(out["tangents"])))
        return out

    def num_leaves(self, which=0) -> int:
        n = C.c_int32()
        _check(self.lib.orc_bvh_num_leaves(self.ctx, which, C.byref(n)))
        return n.value

    def export_bvh(self, which=0):
        n = self.num_leaves(which)
        out = dict(morton=np.zeros(n, np.uint32), sorted_to_orig=np.zeros(n, np.int32),
                   left=np.zeros(n - 1, np.int32), right=np.zeros(n - 1, np.int32),
                   parent_internal=np.zeros(n - 1, np.int32),
                   parent_leaf=np.zeros(n, np.int32),
                   node_aabb=np.zeros((n - 1, 6), np.float32),
                   leaf_aabb=np.zeros((n, 6), np.float32), world_aabb=np.zeros(6, np.float32))
        _check(self.lib.orc_export_bvh(
            self.ctx, which, cabi.uptr(out["morton"]), cabi.iptr(out["sorted_to_orig"]),
            cabi.iptr(out["left"]), cabi.iptr(out["right"]),
            cabi.iptr(out["parent_internal"]), cabi.iptr(out["parent_leaf"]),
            cabi.fptr(out["node_aabb"]), cabi.fptr(out["leaf_aabb"]),
            cabi.fptr(out["world_aabb"])))
        return out

    def trace(self, rays: np.ndarray, which=0):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        out = dict(leaf_id=np.zeros(n, np.int32), tri_id=np.zeros(n, np.int32),
                   t=np.zeros(n, np.float32), bary=np.zeros((n, 3), np.float32))
        _check(self.lib.orc_trace(self.ctx, which, cabi.fptr(rays), C.c_int64(n),
                                  cabi.iptr(out["leaf_id"]), cabi.iptr(out["tri_id"]),
                                  cabi.fptr(out["t"]), cabi.fptr(out["bary"])))
        return out

    def trace_count(self, rays: np.ndarray, which=0) -> dict:
        """V_int / V_leaf of the ordered, culling walk of the CUDA binary-tree kernels
        (oracle.c orc_trace_count)."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        vi, vl = C.c_uint64(), C.c_uint64()
        leaf = np.zeros(n, np.int32)
        _check(self.lib.orc_trace_count(self.ctx, which, cabi.fptr(rays), C.c_int64(n),
                                        C.byref(vi), C.byref(vl), cabi.iptr(leaf)))
        return dict(internal_visits=vi.value, leaf_tests=vl.value, leaf_id=leaf)
