"""Test-side ctypes wrapper of oracle/_ref/libref_replay.so (the reference's own headers
replayed on the host).  TEST INFRASTRUCTURE: never imported by gpu_pathtracer_b200/."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from gpu_pathtracer_b200 import cabi
from gpu_pathtracer_b200.scene import Scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_replay.so")


def build_ref_if_possible():
    """(Re)builds the replay library when the reference sources are present (this
    container); on the GPU box the prebuilt .so that travelled with the snapshot is used."""
    if os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True)
    return os.path.exists(REF_SO)


def available() -> bool:
    return os.path.exists(REF_SO) or build_ref_if_possible()


class RefError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise RefError(f"ref_replay returned {rc}")


class RefReplay:
    """Same surface as gpu_pathtracer_b200.api.PathTracer, backed by the reference's code."""

    def __init__(self, scene: Scene):
        if not available():
            raise RefError("oracle/_ref/libref_replay.so missing and reference sources absent")
        self.lib = C.CDLL(REF_SO)
        self.lib.ref_last_ms.restype = C.c_double
        self.lib.ref_last_rays.restype = C.c_uint64
        self.ctx = C.c_void_p()
        _check(self.lib.ref_create(C.byref(self.ctx)))
        self.scene = scene
        cabi.upload_scene(self.lib, "ref_", self.ctx, scene, _check)
        _check(self.lib.ref_build(self.ctx))
        self.n = scene.num_triangles

    def export_triangles(self):
        n = self.n
        out = dict(tolerance=np.zeros(n, np.float32), aabb=np.zeros((n, 6), np.float32),
                   normal=np.zeros((n, 3), np.float32), area=np.zeros(n, np.float32),
                   flags=np.zeros(n, np.int32), energy=np.zeros((n, 3), np.float32),
                   tangents=np.zeros((n, 9), np.float32))
        _check(self.lib.ref_export_triangles(
            self.ctx, cabi.fptr(out["tolerance"]), cabi.fptr(out["aabb"]),
            cabi.fptr(out["normal"]), cabi.fptr(out["area"]), cabi.iptr(out["flags"]),
            cabi.fptr(out["energy"]), cabi.fptr(out["tangents"])))
        return out

    def num_leaves(self, which=0) -> int:
        n = C.c_int32()
        _check(self.lib.ref_bvh_num_leaves(self.ctx, which, C.byref(n)))
        return n.value

    def export_bvh(self, which=0):
        n = self.num_leaves(which)
        out = dict(morton=np.zeros(n, np.uint32), sorted_to_orig=np.zeros(n, np.int32),
                   left=np.zeros(n - 1, np.int32), right=np.zeros(n - 1, np.int32),
                   parent_internal=np.zeros(n - 1, np.int32),
                   parent_leaf=np.zeros(n, np.int32),
                   node_aabb=np.zeros((n - 1, 6), np.float32),
                   leaf_aabb=np.zeros((n, 6), np.float32), world_aabb=np.zeros(6, np.float32))
        _check(self.lib.ref_export_bvh(
            self.ctx, which, cabi.uptr(out["morton"]), cabi.iptr(out["sorted_to_orig"]),
            cabi.iptr(out["left"]), cabi.iptr(out["right"]),
            cabi.iptr(out["parent_internal"]), cabi.iptr(out["parent_leaf"]),
            cabi.fptr(out["node_aabb"]), cabi.fptr(out["leaf_aabb"]),
            cabi.fptr(out["world_aabb"])))
        return out

    def export_light_tree(self):
        n = self.num_leaves(cabi.GPT_BVH_LIGHT)
        cone = np.zeros((n - 1, 5), np.float32)
        ne = np.zeros((n - 1, 3), np.float32)
        le = np.zeros((n, 3), np.float32)
        _check(self.lib.ref_export_light_tree(self.ctx, cabi.fptr(cone), cabi.fptr(ne),
                                              cabi.fptr(le)))
        return dict(node_cone=cone, node_energy=ne, leaf_energy=le)

    def trace(self, rays: np.ndarray, which=0, nthreads=1):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        out = dict(leaf_id=np.zeros(n, np.int32), tri_id=np.zeros(n, np.int32),
                   t=np.zeros(n, np.float32), point=np.zeros((n, 3), np.float32))
        _check(self.lib.ref_trace(self.ctx, which, cabi.fptr(rays), C.c_int64(n),
                                  cabi.iptr(out["leaf_id"]), cabi.iptr(out["tri_id"]),
                                  cabi.fptr(out["t"]), cabi.fptr(out["point"]),
                                  C.c_int(nthreads)))
        out["ms"] = self.lib.ref_last_ms(self.ctx)
        return out

    def primary_rays(self) -> np.ndarray:
        cam = self.scene.camera
        rays = np.zeros((cam["height"] * cam["width"], 8), np.float32)
        _check(self.lib.ref_primary_rays(self.ctx, cabi.fptr(rays)))
        return rays

    def render_pathtrace(self, window=None, nthreads=None) -> np.ndarray:
        cam = self.scene.camera
        H, W = cam["height"], cam["width"]
        if H < 100:
            # pathtracing.h:298 does `i % (height / 100)`: integer division by zero on a CPU
            raise RefError("the reference's progress printf needs image_height >= 100")
        r0, r1, c0, c1 = window if window is not None else (0, H, 0, W)
        out = np.zeros((H, W, 3), np.float32)
        nt = nthreads or os.cpu_count() or 1
        _check(self.lib.ref_render_pathtrace_window(
            self.ctx, cabi.fptr(out), r0, r1, c0, c1, C.c_int(nt)))
        self.render_ms = self.lib.ref_last_ms(self.ctx)
        self.render_rays = int(self.lib.ref_last_rays(self.ctx))
        return out

    def time_main_bvh_build(self) -> float:
        ms = C.c_double()
        _check(self.lib.ref_time_main_bvh_build(self.ctx, C.byref(ms)))
        return ms.value

    # -- SPPM: the reference's own kernels (oracle/ref_replay.cpp) --------------------------
    def sppm_init(self, nthreads=None):
        _check(self.lib.ref_sppm_init(self.ctx, C.c_int(nthreads or os.cpu_count() or 1)))
        self.sppm_ms = self.lib.ref_last_ms(self.ctx)
        self.sppm_rays = int(self.lib.ref_last_rays(self.ctx))

    def sppm_pass(self, pass_idx, photon_rng=0, stages=7, nthreads=None):
        """stages: 1 photon trace + trees, 2 eye pass, 4 gather + update (bit mask)."""
        _check(self.lib.ref_sppm_pass(self.ctx, C.c_int(pass_idx), C.c_int(photon_rng),
                                      C.c_int(stages), C.c_int(nthreads or os.cpu_count() or 1)))
        self.sppm_ms = self.lib.ref_last_ms(self.ctx)
        self.sppm_rays = int(self.lib.ref_last_rays(self.ctx))

    def sppm_counts(self) -> dict:
        ns, nv, r = C.c_int64(), C.c_int64(), C.c_double()
        _check(self.lib.ref_sppm_counts(self.ctx, C.byref(ns), C.byref(nv), C.byref(r)))
        return dict(n_surface=ns.value, n_volume=nv.value, mean_radius=r.value)

    def sppm_export_hit_points(self) -> dict:
        P = self.scene.camera["height"] * self.scene.camera["width"]
        keys = ("loc_radius", "normal_vradius", "filter", "direct", "indirect", "lm")
        out = {k: np.zeros((P, 4), np.float32) for k in keys}
        _check(self.lib.ref_sppm_export_hit_points(self.ctx, *[cabi.fptr(out[k]) for k in keys]))
        return out

    def sppm_export_photons(self) -> dict:
        n = int(self.scene.params["ppm_num_photon_per_pass"])
        kind = np.zeros(n, np.int32)
        rec = np.zeros((n, 12), np.float32)
        _check(self.lib.ref_sppm_export_photons(self.ctx, cabi.iptr(kind), cabi.fptr(rec)))
        return dict(kind=kind, loc=rec[:, 0:3], prev=rec[:, 3:6], dir=rec[:, 6:9],
                    color=rec[:, 9:12], rec=rec)

    def sppm_gather_only(self, nthreads=None) -> np.ndarray:
        P = self.scene.camera["height"] * self.scene.camera["width"]
        out = np.zeros((P, 4), np.float32)
        _check(self.lib.ref_sppm_gather_only(self.ctx, cabi.fptr(out),
                                             C.c_int(nthreads or os.cpu_count() or 1)))
        return out

    def point_image(self, which: int) -> np.ndarray:
        """The reference's create_point_image (point_operations.h:19-75), threads in index order:
        0 hit points, 1 / 2 surface / volume photons of the last photon pass, 3 SSS points."""
        cam = self.scene.camera
        out = np.zeros((cam["height"], cam["width"], 3), np.float32)
        _check(self.lib.ref_point_image(self.ctx, C.c_int(which), cabi.fptr(out)))
        return out

    def sppm_resolve(self, mode=0, passes=1) -> np.ndarray:
        """mode 0 global, 1 direct, 2 indirect (the C-ABI's numbering)."""
        cam = self.scene.camera
        out = np.zeros((cam["height"], cam["width"], 3), np.float32)
        ref_mode = {0: 2, 1: 0, 2: 1}[mode]
        _check(self.lib.ref_sppm_resolve(self.ctx, C.c_int(ref_mode), C.c_int(passes),
                                         cabi.fptr(out)))
        return out

    # -- SSS pre-pass (src/main.cu:555-824, 1213-1232) -------------------------------------
    def sss_export_points(self, first_pass=True, nthreads=None) -> dict:
        n = C.c_int64()
        nt = C.c_int(nthreads or os.cpu_count() or 1)
        _check(self.lib.ref_sss_export_points(self.ctx, 0, None, None, C.byref(n), nt))
        obj = np.zeros(n.value, np.int32)
        rec = np.zeros((n.value, 12), np.float32)
        if n.value:
            _check(self.lib.ref_sss_export_points(self.ctx, int(first_pass), cabi.iptr(obj),
                                                  cabi.fptr(rec), C.byref(n), nt))
        return dict(object_idx=obj, loc=rec[:, 0:3], filter=rec[:, 3:6], normal=rec[:, 6:9],
                    color=rec[:, 9:12], rec=rec)
