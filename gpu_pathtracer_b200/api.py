"""Host-side mirror of the reference's launch sequence, bound to libgpt.so over ctypes.

``PathTracer`` plays the role of the part of ``main()`` that follows the loaders
(src/main.cu:474-1870): it takes the flat SoA the reference's OBJ/MTL loader produces (a
``Scene``), hands it to the C-ABI (include/gpt.h) and exposes build / trace / render.

There is NO CPU fallback: if the CUDA library is missing or no GPU is visible every call
raises ``GptError``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import cabi
from .scene import Scene

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GPT_LIB", os.path.join(_HERE, "csrc", "libgpt.so"))

# every symbol include/gpt.h declares
SYMBOLS = """gpt_create gpt_destroy gpt_last_error gpt_version gpt_set_stream gpt_set_mesh
gpt_set_materials gpt_set_textures gpt_set_background gpt_set_camera gpt_set_params gpt_build
gpt_rebuild_main_bvh gpt_export_triangles gpt_bvh_num_leaves gpt_export_bvh gpt_trace
gpt_trace_device gpt_trace_count_device gpt_render_pathtrace gpt_render_pathtrace_device
gpt_sppm_init gpt_sppm_pass gpt_sppm_flux_device gpt_sppm_update gpt_sppm_resolve
gpt_sppm_stats gpt_sppm_export_hit_points gpt_sppm_export_photons gpt_sppm_import_photons
gpt_sppm_pass_stages gpt_sppm_import_hit_points gpt_sss_export_points gpt_tonemap_u8 gpt_write_ppm_u8
gpt_save_image gpt_sppm_debug_image gpt_sss_debug_image gpt_stats
gpt_synchronize gpt_set_profiling gpt_set_wavefront_groups gpt_set_trace_tail gpt_set_trace_leaf_batching gpt_set_pipeline gpt_set_trace_bvh gpt_set_fast_rng_lanes gpt_sppm_debug_eye_shard
gpt_trace_static gpt_comm_unique_id gpt_comm_init gpt_comm_info gpt_comm_destroy gpt_sppm_allreduce gpt_gather_frame
gpt_scene_files_parse gpt_scene_files_view gpt_scene_files_material_name
gpt_scene_files_texture_name gpt_scene_files_free gpt_load_scene_files""".split()


class GptError(RuntimeError):
    pass


_lib = None


def write_ppm_u8(path: str, rgb_u8: np.ndarray):
    """ASCII PPM byte-identical to the reference's save_image (host only, no GPU needed)."""
    lib = load_library()
    rgb_u8 = np.ascontiguousarray(rgb_u8, dtype=np.uint8)
    h, w = rgb_u8.shape[:2]
    rc = lib.gpt_write_ppm_u8(path.encode(), rgb_u8.ctypes.data_as(C.POINTER(C.c_uint8)),
                              C.c_int32(w), C.c_int32(h))
    if rc != 0:
        raise GptError(f"gpt_write_ppm_u8 failed with {rc}")


def load_library():
    """dlopen the in-tree libgpt.so; raise loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GptError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` or `make -C gpu_pathtracer_b200/csrc`. There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.gpt_last_error.restype = C.c_char_p
    lib.gpt_version.restype = C.c_char_p
    lib.gpt_destroy.restype = None
    _lib = lib
    return lib


def load_scene_files(master_path: str, threads: int = 0, reader=None) -> Scene:
    """The scene packer (gpt_scene_files_parse): master file + OBJ + MTL + textures -> Scene.
    Host work inside libgpt.so; needs no GPU.  ``reader(user, path, h*, w*, rgb**)`` decodes
    image formats the library does not (JPEG: the host's libjpeg)."""
    lib = load_library()
    lib.gpt_scene_files_free.restype = None
    h, v = cabi.parse_scene_files(lib, "gpt_", master_path, threads, reader)
    try:
        return cabi.scene_from_view(lib, "gpt_", h, v)
    finally:
        lib.gpt_scene_files_free(h)


class PathTracer:
    def __init__(self, scene: Scene | None = None, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        rc = self.lib.gpt_create(C.c_int(device), C.byref(self.ctx))
        if rc != 0:
            raise GptError({-3: "no CUDA device visible (GPT_E_NODEVICE): this library has "
                                "no CPU path"}.get(rc, f"gpt_create failed with {rc}"))
        self.scene = None
        if scene is not None:
            self.set_scene(scene)

    # -- plumbing ---------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            msg = self.lib.gpt_last_error(self.ctx)
            raise GptError(f"gpt error {rc}: {msg.decode() if msg else ''}")

    def close(self):
        if self.ctx:
            self.lib.gpt_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        self._check(self.lib.gpt_set_stream(self.ctx, C.c_uint64(cuda_stream)))

    def synchronize(self):
        self._check(self.lib.gpt_synchronize(self.ctx))

    # -- scene ------------------------------------------------------------------------
    def set_scene(self, scene: Scene):
        self.scene = scene
        cabi.upload_scene(self.lib, "gpt_", self.ctx, scene, self._check)

    def load_files(self, master_path: str, threads: int = 0, reader=None):
        """Parses the reference's input files and uploads them in one go (main.cu:165-640):
        gpt_scene_files_parse + gpt_load_scene_files, no numpy round trip of the mesh."""
        self.lib.gpt_scene_files_free.restype = None
        h, v = cabi.parse_scene_files(self.lib, "gpt_", master_path, threads, reader)
        try:
            self._check(self.lib.gpt_load_scene_files(self.ctx, h))
            self.scene = cabi.scene_from_view(self.lib, "gpt_", h, v)
        finally:
            self.lib.gpt_scene_files_free(h)
        return self.scene

    def set_params(self, **kw):
        self.scene.params.update(kw)
        p = cabi.pack_params(self.scene.params)
        self._check(self.lib.gpt_set_params(self.ctx, C.byref(p)))

    def build(self):
        self._check(self.lib.gpt_build(self.ctx))
        return self.stats()

    def rebuild_main_bvh(self) -> float:
        self._check(self.lib.gpt_rebuild_main_bvh(self.ctx))
        return self.stats()["build_ms"]

    def set_wavefront_groups(self, groups=0):
        self._check(self.lib.gpt_set_wavefront_groups(self.ctx, C.c_int(int(groups))))

    def set_trace_tail(self, mode=-1):
        """1 / 0: cooperative tail of the persistent trace kernels on / off; -1: default."""
        self._check(self.lib.gpt_set_trace_tail(self.ctx, C.c_int(int(mode))))

    def set_trace_leaf_batching(self, mode=-1):
        """gpt_set_trace_leaf_batching: 1 batch a warp's triangle tests (deferred leaf tests),
        0 test every leaf where it is found, -1 default.  Bit-identical hits."""
        self._check(self.lib.gpt_set_trace_leaf_batching(self.ctx, C.c_int(int(mode))))

    # -- multi-GPU: the library's own NCCL communicator (include/gpt.h "multi-GPU") -----------
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = load_library().gpt_comm_unique_id(buf)
        if rc != 0:
            raise GptError(f"gpt_comm_unique_id failed with {rc} (libnccl.so.2 not loadable?)")
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, nranks: int):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._check(self.lib.gpt_comm_init(self.ctx, buf, C.c_int(rank), C.c_int(nranks)))

    def comm_info(self) -> dict:
        r, n, v = C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.gpt_comm_info(self.ctx, C.byref(r), C.byref(n), C.byref(v)))
        return dict(rank=r.value, nranks=n.value, nccl_version=v.value)

    def comm_destroy(self):
        self._check(self.lib.gpt_comm_destroy(self.ctx))

    def sppm_allreduce(self):
        """The per-pass all-reduce of the hit-point accumulators, on the context's stream."""
        self._check(self.lib.gpt_sppm_allreduce(self.ctx))

    def gather_frame(self, d_rgb: int, root: int = 0):
        self._check(self.lib.gpt_gather_frame(self.ctx, C.c_void_p(d_rgb), C.c_int(root)))

    def set_fast_rng_lanes(self, lanes=0):
        """rng_mode 1: independent sample streams per pixel (0 = default 8)."""
        self._check(self.lib.gpt_set_fast_rng_lanes(self.ctx, C.c_int(int(lanes))))

    def set_params(self, **kw):
        """Update render scalars (GptParams) of the current scene and re-send them."""
        self.scene.params.update(kw)
        p = cabi.pack_params(self.scene.params)
        self._check(self.lib.gpt_set_params(self.ctx, C.byref(p)))

    def set_trace_bvh(self, mode=-1):
        """gpt_set_trace_bvh: 1 walk the 4-wide tree, 0 the binary LBVH, -1 default ('wide' /
        'binary' accepted too).  Bit-identical hits."""
        mode = {"wide": 1, "binary": 0, "default": -1}.get(mode, mode)
        self._check(self.lib.gpt_set_trace_bvh(self.ctx, C.c_int(int(mode))))

    def set_pipeline(self, mode=-1):
        """gpt_set_pipeline: 0 rounds of launches, 1 streaming (two persistent kernels),
        -1 default; 'rounds' / 'stream' accepted too.  Bit-identical frames."""
        mode = {"rounds": 0, "stream": 1, "default": -1}.get(mode, mode)
        self._check(self.lib.gpt_set_pipeline(self.ctx, C.c_int(int(mode))))

    def set_profiling(self, time_stages=False, count_visits=False):
        self._check(self.lib.gpt_set_profiling(self.ctx, C.c_int(int(time_stages)),
                                               C.c_int(int(count_visits))))

    def stats(self) -> dict:
        s = cabi.GptStats()
        self._check(self.lib.gpt_stats(self.ctx, C.byref(s)))
        return {n: getattr(s, n) for n, _ in cabi.GptStats._fields_}

    # -- parity exports -----------------------------------------------------------------
    def export_triangles(self) -> dict:
        n = self.scene.num_triangles
        out = dict(tolerance=np.zeros(n, np.float32), aabb=np.zeros((n, 6), np.float32),
                   normal=np.zeros((n, 3), np.float32), area=np.zeros(n, np.float32),
                   flags=np.zeros(n, np.int32), energy=np.zeros((n, 3), np.float32),
                   tangents=np.zeros((n, 9), np.float32))
        self._check(self.lib.gpt_export_triangles(
            self.ctx, cabi.fptr(out["tolerance"]), cabi.fptr(out["aabb"]),
            cabi.fptr(out["normal"]), cabi.fptr(out["area"]), cabi.iptr(out["flags"]),
            cabi.fptr(out["energy"]), cabi.fptr(out["tangents"])))
        return out

    def num_leaves(self, which=0) -> int:
        n = C.c_int32()
        self._check(self.lib.gpt_bvh_num_leaves(self.ctx, which, C.byref(n)))
        return n.value

    def export_bvh(self, which=0) -> dict:
        n = self.num_leaves(which)
        out = dict(morton=np.zeros(n, np.uint32), sorted_to_orig=np.zeros(n, np.int32),
                   left=np.zeros(n - 1, np.int32), right=np.zeros(n - 1, np.int32),
                   parent_internal=np.zeros(n - 1, np.int32),
                   parent_leaf=np.zeros(n, np.int32),
                   node_aabb=np.zeros((n - 1, 6), np.float32),
                   leaf_aabb=np.zeros((n, 6), np.float32), world_aabb=np.zeros(6, np.float32))
        self._check(self.lib.gpt_export_bvh(
            self.ctx, which, cabi.uptr(out["morton"]), cabi.iptr(out["sorted_to_orig"]),
            cabi.iptr(out["left"]), cabi.iptr(out["right"]),
            cabi.iptr(out["parent_internal"]), cabi.iptr(out["parent_leaf"]),
            cabi.fptr(out["node_aabb"]), cabi.fptr(out["leaf_aabb"]),
            cabi.fptr(out["world_aabb"])))
        return out

    # -- trace --------------------------------------------------------------------------
    def trace(self, rays: np.ndarray, which=0) -> dict:
        """rays: (n, 8) float32 {ox,oy,oz,tmax,dx,dy,dz,0}, host memory."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        out = dict(leaf_id=np.zeros(n, np.int32), tri_id=np.zeros(n, np.int32),
                   t=np.zeros(n, np.float32), bary=np.zeros((n, 3), np.float32))
        self._check(self.lib.gpt_trace(self.ctx, which, cabi.fptr(rays), C.c_int64(n),
                                       cabi.iptr(out["leaf_id"]), cabi.iptr(out["tri_id"]),
                                       cabi.fptr(out["t"]), cabi.fptr(out["bary"])))
        return out

    TRACE_STATIC_MAX_RAYS = 131072  # GPT_TRACE_STATIC_MAX_RAYS (include/gpt.h)

    def trace_static(self, rays: np.ndarray, which=0) -> dict:
        """gpt_trace_static: gpt_trace with its staging and scratch in a static device array --
        no allocation of any kind, at most TRACE_STATIC_MAX_RAYS rays per call."""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        n = rays.shape[0]
        out = dict(leaf_id=np.zeros(n, np.int32), tri_id=np.zeros(n, np.int32),
                   t=np.zeros(n, np.float32), bary=np.zeros((n, 3), np.float32))
        self._check(self.lib.gpt_trace_static(self.ctx, which, cabi.fptr(rays), C.c_int64(n),
                                              cabi.iptr(out["leaf_id"]), cabi.iptr(out["tri_id"]),
                                              cabi.fptr(out["t"]), cabi.fptr(out["bary"])))
        return out

    def trace_device(self, d_rays: int, n: int, d_hits: int, which=0):
        """Device pointers (ints); asynchronous on the context's stream."""
        self._check(self.lib.gpt_trace_device(self.ctx, which, C.c_void_p(d_rays),
                                              C.c_int64(n), C.c_void_p(d_hits)))

    def trace_count_device(self, d_rays: int, n: int, which=0) -> dict:
        self._check(self.lib.gpt_trace_count_device(self.ctx, which, C.c_void_p(d_rays),
                                                    C.c_int64(n)))
        return self.stats()

    # -- render -------------------------------------------------------------------------
    def render_pathtrace(self, out: np.ndarray | None = None) -> np.ndarray:
        """Full frame into a host array (``out``: float32 [H, W, 3], C order; page-locked
        memory makes the read-back a plain DMA).  Every pixel is written."""
        cam = self.scene.camera
        shape = (cam["height"], cam["width"], 3)
        if out is None:
            out = np.empty(shape, np.float32)
        elif out.shape != shape:
            raise ValueError(f"out must be float32 {shape}, got {out.dtype} {out.shape}")
        self._check(self.lib.gpt_render_pathtrace(self.ctx, cabi.fptr(out)))
        return out

    def render_pathtrace_device(self, d_out: int, rank: int = 0, nranks: int = 1):
        self._check(self.lib.gpt_render_pathtrace_device(self.ctx, C.c_void_p(d_out),
                                                         C.c_int(rank), C.c_int(nranks)))

    # -- output stage (src/util/image_util.h:50-72) and debug splats -------------------------
    def tonemap_u8(self, rgb: np.ndarray) -> np.ndarray:
        rgb = np.ascontiguousarray(rgb, dtype=np.float32)
        out = np.zeros(rgb.shape, np.uint8)
        self._check(self.lib.gpt_tonemap_u8(self.ctx, cabi.fptr(rgb), C.c_int64(rgb.size),
                                            out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def save_image(self, rgb: np.ndarray, path: str):
        rgb = np.ascontiguousarray(rgb, dtype=np.float32)
        h, w = rgb.shape[:2]
        self._check(self.lib.gpt_save_image(self.ctx, cabi.fptr(rgb), C.c_int32(w), C.c_int32(h),
                                            path.encode()))

    def sppm_debug_image(self, which: int) -> np.ndarray:
        cam = self.scene.camera
        out = np.zeros((cam["height"], cam["width"], 3), np.float32)
        self._check(self.lib.gpt_sppm_debug_image(self.ctx, C.c_int(which), cabi.fptr(out)))
        return out

    def sss_debug_image(self) -> np.ndarray:
        cam = self.scene.camera
        out = np.zeros((cam["height"], cam["width"], 3), np.float32)
        self._check(self.lib.gpt_sss_debug_image(self.ctx, cabi.fptr(out)))
        return out

    # -- SSS (src/main.cu:642-824, 1213-1232) ---------------------------------------------
    def sss_export_points(self, first_pass: bool = True) -> dict:
        n = C.c_int64()
        self._check(self.lib.gpt_sss_export_points(self.ctx, 0, None, None, C.c_int64(0),
                                                   C.byref(n)))
        obj = np.zeros(n.value, np.int32)
        rec = np.zeros((n.value, 12), np.float32)
        if n.value:
            self._check(self.lib.gpt_sss_export_points(
                self.ctx, C.c_int(int(first_pass)), cabi.iptr(obj), cabi.fptr(rec),
                C.c_int64(n.value), C.byref(n)))
        return dict(object_idx=obj, loc=rec[:, 0:3], filter=rec[:, 3:6], normal=rec[:, 6:9],
                    color=rec[:, 9:12], rec=rec)

    # -- SPPM (render_mode 2, src/main.cu:1282-1870) ---------------------------------------
    def sppm_init(self):
        self._check(self.lib.gpt_sppm_init(self.ctx))

    def sppm_pass(self, pass_idx: int, photon_begin: int | None = None,
                  photon_end: int | None = None):
        if photon_begin is None:
            photon_begin, photon_end = 0, int(self.scene.params["ppm_num_photon_per_pass"])
        self._check(self.lib.gpt_sppm_pass(self.ctx, C.c_int(pass_idx), C.c_int64(photon_begin),
                                           C.c_int64(photon_end)))

    def sppm_flux_device(self):
        """(device pointer, float count) of the per-pass accumulator buffer to all-reduce."""
        p, n = C.c_void_p(), C.c_int64()
        self._check(self.lib.gpt_sppm_flux_device(self.ctx, C.byref(p), C.byref(n)))
        return p.value, n.value

    def sppm_update(self, pass_idx: int):
        self._check(self.lib.gpt_sppm_update(self.ctx, C.c_int(pass_idx)))

    def sppm_resolve(self, mode: int = 0, passes: int | None = None) -> np.ndarray:
        cam = self.scene.camera
        passes = passes or int(self.scene.params["ppm_num_pass"])
        out = np.zeros((cam["height"], cam["width"], 3), np.float32)
        self._check(self.lib.gpt_sppm_resolve(self.ctx, C.c_int(mode), C.c_int(passes),
                                              cabi.fptr(out)))
        return out

    def sppm_stats(self) -> dict:
        ns, nv, r, rays = C.c_int64(), C.c_int64(), C.c_double(), C.c_uint64()
        self._check(self.lib.gpt_sppm_stats(self.ctx, C.byref(ns), C.byref(nv), C.byref(r),
                                            C.byref(rays)))
        return dict(n_surface=ns.value, n_volume=nv.value, mean_radius=r.value,
                    rays_total=rays.value)

    def sppm_export_hit_points(self) -> dict:
        P = self.scene.camera["height"] * self.scene.camera["width"]
        out = {k: np.zeros((P, 4), np.float32)
               for k in ("loc_radius", "normal_vradius", "filter", "direct", "indirect")}
        out["flux_lm"] = np.zeros((2 * P, 4), np.float32)
        self._check(self.lib.gpt_sppm_export_hit_points(
            self.ctx, *[cabi.fptr(out[k]) for k in ("loc_radius", "normal_vradius", "filter",
                                                    "direct", "indirect", "flux_lm")]))
        out["flux"], out["lm"] = out["flux_lm"][:P], out["flux_lm"][P:]
        return out

    def sppm_export_photons(self) -> dict:
        n = C.c_int64()
        self._check(self.lib.gpt_sppm_export_photons(self.ctx, None, None, C.c_int64(0),
                                                     C.byref(n)))
        kind = np.zeros(n.value, np.int32)
        rec = np.zeros((n.value, 12), np.float32)
        self._check(self.lib.gpt_sppm_export_photons(self.ctx, cabi.iptr(kind), cabi.fptr(rec),
                                                     C.c_int64(n.value), C.byref(n)))
        return dict(kind=kind, loc=rec[:, 0:3], prev=rec[:, 3:6], dir=rec[:, 6:9],
                    color=rec[:, 9:12], rec=rec)

    def sppm_import_photons(self, kind: np.ndarray, rec12: np.ndarray):
        rec12 = np.ascontiguousarray(rec12, dtype=np.float32).reshape(-1, 12)
        kind = np.ascontiguousarray(kind, dtype=np.int32)
        self._check(self.lib.gpt_sppm_import_photons(self.ctx, cabi.iptr(kind), cabi.fptr(rec12),
                                                     C.c_int64(rec12.shape[0])))

    def sppm_pass_stages(self, pass_idx: int, stages: int, photon_begin: int | None = None,
                         photon_end: int | None = None):
        if photon_begin is None:
            photon_begin, photon_end = 0, int(self.scene.params["ppm_num_photon_per_pass"])
        self._check(self.lib.gpt_sppm_pass_stages(self.ctx, C.c_int(pass_idx),
                                                  C.c_int64(photon_begin), C.c_int64(photon_end),
                                                  C.c_int(stages)))

    def sppm_debug_eye_shard(self, rank: int = 0, nranks: int = 1):
        """Test hook: tile ownership of the sharded eye pass without a communicator."""
        self._check(self.lib.gpt_sppm_debug_eye_shard(self.ctx, C.c_int(rank), C.c_int(nranks)))

    def sppm_import_hit_points(self, loc_radius=None, normal_vradius=None, filter=None):
        arrs = [None if a is None else np.ascontiguousarray(a, dtype=np.float32)
                for a in (loc_radius, normal_vradius, filter)]
        self._check(self.lib.gpt_sppm_import_hit_points(
            self.ctx, *[None if a is None else cabi.fptr(a) for a in arrs]))

    def render_sppm(self, passes: int | None = None, mode: int = 0) -> np.ndarray:
        """The whole render_mode 2 frame on one GPU."""
        passes = passes or int(self.scene.params["ppm_num_pass"])
        self.sppm_init()
        for p in range(passes):
            self.sppm_pass(p)
            self.sppm_update(p)
        return self.sppm_resolve(mode, passes)
