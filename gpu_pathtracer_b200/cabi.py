"""ctypes mirror of include/gpt.h (struct layouts + helpers to fill them from numpy).

Shared by the product wrapper (api.py) and by the test-side wrapper of the host replay
(tests/refreplay.py), which exports the same argument structs under a ``ref_`` prefix.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .scene import MAT_FLOAT_FIELDS, MAT_INT_FIELDS, Scene

PF = C.POINTER(C.c_float)
PI = C.POINTER(C.c_int32)
PU = C.POINTER(C.c_uint32)

GPT_BVH_MAIN, GPT_BVH_TRANSPARENT, GPT_BVH_LIGHT = 0, 1, 2
GPT_TRI_LIGHT, GPT_TRI_TRANSPARENT, GPT_TRI_SSS = 1, 2, 4

MESH_FLOAT = "x y z x_norm y_norm z_norm x_tex y_tex".split()
MESH_INT = ("point_1_idx point_2_idx point_3_idx norm_1_idx norm_2_idx norm_3_idx "
            "tex_1_idx tex_2_idx tex_3_idx material_idx triangle_object_idx "
            "object_num_primitives object_primitive_offset_idx").split()


class GptMesh(C.Structure):
    _fields_ = ([(n, C.c_int32) for n in ("num_vertices", "num_normals", "num_texcoords",
                                          "num_triangles", "num_objects")]
                + [(n, PF) for n in MESH_FLOAT] + [(n, PI) for n in MESH_INT])


class GptMaterials(C.Structure):
    _fields_ = ([("num_materials", C.c_int32)] + [(n, PF) for n in MAT_FLOAT_FIELDS]
                + [(n, PI) for n in MAT_INT_FIELDS])


class GptCamera(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("center", C.c_float * 3), ("up", C.c_float * 3),
                ("fovy", C.c_float), ("width", C.c_int32), ("height", C.c_int32),
                ("aperture", C.c_float), ("focus_dist", C.c_float)]


class GptParams(C.Structure):
    _fields_ = [("render_mode", C.c_int32), ("pathtracing_sample_size", C.c_int32),
                ("pathtracing_level", C.c_int32), ("dof_sample_size", C.c_int32),
                ("sss_pts_per_object", C.c_int32), ("hittable_pdf_weight", C.c_float),
                ("sky_emission", C.c_float * 3), ("ppm_num_photon_per_pass", C.c_int32),
                ("ppm_num_pass", C.c_int32), ("ppm_max_bounce", C.c_int32),
                ("ppm_alpha", C.c_float), ("ppm_radius_scaling_factor", C.c_float),
                ("rng_mode", C.c_int32)]


class GptStats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("build_all_ms", C.c_double),
                ("trace_ms", C.c_double), ("render_ms", C.c_double),
                ("rays_traced", C.c_uint64), ("internal_visits", C.c_uint64),
                ("leaf_tests", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("shade_events", C.c_uint64), ("trace_kernel_ms", C.c_double),
                ("shade_kernel_ms", C.c_double), ("trace_launches", C.c_uint64)]


def fptr(a: np.ndarray):
    assert a.dtype == np.float32 and a.flags.c_contiguous, (a.dtype, a.flags)
    return a.ctypes.data_as(PF)


def iptr(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags.c_contiguous, (a.dtype, a.flags)
    return a.ctypes.data_as(PI)


def uptr(a: np.ndarray):
    assert a.dtype == np.uint32 and a.flags.c_contiguous
    return a.ctypes.data_as(PU)


def pack_mesh(mesh: dict) -> GptMesh:
    m = GptMesh()
    m.num_vertices = mesh["x"].shape[0]
    m.num_normals = mesh["x_norm"].shape[0]
    m.num_texcoords = mesh["x_tex"].shape[0]
    m.num_triangles = mesh["point_1_idx"].shape[0]
    m.num_objects = mesh["object_num_primitives"].shape[0]
    for n in MESH_FLOAT:
        setattr(m, n, fptr(mesh[n]))
    for n in MESH_INT:
        setattr(m, n, iptr(mesh[n]))
    m._keep = mesh  # the arrays must outlive the struct
    return m


def pack_materials(mats: dict) -> GptMaterials:
    m = GptMaterials()
    m.num_materials = int(mats["num_materials"])
    for n in MAT_FLOAT_FIELDS:
        setattr(m, n, fptr(mats[n]))
    for n in MAT_INT_FIELDS:
        setattr(m, n, iptr(mats[n]))
    m._keep = mats
    return m


def pack_camera(cam: dict) -> GptCamera:
    c = GptCamera()
    c.eye[:] = [float(v) for v in cam["eye"]]
    c.center[:] = [float(v) for v in cam["center"]]
    c.up[:] = [float(v) for v in cam["up"]]
    c.fovy = float(cam["fovy"])
    c.width, c.height = int(cam["width"]), int(cam["height"])
    c.aperture, c.focus_dist = float(cam["aperture"]), float(cam["focus_dist"])
    return c


def pack_params(p: dict) -> GptParams:
    q = GptParams()
    for name, _ in GptParams._fields_:
        if name == "sky_emission":
            q.sky_emission[:] = [float(v) for v in p["sky_emission"]]
        else:
            setattr(q, name, p[name])
    return q


def upload_scene(lib, prefix: str, ctx, scene: Scene, check):
    """Calls <prefix>set_* for every part of ``scene`` on either library."""
    mesh, mats = pack_mesh(scene.mesh), pack_materials(scene.materials)
    cam, par = pack_camera(scene.camera), pack_params(scene.params)
    g = lambda n: getattr(lib, prefix + n)
    check(g("set_mesh")(ctx, C.byref(mesh)))
    check(g("set_materials")(ctx, C.byref(mats)))
    m = scene.materials
    check(g("set_textures")(ctx, fptr(m["tex_r"]), fptr(m["tex_g"]), fptr(m["tex_b"]),
                            C.c_int64(m["tex_r"].shape[0])))
    b = scene.bg
    check(g("set_background")(ctx, fptr(b["r"]), fptr(b["g"]), fptr(b["b"]),
                              C.c_int32(b["height"]), C.c_int32(b["width"])))
    check(g("set_camera")(ctx, C.byref(cam)))
    check(g("set_params")(ctx, C.byref(par)))
