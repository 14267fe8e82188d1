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
                ("shade_kernel_ms", C.c_double), ("trace_launches", C.c_uint64),
                ("build_wide_ms", C.c_double), ("wide_node_visits", C.c_uint64),
                ("wide_leaf_tests", C.c_uint64)]


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


# ---- scene packer (include/gpt.h "scene packer") -------------------------------------------
class GptSceneFilesView(C.Structure):
    _fields_ = [("mesh", GptMesh), ("materials", GptMaterials), ("camera", GptCamera),
                ("params", GptParams),
                ("tex_r", PF), ("tex_g", PF), ("tex_b", PF), ("texture_length", C.c_int64),
                ("bg_r", PF), ("bg_g", PF), ("bg_b", PF),
                ("bg_height", C.c_int32), ("bg_width", C.c_int32),
                ("num_textures", C.c_int32),
                ("texture_offset", PI), ("texture_height", PI), ("texture_width", PI),
                ("input_folder_path", C.c_char_p), ("obj_filename", C.c_char_p),
                ("texture_bg_path", C.c_char_p), ("image_output_path", C.c_char_p),
                ("ppm_image_output_iteration", C.c_int32),
                ("parse_ms", C.c_double), ("obj_parse_ms", C.c_double),
                ("obj_bytes", C.c_int64), ("threads", C.c_int32)]


IMAGE_READER = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, PI, PI,
                           C.POINTER(C.POINTER(C.c_uint8)))


def _np(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def scene_from_view(lib, prefix: str, handle, v: GptSceneFilesView) -> Scene:
    """Copies a parsed scene (ours or the reference loader's) out of C memory into a Scene."""
    m, q = v.mesh, v.materials
    sizes = dict(x=m.num_vertices, y=m.num_vertices, z=m.num_vertices,
                 x_norm=m.num_normals, y_norm=m.num_normals, z_norm=m.num_normals,
                 x_tex=m.num_texcoords, y_tex=m.num_texcoords,
                 object_num_primitives=m.num_objects, object_primitive_offset_idx=m.num_objects)
    mesh = {n: _np(getattr(m, n), sizes[n], np.float32) for n in MESH_FLOAT}
    for n in MESH_INT:
        mesh[n] = _np(getattr(m, n), sizes.get(n, m.num_triangles), np.int32)
    name = getattr(lib, prefix + "scene_files_material_name")
    name.restype = C.c_char_p
    tname = getattr(lib, prefix + "scene_files_texture_name")
    tname.restype = C.c_char_p
    mats = {"num_materials": int(q.num_materials),
            "material_names": [name(handle, C.c_int32(i)).decode() for i in range(q.num_materials)]}
    for n in MAT_FLOAT_FIELDS:
        mats[n] = _np(getattr(q, n), q.num_materials, np.float32)
    for n in MAT_INT_FIELDS:
        mats[n] = _np(getattr(q, n), q.num_materials, np.int32)
    for c in "rgb":
        mats["tex_" + c] = _np(getattr(v, "tex_" + c), v.texture_length, np.float32)
    mats["texture_names"] = [tname(handle, C.c_int32(i)).decode() for i in range(v.num_textures)]
    for n in ("texture_offset", "texture_height", "texture_width"):
        mats[n] = _np(getattr(v, n), v.num_textures, np.int32)
    mats["images"] = []
    nbg = v.bg_height * v.bg_width
    bg = dict(r=_np(v.bg_r, nbg, np.float32), g=_np(v.bg_g, nbg, np.float32),
              b=_np(v.bg_b, nbg, np.float32), height=int(v.bg_height), width=int(v.bg_width))
    c = v.camera
    cam = dict(eye=tuple(c.eye), center=tuple(c.center), up=tuple(c.up), fovy=c.fovy,
               width=c.width, height=c.height, aperture=c.aperture, focus_dist=c.focus_dist)
    par = {n: getattr(v.params, n) for n, _ in GptParams._fields_ if n != "sky_emission"}
    par["sky_emission"] = tuple(v.params.sky_emission)
    sc = Scene(mesh=mesh, materials=mats, camera=cam, params=par, bg=bg,
               name=(v.obj_filename or b"scene").decode())
    sc.files = dict(input_folder_path=(v.input_folder_path or b"").decode(),
                    obj_filename=(v.obj_filename or b"").decode(),
                    texture_bg_path=(v.texture_bg_path or b"").decode(),
                    image_output_path=(v.image_output_path or b"").decode(),
                    ppm_image_output_iteration=int(v.ppm_image_output_iteration),
                    parse_ms=float(v.parse_ms), obj_parse_ms=float(v.obj_parse_ms),
                    obj_bytes=int(v.obj_bytes), threads=int(v.threads))
    return sc


def parse_scene_files(lib, prefix: str, master_path: str, threads: int = 0, reader=None):
    """<prefix>scene_files_parse -> (handle, view).  The caller frees the handle."""
    err = C.create_string_buffer(1024)
    h = C.c_void_p()
    if prefix == "gpt_":
        cb = IMAGE_READER(reader) if reader is not None else C.cast(None, IMAGE_READER)
        rc = lib.gpt_scene_files_parse(master_path.encode(), C.c_int(threads), cb, None,
                                       C.byref(h), err, C.c_int(len(err)))
    else:
        rc = getattr(lib, prefix + "scene_files_parse")(master_path.encode(), C.byref(h), err,
                                                       C.c_int(len(err)))
    if rc != 0:
        raise ValueError(f"scene loader: {err.value.decode(errors='replace')}")
    v = GptSceneFilesView()
    getattr(lib, prefix + "scene_files_view")(h, C.byref(v))
    return h, v
