"""The scene packer: gpt_scene_files_parse of libgpt.so against the
reference's OWN loaders (src/input/*.h compiled into oracle/_ref/libref_loader.so), and a
round trip through write_scene_files.  Host code on both sides: no GPU needed; every array is
compared BIT for bit (the floats go through from_chars on our side and stof on the
reference's; both are correctly rounded)."""
import ctypes as C
import os

import numpy as np
import pytest

from gpu_pathtracer_b200 import cabi, scene as S
from gpu_pathtracer_b200.api import load_library, load_scene_files

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LOADER = os.path.join(ROOT, "oracle", "_ref", "libref_loader.so")


def _ref_load(master):
    if not os.path.exists(REF_LOADER):
        import subprocess
        if os.path.isdir("/root/reference/src"):
            subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle"),
                            "_ref/libref_loader.so"], check=True)
    if not os.path.exists(REF_LOADER):
        pytest.skip("oracle/_ref/libref_loader.so not built and reference sources absent")
    lib = C.CDLL(REF_LOADER)
    lib.ref_scene_files_free.restype = None
    h, v = cabi.parse_scene_files(lib, "ref_", master)
    try:
        return cabi.scene_from_view(lib, "ref_", h, v)
    finally:
        lib.ref_scene_files_free(h)


def _same_bits(a, b, what):
    assert a.shape == b.shape, (what, a.shape, b.shape)
    if a.dtype == np.float32:
        a, b = a.view(np.uint32), b.view(np.uint32)
    bad = np.nonzero(a != b)[0]
    assert bad.size == 0, (what, bad[:5], a[bad[:5]], b[bad[:5]])


def _assert_same_scene(got: S.Scene, want: S.Scene):
    for n in cabi.MESH_FLOAT + cabi.MESH_INT:
        _same_bits(got.mesh[n], want.mesh[n], n)
    assert got.materials["material_names"] == want.materials["material_names"]
    assert got.materials["texture_names"] == want.materials["texture_names"]
    for n in (S.MAT_FLOAT_FIELDS + S.MAT_INT_FIELDS + ["tex_r", "tex_g", "tex_b", "texture_offset",
                                                       "texture_height", "texture_width"]):
        _same_bits(got.materials[n], want.materials[n], n)
    for c in "rgb":
        _same_bits(got.bg[c], want.bg[c], "bg_" + c)
    assert (got.bg["height"], got.bg["width"]) == (want.bg["height"], want.bg["width"])
    for k, v in want.camera.items():
        assert np.array_equal(np.float32(got.camera[k]).view(np.uint32),
                              np.float32(v).view(np.uint32)), k
    for k, v in want.params.items():
        if k != "rng_mode":
            assert np.array_equal(np.float32(got.params[k]), np.float32(v)), k
    for k in ("input_folder_path", "obj_filename", "texture_bg_path", "image_output_path",
              "ppm_image_output_iteration"):
        assert got.files[k] == want.files[k], k


@pytest.mark.parametrize("kind,kw", [
    ("small", dict(n=24, width=64, height=48, textured=True)),
    ("small", dict(n=12, width=32, height=32, sss_pts=50)),
    ("C3", dict(width=64, height=48)),
    ("C4", dict(width=48, height=32)),
])
def test_loader_matches_the_reference_loader(tmp_path, kind, kw):
    sc = S.make_scene(kind, **kw)
    master = S.write_scene_files(sc, str(tmp_path), "scene")
    want = _ref_load(master)
    for threads in (1, 3):
        got = load_scene_files(master, threads=threads)
        _assert_same_scene(got, want)


def test_round_trip_through_the_scene_writer(tmp_path):
    """write_scene_files -> loader gives back the generator's SoA: geometry and indices bit
    for bit, materials shifted by the loader's built-in Default_123 at slot 0."""
    sc = S.make_scene("small", n=20, width=64, height=48, textured=True)
    got = load_scene_files(S.write_scene_files(sc, str(tmp_path), "rt"), threads=2)
    for n in cabi.MESH_FLOAT:
        _same_bits(got.mesh[n], sc.mesh[n], n)
    for n in cabi.MESH_INT:
        if n == "material_idx":
            _same_bits(got.mesh[n], sc.mesh[n] + 1, n)
        elif n not in ("triangle_object_idx", "object_num_primitives", "object_primitive_offset_idx"):
            _same_bits(got.mesh[n], sc.mesh[n], n)
    # objects = runs of equal material in face order
    runs = 1 + int(np.count_nonzero(np.diff(sc.mesh["material_idx"])))
    assert got.mesh["object_num_primitives"].shape[0] == runs
    assert got.mesh["object_num_primitives"].sum() == sc.num_triangles
    assert got.materials["num_materials"] == sc.materials["num_materials"] + 1
    assert got.materials["material_names"][0] == "Default_123"
    for n in S.MAT_FLOAT_FIELDS:
        _same_bits(got.materials[n][1:], sc.materials[n], n)
    for c in "rgb":
        _same_bits(got.materials["tex_" + c], sc.materials["tex_" + c], "tex_" + c)
    assert got.camera["width"] == 64 and got.camera["height"] == 48


QUIRKY_MTL = """# comment
Kd 0.25 0.5 0.75
newmtl a
Kd 1 1 1\r
d 1.5
Ns -3
Ni 0.5
priority -7
newmtl   b\t
Tr 2.5
Ns 2000000
Ni 5000
priority 9
map_Bump -bm 0.125 bump.img
map_Kd kd.img
newmtl c
d 0.25
Ke +1.5 .5 1e1
"""

QUIRKY_OBJ = """mtllib q.mtl q2.mtl q.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0.5 0.5 1.25e-1
vn 0 0 1
vt 0.5 0.5
vt 1 0
usemtl a
f 1 2 3 4 5
usemtl b
f 1/1 2/2 3/1
f 1//1 2//1 3//1
usemtl c
  f   1/2/1\t2/1/1   3/2/1   4/1/1  \r
usemtl d
f 3 4 5
"""


def _write_img(path, h, w, seed):
    rng = np.random.default_rng(seed)
    S._write_image(path, rng.integers(0, 256, (h, w, 3), dtype=np.uint8))


def test_loader_quirks_match_the_reference(tmp_path):
    """Fan triangulation, per-face vt/vn detection from the first corner, \\r line ends, tabs,
    the inverted clamp() arguments, priority clamped from below only, a property line before
    the first newmtl, a second material file, a repeated mtllib, textures shared by name."""
    d = os.path.join(str(tmp_path), "")
    open(d + "q.mtl", "w", newline="").write(QUIRKY_MTL)
    open(d + "q2.mtl", "w").write("Ks 0.1 0.2 0.3\nnewmtl d\nmap_Ks kd.img\nmap_Ke second.img\n")
    open(d + "q.obj", "w", newline="").write(QUIRKY_OBJ)
    _write_img(d + "bump.img", 3, 5, 1)
    _write_img(d + "kd.img", 4, 2, 2)
    _write_img(d + "second.img", 1, 7, 3)
    _write_img(d + "bg.img", 2, 3, 4)
    master = d + "master.txt"
    with open(master, "w") as f:
        f.write(f"input_folder_path\t{d}\nobj_filename\tq.obj\ntexture_bg_path\tbg.img\n"
                "image_output_path\tout.ppm\nimage_width\t40\nimage_height\t30\nrender_mode\t1\n"
                "pathtracing_sample_size\t3\npathtracing_level\t4\ndof_sample_size\t5\n"
                "aperture\t0.5\nfocus_dist\t2.5\neye_x\t1\neye_y\t2\neye_z\t3\ncenter_x\t0\n"
                "center_y\t0\ncenter_z\t0\nup_x\t0\nup_y\t1\nup_z\t0\nfovy\t33.3\n"
                "sky_emission_r\t.1\nsky_emission_g\t.2\nsky_emission_b\t.3\n"
                "sss_pts_per_object\t7\nhittable_pdf_weight\t0.25\nppm_num_photon_per_pass\t11\n"
                "ppm_num_pass\t2\nppm_max_bounce\t3\nppm_alpha\t0.6\n"
                "ppm_radius_scaling_factor\t1.5\nppm_image_output_iteration\t4\n")
    want = _ref_load(master)
    got = load_scene_files(master, threads=1)
    _assert_same_scene(got, want)
    m = got.materials
    assert m["material_names"] == ["Default_123", "a", "b", "c", "d"]
    # what the quirks amount to (also pinned against the reference above)
    assert list(m["kd_x"][:2]) == [np.float32(0.25), 1.0]          # line before newmtl
    assert m["t_r"][1] == np.float32(1.0 - 1.5) and m["t_r"][2] == 2.5  # lower bound only
    assert m["n_s"][1] == 0.0 and m["n_s"][2] == 2000000.0
    assert m["n_i"][1] == 1.0 and m["n_i"][2] == 5000.0
    assert m["material_priority"][1] == -1 and m["material_priority"][2] == 9
    assert m["ks_x"][3] == np.float32(0.1)                        # q2.mtl edits material c
    assert got.params["dof_sample_size"] == 1                     # aperture came later
    assert got.mesh["point_1_idx"].shape[0] == 3 + 1 + 1 + 2 + 1  # fans
    assert list(got.mesh["object_num_primitives"]) == [3, 2, 2, 1]
    assert m["texture_names"] == ["Default_texture", "bump.img", "kd.img", "second.img"]
    assert list(m["texture_offset"]) == [0, 1, 16, 24]


def test_loader_errors_are_reported_not_fatal(tmp_path):
    d = os.path.join(str(tmp_path), "")
    _write_img(d + "bg.img", 1, 1, 0)
    base = f"input_folder_path\t{d}\ntexture_bg_path\tbg.img\nimage_width\t4\nimage_height\t4\n"

    def run(obj_text, master_extra="obj_filename\tx.obj\n"):
        open(d + "x.obj", "w").write(obj_text)
        open(d + "m.txt", "w").write(base + master_extra)
        return load_scene_files(d + "m.txt")

    with pytest.raises(ValueError, match="face before the first usemtl"):
        run("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(ValueError, match="does not have"):
        run("v 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl a\nf 1 2 9\n")
    with pytest.raises(ValueError, match="not a number"):
        run("v 0 zero 0\n")
    with pytest.raises(ValueError, match="no obj_filename"):
        run("", master_extra="")
    with pytest.raises(ValueError, match="cannot open"):
        load_scene_files(d + "does_not_exist.txt")
    open(d + "t.mtl", "w").write("newmtl a\nmap_Kd photo.jpg\n")
    open(d + "photo.jpg", "wb").write(b"\xff\xd8\xff\xe0 not really")
    with pytest.raises(ValueError, match="gpt_image_reader"):
        run("mtllib t.mtl\nv 0 0 0\n")
    # ... and with a reader the same file loads
    def reader(user, path, h, w, rgb):
        libc = C.CDLL(None)
        libc.malloc.restype = C.c_void_p
        h[0], w[0] = 2, 2
        p = libc.malloc(12)
        C.memmove(p, bytes(range(100, 112)), 12)
        rgb[0] = C.cast(p, C.POINTER(C.c_uint8))
        return 0
    open(d + "m.txt", "w").write(base.replace("bg.img", "photo.jpg") + "obj_filename\tx.obj\n")
    sc = load_scene_files(d + "m.txt", reader=reader)
    assert sc.materials["texture_names"] == ["Default_texture", "photo.jpg"]
    # rows flipped: atlas texel 1 is the first pixel of the LAST scan line
    assert sc.materials["tex_r"][1] == np.float32(106 / 255.0)
    assert sc.bg["height"] == 2 and sc.bg["r"][0] == np.float32(106 / 255.0)


def test_parallel_obj_parse_is_thread_count_invariant(tmp_path):
    """A 100K-triangle OBJ (several MB): 1, 2, 5 and 'all' threads give identical arrays, and
    the library exports every scene-packer symbol."""
    sc = S.make_scene("C1", width=32, height=32)
    master = S.write_scene_files(sc, str(tmp_path), "c1")
    base = load_scene_files(master, threads=1)
    assert base.files["threads"] == 1 and base.files["obj_bytes"] > 1 << 20
    for t in (2, 5, 0):
        got = load_scene_files(master, threads=t)
        for n in cabi.MESH_FLOAT + cabi.MESH_INT:
            _same_bits(got.mesh[n], base.mesh[n], (t, n))
    lib = load_library()
    for sym in ("gpt_scene_files_parse", "gpt_scene_files_view", "gpt_scene_files_free",
                "gpt_load_scene_files", "gpt_scene_files_material_name"):
        assert hasattr(lib, sym)


def _random_case(rng, d):
    """A random but well-formed OBJ + 1-2 MTL files using every construct the loaders know:
    polygons of 3-6 corners in the four corner forms, optional vn / vt, several usemtl runs,
    odd whitespace, CRLF, numbers in every spelling stof accepts."""
    def num(lo=-50.0, hi=50.0):
        v = rng.uniform(lo, hi)
        form = rng.integers(0, 6)
        if form == 0:
            return "%.9g" % np.float32(v)
        if form == 1:
            return "%d" % int(v)
        if form == 2:
            return "%.3e" % v
        if form == 3:
            return ("+" if v >= 0 else "") + "%.4f" % v
        if form == 4:
            return ("%.5f" % (v / 100.0)).replace("0.", ".", 1)
        return "%.12f" % v
    sep = lambda: rng.choice([" ", "  ", "\t", " \t "])
    eol = rng.choice(["\n", "\r\n"])
    nv, nn, nt = int(rng.integers(4, 40)), int(rng.integers(0, 8)), int(rng.integers(0, 8))
    mats = ["m%d" % i for i in range(int(rng.integers(1, 5)))]
    files = ["a.mtl"] + (["b.mtl"] if rng.random() < 0.5 else [])
    tex = []
    for fi, fn in enumerate(files):
        with open(d + fn, "w", newline="") as f:
            if fi == 1 and rng.random() < 0.5:
                f.write("Ks" + sep() + " ".join(num(0, 1) for _ in range(3)) + eol)
            for m in (mats if fi == 0 else ["x%d" % i for i in range(int(rng.integers(1, 3)))]):
                f.write("newmtl" + sep() + m + eol)
                for key in ("Ka", "Kd", "Ks", "Ke", "Tf"):
                    if rng.random() < 0.6:
                        f.write(key + sep() + sep().join(num(0, 2) for _ in range(3)) + eol)
                for key, lo, hi in (("d", -0.5, 1.5), ("Tr", -0.5, 1.5), ("Ns", -10, 2e6), ("Ni", 0, 3),
                                    ("path_length", 0, 1), ("scattering_coef", 0, 1),
                                    ("absorption_coef", 0, 1), ("g", -1, 1)):
                    if rng.random() < 0.4:
                        f.write(key + sep() + num(lo, hi) + eol)
                if rng.random() < 0.4:
                    f.write("priority" + sep() + str(int(rng.integers(-9, 9))) + eol)
                for key in ("map_Kd", "map_Ks", "map_Ke", "map_Ns"):
                    if rng.random() < 0.25:
                        name = "t%d.img" % int(rng.integers(0, 3))
                        tex.append(name)
                        f.write(key + sep() + name + eol)
                if rng.random() < 0.25:
                    name = "t%d.img" % int(rng.integers(0, 3))
                    tex.append(name)
                    f.write("map_Bump -bm " + num(0, 1) + " " + name + eol)
                f.write(eol)
    for name in set(tex):
        _write_img(d + name, int(rng.integers(1, 6)), int(rng.integers(1, 6)), int(rng.integers(1 << 30)))
    _write_img(d + "bg.img", int(rng.integers(1, 4)), int(rng.integers(1, 4)), 7)
    with open(d + "r.obj", "w", newline="") as f:
        f.write("# generated" + eol + "mtllib " + " ".join(files) + eol)
        for _ in range(nv):
            f.write("v" + sep() + sep().join(num() for _ in range(3)) + eol)
        for _ in range(nn):
            f.write("vn" + sep() + sep().join(num(-1, 1) for _ in range(3)) + eol)
        for _ in range(nt):
            f.write("vt" + sep() + sep().join(num(0, 1) for _ in range(2)) + eol)
        f.write(eol + "o thing" + eol + "s off" + eol)
        for _ in range(int(rng.integers(1, 6))):
            f.write("usemtl" + sep() + str(rng.choice(mats)) + eol)
            for _ in range(int(rng.integers(0, 12))):
                k = int(rng.integers(3, 7))
                form = int(rng.integers(0, 4))
                if form in (1, 3) and nt == 0:
                    form = 0
                if form in (2, 3) and nn == 0:
                    form = 0
                corners = []
                for _ in range(k):
                    p = int(rng.integers(1, nv + 1))
                    t = int(rng.integers(1, max(nt, 1) + 1))
                    n = int(rng.integers(1, max(nn, 1) + 1))
                    corners.append(["%d" % p, "%d/%d" % (p, t), "%d//%d" % (p, n),
                                    "%d/%d/%d" % (p, t, n)][form])
                f.write(rng.choice(["f", "  f"]) + sep() + sep().join(corners) + rng.choice(["", " "]) + eol)
    with open(d + "m.txt", "w") as f:
        f.write(f"input_folder_path\t{d}\nobj_filename\tr.obj\ntexture_bg_path\tbg.img\n"
                f"image_width\t{int(rng.integers(1, 99))}\nimage_height\t{int(rng.integers(1, 99))}\n"
                f"aperture\t{num(-1, 1)}\ndof_sample_size\t{int(rng.integers(1, 9))}\n"
                f"fovy\t{num(10, 90)}\neye_x\t{num()}\nrender_mode\t{int(rng.integers(1, 3))}\n")
    return d + "m.txt"


@pytest.mark.parametrize("seed", range(12))
def test_loader_matches_the_reference_loader_on_random_files(tmp_path, seed):
    rng = np.random.default_rng(1984 + seed)
    d = os.path.join(str(tmp_path), "")
    master = _random_case(rng, d)
    want = _ref_load(master)
    got = load_scene_files(master, threads=int(rng.integers(1, 4)))
    _assert_same_scene(got, want)
