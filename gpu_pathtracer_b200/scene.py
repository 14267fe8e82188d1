"""Procedural scenes in the reference's flat SoA layout.

The reference ships no scene files, so every config of BASELINE.json is
generated here, deterministically, straight into the arrays that cross the C-ABI
(include/gpt.h GptMesh / GptMaterials, i.e. src/main.cu:386-471 and :139-156).  The same
dictionary can be written out as OBJ + MTL + master file (``write_obj``) so the reference's
own loader reads the identical floats (printed with %.9g).

Everything is numpy; nothing here touches the GPU.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field

import numpy as np

F32 = np.float32
I32 = np.int32


# --------------------------------------------------------------------------------------
# deterministic lattice noise (integer hash; independent of numpy's RNG implementation)
# --------------------------------------------------------------------------------------
def _hash01(ix: np.ndarray, iy: np.ndarray, seed: int) -> np.ndarray:
    h = (ix.astype(np.uint64) * np.uint64(0x9E3779B1)) ^ (
        iy.astype(np.uint64) * np.uint64(0x85EBCA77)) ^ np.uint64(seed * 0xC2B2AE3D & 0xFFFFFFFF)
    h &= np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    return h.astype(np.float64) / 4294967296.0


def value_noise(u: np.ndarray, v: np.ndarray, seed: int, octaves: int = 3) -> np.ndarray:
    out = np.zeros_like(u, dtype=np.float64)
    amp, freq = 1.0, 1.0
    for o in range(octaves):
        x, y = u * freq, v * freq
        x0, y0 = np.floor(x), np.floor(y)
        fx, fy = x - x0, y - y0
        sx, sy = fx * fx * (3 - 2 * fx), fy * fy * (3 - 2 * fy)
        ix, iy = x0.astype(np.int64) & 0xFFFF, y0.astype(np.int64) & 0xFFFF
        a = _hash01(ix, iy, seed + o)
        b = _hash01(ix + 1, iy, seed + o)
        c = _hash01(ix, iy + 1, seed + o)
        d = _hash01(ix + 1, iy + 1, seed + o)
        out += amp * ((a * (1 - sx) + b * sx) * (1 - sy) + (c * (1 - sx) + d * sx) * sy)
        amp *= 0.5
        freq *= 2.0
    return out


# --------------------------------------------------------------------------------------
# mesh / material builders
# --------------------------------------------------------------------------------------
class MeshBuilder:
    """Accumulates objects into the reference's SoA (0-based indices)."""

    def __init__(self):
        self.v, self.vn, self.vt = [], [], []
        self.p, self.n, self.t = [], [], []
        self.mat, self.obj = [], []
        self.obj_num, self.obj_off = [], []
        self._nv = self._nn = self._nt = self._ntri = 0

    def add(self, verts, faces, material, normals=None, uvs=None):
        verts = np.asarray(verts, dtype=F32).reshape(-1, 3)
        faces = np.asarray(faces, dtype=np.int64).reshape(-1, 3)
        k = len(faces)
        self.v.append(verts)
        self.p.append((faces + self._nv).astype(I32))
        if normals is not None:
            normals = np.asarray(normals, dtype=F32).reshape(-1, 3)
            self.vn.append(normals)
            self.n.append((faces + self._nn).astype(I32))
            self._nn += len(normals)
        else:
            # the loader leaves index 0 where a face has no normal; vn[0] must then be a
            # zero vector so Triangle falls back to the geometric normal (triangle.h:292-300)
            self.n.append(np.full((k, 3), -1, dtype=I32))
        if uvs is not None:
            uvs = np.asarray(uvs, dtype=F32).reshape(-1, 2)
            self.vt.append(uvs)
            self.t.append((faces + self._nt).astype(I32))
            self._nt += len(uvs)
        else:
            self.t.append(np.full((k, 3), -1, dtype=I32))
        m = np.asarray(material, dtype=I32)
        self.mat.append(np.broadcast_to(m, (k,)).astype(I32))
        self.obj.append(np.full(k, len(self.obj_num), dtype=I32))
        self.obj_off.append(self._ntri)
        self.obj_num.append(k)
        self._nv += len(verts)
        self._ntri += k

    def finish(self) -> dict:
        # slot 0 of vn / vt is the "absent" entry (zero normal, zero uv)
        vn = np.concatenate([np.zeros((1, 3), F32)] + self.vn) if True else None
        vt = np.concatenate([np.zeros((1, 2), F32)] + self.vt)
        v = np.concatenate(self.v)
        p = np.concatenate(self.p)
        n = np.concatenate(self.n) + 1  # shift: -1 (absent) -> 0, real -> +1
        t = np.concatenate(self.t) + 1
        c = np.ascontiguousarray
        return dict(
            x=c(v[:, 0]), y=c(v[:, 1]), z=c(v[:, 2]),
            x_norm=c(vn[:, 0]), y_norm=c(vn[:, 1]), z_norm=c(vn[:, 2]),
            x_tex=c(vt[:, 0]), y_tex=c(vt[:, 1]),
            point_1_idx=c(p[:, 0]), point_2_idx=c(p[:, 1]), point_3_idx=c(p[:, 2]),
            norm_1_idx=c(n[:, 0].astype(I32)), norm_2_idx=c(n[:, 1].astype(I32)),
            norm_3_idx=c(n[:, 2].astype(I32)),
            tex_1_idx=c(t[:, 0].astype(I32)), tex_2_idx=c(t[:, 1].astype(I32)),
            tex_3_idx=c(t[:, 2].astype(I32)),
            material_idx=c(np.concatenate(self.mat)),
            triangle_object_idx=c(np.concatenate(self.obj)),
            object_num_primitives=np.asarray(self.obj_num, dtype=I32),
            object_primitive_offset_idx=np.asarray(self.obj_off, dtype=I32),
        )


MAT_FLOAT_FIELDS = (
    "ka_x ka_y ka_z kd_x kd_y kd_z ks_x ks_y ks_z ke_x ke_y ke_z tf_x tf_y tf_z "
    "path_length t_r n_s n_i bm scattering_coef absorption_coef g").split()
MAT_INT_FIELDS = (
    "material_priority height_diffuse width_diffuse offset_diffuse height_specular "
    "width_specular offset_specular height_emission width_emission offset_emission "
    "height_n_s width_n_s offset_n_s height_bump width_bump offset_bump").split()


class MaterialBuilder:
    """Material table + planar texture atlas (offset 0 = 1x1 white default,
    src/input/read_image_util.h:135-144)."""

    def __init__(self):
        self.rows = []
        self.names = []
        self.tex = [np.ones((1, 3), F32)]
        self.tex_len = 1
        self.images = []

    def add_texture(self, img: np.ndarray):
        """img: (H, W, 3) float in [0,1], row 0 = v=0 (the loader flips rows).  Values are
        quantised to k/255 exactly as the reference's 8-bit image loader produces them
        (`1.0 * lum / 255.0`, src/input/read_image_util.h:58-66), so a scene written to
        files by write_scene_files loads back bit-identically."""
        q = np.clip(np.rint(np.asarray(img, dtype=np.float64) * 255.0), 0, 255)
        img = (q / 255.0).astype(F32)
        h, w, _ = img.shape
        off = self.tex_len
        self.tex.append(img.reshape(-1, 3))
        self.tex_len += h * w
        self.images.append((h, w, off, q.astype(np.uint8)))
        return (h, w, off)

    def add(self, name="m", ka=(0, 0, 0), kd=(0, 0, 0), ks=(0, 0, 0), ke=(0, 0, 0),
            tf=(1, 1, 1), path_length=0.0, t_r=0.0, n_s=0.0, n_i=1.0, bm=1.0,
            scattering_coef=0.0, absorption_coef=0.0, g=0.0, priority=-1,
            map_kd=None, map_ks=None, map_ke=None, map_ns=None, map_bump=None):
        d = (1, 1, 0)
        r = dict(ka_x=ka[0], ka_y=ka[1], ka_z=ka[2], kd_x=kd[0], kd_y=kd[1], kd_z=kd[2],
                 ks_x=ks[0], ks_y=ks[1], ks_z=ks[2], ke_x=ke[0], ke_y=ke[1], ke_z=ke[2],
                 tf_x=tf[0], tf_y=tf[1], tf_z=tf[2], path_length=path_length, t_r=t_r,
                 n_s=n_s, n_i=n_i, bm=bm, scattering_coef=scattering_coef,
                 absorption_coef=absorption_coef, g=g, material_priority=priority)
        for key, m in (("diffuse", map_kd), ("specular", map_ks), ("emission", map_ke),
                       ("n_s", map_ns), ("bump", map_bump)):
            h, w, o = m if m is not None else d
            r["height_" + key], r["width_" + key], r["offset_" + key] = h, w, o
        self.rows.append(r)
        self.names.append(name)
        return len(self.rows) - 1

    def finish(self) -> dict:
        out = {"num_materials": len(self.rows), "material_names": list(self.names)}
        for f in MAT_FLOAT_FIELDS:
            out[f] = np.asarray([r[f] for r in self.rows], dtype=F32)
        for f in MAT_INT_FIELDS:
            out[f] = np.asarray([r[f] for r in self.rows], dtype=I32)
        tex = np.concatenate(self.tex)
        out["tex_r"] = np.ascontiguousarray(tex[:, 0])
        out["tex_g"] = np.ascontiguousarray(tex[:, 1])
        out["tex_b"] = np.ascontiguousarray(tex[:, 2])
        out["images"] = list(self.images)  # (h, w, offset, uint8 HxWx3) for write_scene_files
        return out


# --------------------------------------------------------------------------------------
# shapes
# --------------------------------------------------------------------------------------
def terrain(n: int, seed: int = 1984, extent: float = 10.0, height: float = 3.0):
    """n x n vertex grid on [-extent, extent]^2 (x, z), y = value noise; 2 triangles per
    cell, vertex normals and uvs.  2 (n-1)^2 triangles."""
    g = np.linspace(-extent, extent, n, dtype=np.float64)
    xx, zz = np.meshgrid(g, g, indexing="xy")
    u = (xx + extent) / (2 * extent)
    v = (zz + extent) / (2 * extent)
    yy = height * (value_noise(u * 4.0, v * 4.0, seed) / 1.75 - 0.5)
    verts = np.stack([xx, yy, zz], axis=-1).reshape(-1, 3).astype(F32)
    # normals from central differences of the (float64) height field
    dydx = np.gradient(yy, g, axis=1)
    dydz = np.gradient(yy, g, axis=0)
    nrm = np.stack([-dydx, np.ones_like(yy), -dydz], axis=-1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    normals = nrm.reshape(-1, 3).astype(F32)
    uvs = np.stack([u, v], axis=-1).reshape(-1, 2).astype(F32)
    i, j = np.meshgrid(np.arange(n - 1), np.arange(n - 1), indexing="xy")
    a = (j * n + i).reshape(-1)
    b, c, d = a + 1, a + n, a + n + 1
    # counter-clockwise seen from +y
    faces = np.stack([np.stack([a, c, b], -1), np.stack([b, c, d], -1)], 1).reshape(-1, 3)
    return verts, faces, normals, uvs


def quad(p0, p1, p2, p3):
    verts = np.asarray([p0, p1, p2, p3], dtype=F32)
    faces = np.asarray([[0, 1, 2], [0, 2, 3]])
    uvs = np.asarray([[0, 0], [1, 0], [1, 1], [0, 1]], dtype=F32)
    return verts, faces, uvs


def icosphere(center, radius, subdiv=1):
    t = (1.0 + 5 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                  [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]],
                 dtype=np.float64)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9],
                  [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2],
                  [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]])
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    for _ in range(subdiv):
        cache, vl, nf = {}, list(map(tuple, v)), []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                m = (np.array(vl[a]) + np.array(vl[b])) / 2
                m /= np.linalg.norm(m)
                vl.append(tuple(m))
                cache[k] = len(vl) - 1
            return cache[k]

        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        v, f = np.array(vl), np.array(nf)
    verts = (v * radius + np.asarray(center, dtype=np.float64)).astype(F32)
    return verts, f, v.astype(F32)


def box(lo, hi, inward=False):
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    c = np.array([[lo[0], lo[1], lo[2]], [hi[0], lo[1], lo[2]], [hi[0], hi[1], lo[2]],
                  [lo[0], hi[1], lo[2]], [lo[0], lo[1], hi[2]], [hi[0], lo[1], hi[2]],
                  [hi[0], hi[1], hi[2]], [lo[0], hi[1], hi[2]]], dtype=F32)
    f = np.array([[0, 2, 1], [0, 3, 2], [4, 5, 6], [4, 6, 7], [0, 1, 5], [0, 5, 4],
                  [3, 7, 6], [3, 6, 2], [0, 4, 7], [0, 7, 3], [1, 2, 6], [1, 6, 5]])
    if inward:
        f = f[:, ::-1]
    return c, f


def grid_quad(p0, du, dv, nu, nv):
    """Tessellated parallelogram p0 + s du + t dv with nu x nv cells."""
    p0, du, dv = (np.asarray(a, np.float64) for a in (p0, du, dv))
    s, t = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    verts = (p0 + s[..., None] * du + t[..., None] * dv).reshape(-1, 3).astype(F32)
    uvs = np.stack([s, t], -1).reshape(-1, 2).astype(F32)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="xy")
    a = (j * (nu + 1) + i).reshape(-1)
    b, c, d = a + 1, a + nu + 1, a + nu + 2
    faces = np.stack([np.stack([a, b, c], -1), np.stack([b, d, c], -1)], 1).reshape(-1, 3)
    return verts, faces, uvs


# --------------------------------------------------------------------------------------
# scenes
# --------------------------------------------------------------------------------------
@dataclass
class Scene:
    """Everything that crosses the C-ABI, as numpy arrays / plain scalars."""
    mesh: dict
    materials: dict
    camera: dict
    params: dict
    bg: dict = field(default_factory=lambda: dict(
        r=np.ones(1, F32), g=np.ones(1, F32), b=np.ones(1, F32), height=1, width=1))
    name: str = "scene"
    files: dict = field(default_factory=dict)  # paths + parse timings when loaded from files

    @property
    def num_triangles(self) -> int:
        return int(self.mesh["point_1_idx"].shape[0])

    def pinned(self) -> "Scene":
        """The same scene with every array in page-locked host memory, so that the uploads of
        gpt_set_mesh / gpt_set_textures / gpt_set_background are real DMA transfers instead of
        the driver's staged pageable copies (C2: 86 MB in 8.4 ms pageable).  The reference
        keeps these arrays in cudaMallocManaged memory (src/main.cu:386-450); a host that
        fills them for this library can allocate them page-locked the same way."""
        import torch

        def pin(v):
            if not isinstance(v, np.ndarray) or v.size == 0:
                return v
            return torch.from_numpy(np.ascontiguousarray(v)).pin_memory().numpy()

        return Scene(mesh={k: pin(v) for k, v in self.mesh.items()},
                     materials={k: pin(v) for k, v in self.materials.items()},
                     camera=dict(self.camera), params=dict(self.params),
                     bg={k: pin(v) for k, v in self.bg.items()}, name=self.name,
                     files=dict(self.files))


def default_params(**kw) -> dict:
    p = dict(render_mode=1, pathtracing_sample_size=16, pathtracing_level=5,
             dof_sample_size=1, sss_pts_per_object=0, hittable_pdf_weight=0.5,
             sky_emission=(0.5, 0.5, 0.5), ppm_num_photon_per_pass=100000, ppm_num_pass=4,
             ppm_max_bounce=8, ppm_alpha=0.7, ppm_radius_scaling_factor=2.0, rng_mode=0)
    p.update(kw)
    return p


def default_camera(width, height, **kw) -> dict:
    c = dict(eye=(0.0, 7.0, 16.0), center=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), fovy=45.0,
             width=width, height=height, aperture=0.0, focus_dist=10.0)
    c.update(kw)
    return c


def _add_lights(mb: MeshBuilder, mat_light: int, y=8.0, size=2.0, sep=4.0):
    # two emissive quads (4 triangles) facing down: the reference needs >= 2 emissive
    # triangles (src/main.cu:1124-1125)
    for cx in (-sep, sep):
        v, f, uv = quad((cx - size, y, -size), (cx + size, y, -size),
                        (cx + size, y, size), (cx - size, y, size))
        mb.add(v, f, mat_light, uvs=uv)


def terrain_scene(n: int = 225, width: int = 512, height: int = 512, spp: int = 16,
                  level: int = 5, seed: int = 1984, name: str = "C1") -> Scene:
    """Config 1: single diffuse terrain + floor + two area lights."""
    mats = MaterialBuilder()
    m_terrain = mats.add("terrain", kd=(0.7, 0.7, 0.7))
    m_floor = mats.add("floor", kd=(0.5, 0.5, 0.5))
    m_light = mats.add("light", kd=(0.0, 0.0, 0.0), ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    v, f, nrm, uv = terrain(n, seed)
    mb.add(v, f, m_terrain, normals=nrm, uvs=uv)
    v, f, uv = quad((-12, -2.0, -12), (-12, -2.0, 12), (12, -2.0, 12), (12, -2.0, -12))
    mb.add(v, f, m_floor, uvs=uv)
    _add_lights(mb, m_light)
    return Scene(mb.finish(), mats.finish(), default_camera(width, height),
                 default_params(pathtracing_sample_size=spp, pathtracing_level=level),
                 name=name)


def checker_texture(size=1024, cells=32):
    i, j = np.meshgrid(np.arange(size), np.arange(size), indexing="ij")
    c = (((i * cells) // size + (j * cells) // size) & 1).astype(F32)
    img = np.stack([0.25 + 0.7 * c, 0.3 + 0.5 * c, 0.85 - 0.6 * c], -1)
    return img.astype(F32)


def bump_texture(size=1024, seed=7):
    g = np.linspace(0, 16, size, dtype=np.float64)
    u, v = np.meshgrid(g, g, indexing="xy")
    a = value_noise(u, v, seed, 2) / 1.5
    b = value_noise(u + 37.0, v + 11.0, seed + 5, 2) / 1.5
    return np.stack([a, b, np.full_like(a, 0.5)], -1).astype(F32)


def mixed_scene(n: int = 708, width: int = 1920, height: int = 1080, spp: int = 64,
                level: int = 5, tex: int = 1024, n_spheres: int = 64, seed: int = 1984,
                name: str = "C2") -> Scene:
    """Config 2: terrain in three material bands (textured+bump diffuse, metal, plain),
    plus dielectric icospheres resting on it."""
    mats = MaterialBuilder()
    t_kd = mats.add_texture(checker_texture(tex))
    t_bm = mats.add_texture(bump_texture(tex))
    m_tex = mats.add("checker", kd=(0.9, 0.9, 0.9), map_kd=t_kd, map_bump=t_bm, bm=0.3)
    m_metal = mats.add("metal", kd=(0.05, 0.05, 0.05), ks=(0.9, 0.9, 0.9), n_s=500.0)
    m_plain = mats.add("plain", kd=(0.7, 0.6, 0.5))
    m_glass = mats.add("glass", tf=(1, 1, 1), t_r=1.0, n_i=1.5, n_s=1e6, priority=1)
    m_floor = mats.add("floor", kd=(0.5, 0.5, 0.5))
    m_light = mats.add("light", ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    v, f, nrm, uv = terrain(n, seed)
    # band by the x coordinate of the first vertex of each face
    cx = v[f[:, 0], 0]
    band = np.where(cx < -10.0 / 3, m_tex, np.where(cx < 10.0 / 3, m_metal, m_plain))
    mb.add(v, f, band.astype(I32), normals=nrm, uvs=uv)
    v2, f2, uv2 = quad((-12, -2.0, -12), (-12, -2.0, 12), (12, -2.0, 12), (12, -2.0, -12))
    mb.add(v2, f2, m_floor, uvs=uv2)
    _add_lights(mb, m_light)
    # spheres on an 8 x 8 lattice, resting 0.02 above the local terrain height
    k = int(round(n_spheres ** 0.5))
    for a in range(k):
        for b in range(k):
            sx = -8.0 + 16.0 * (a + 0.5) / k
            sz = -8.0 + 16.0 * (b + 0.5) / k
            gi = int(round((sx + 10.0) / 20.0 * (n - 1)))
            gj = int(round((sz + 10.0) / 20.0 * (n - 1)))
            sy = float(v[gj * n + gi, 1]) + 0.6 + 0.02
            sv, sf, sn = icosphere((sx, sy, sz), 0.6, 2)
            mb.add(sv, sf, m_glass, normals=sn)
    return Scene(mb.finish(), mats.finish(), default_camera(width, height),
                 default_params(pathtracing_sample_size=spp, pathtracing_level=level),
                 name=name)


def bottle_scene(width=1920, height=1080, spp=8, dof=8, level=8, name="C3") -> Scene:
    """Config 3: nested dielectrics (glass shell > liquid > diffuse object) + depth of field."""
    mats = MaterialBuilder()
    m_floor = mats.add("floor", kd=(0.6, 0.6, 0.6))
    m_glass = mats.add("glass", t_r=1.0, n_i=1.5, n_s=1e6, priority=1)
    m_liquid = mats.add("liquid", tf=(0.8, 0.9, 1.0), t_r=1.0, n_i=1.33, n_s=1e6, priority=2)
    m_obj = mats.add("ship", kd=(0.8, 0.3, 0.2), ks=(0.2, 0.2, 0.2), n_s=50.0)
    m_light = mats.add("light", ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    v, f, uv = grid_quad((-12, 0, -12), (0, 0, 24), (24, 0, 0), 64, 64)
    mb.add(v, f, m_floor, uvs=uv)
    sv, sf, sn = icosphere((0, 2.0, 0), 2.0, 4)
    mb.add(sv, sf, m_glass, normals=sn)
    sv, sf, sn = icosphere((0, 2.0, 0), 1.8, 4)
    mb.add(sv, sf, m_liquid, normals=sn)
    sv, sf, sn = icosphere((0, 1.8, 0), 0.8, 3)
    mb.add(sv, sf, m_obj, normals=sn)
    _add_lights(mb, m_light)
    cam = default_camera(width, height, eye=(0.0, 3.5, 9.0), center=(0.0, 2.0, 0.0),
                         aperture=0.3, focus_dist=9.1)
    return Scene(mb.finish(), mats.finish(), cam,
                 default_params(pathtracing_sample_size=spp, dof_sample_size=dof,
                                pathtracing_level=level), name=name)


def sky_texture(height: int = 32, width: int = 64, seed: int = 5) -> dict:
    """A non-constant equirectangular background (horizon glow + value noise), quantised to
    k/255 like an 8-bit image the reference's loader would have read."""
    v, u = np.meshgrid((np.arange(height) + 0.5) / height, (np.arange(width) + 0.5) / width,
                       indexing="ij")
    n = value_noise(u * 6, v * 3, seed)
    img = np.stack([0.35 + 0.6 * np.exp(-((v - 0.5) * 4) ** 2) * (0.6 + 0.4 * n),
                    0.45 + 0.4 * n, 0.95 - 0.5 * v], -1)
    img = (np.clip(np.rint(img * 255.0), 0, 255) / 255.0).astype(F32)
    return dict(r=np.ascontiguousarray(img[..., 0].ravel()), g=np.ascontiguousarray(img[..., 1].ravel()),
                b=np.ascontiguousarray(img[..., 2].ravel()), height=height, width=width)


def small_scene(n: int = 24, width: int = 128, height: int = 128, spp: int = 2,
                level: int = 4, glass: bool = True, metal: bool = True,
                textured: bool = False, sss_pts: int = 0, all_maps: bool = False,
                sky: bool = False, name: str = "small") -> Scene:
    """Small everything-in-one scene for parity tests that the host replay finishes fast.
    all_maps: the metal sphere gets a specular-colour map (map_Ks) and an exponent map (map_Ns),
    the lights an emission map (map_Ke) -- Material::_get_texture for every slot
    (material.h:653-727).  sky: a 64 x 32 non-constant equirectangular background
    (_get_sky_color, pathtracing.h:69-82)."""
    mats = MaterialBuilder()
    kw = {}
    if textured:
        kw = dict(map_kd=mats.add_texture(checker_texture(64, 8)),
                  map_bump=mats.add_texture(bump_texture(64)), bm=0.3)
    m_terrain = mats.add("terrain", kd=(0.7, 0.7, 0.7), **kw)
    mkw, lkw = {}, {}
    if all_maps:
        ks_img = 0.35 + 0.65 * checker_texture(32, 4)
        ks_img[..., 2] *= 0.6
        ns_img = np.repeat(value_noise(*np.meshgrid(np.linspace(0, 4, 32), np.linspace(0, 4, 32)), 3)[..., None], 3, -1)
        mkw = dict(map_ks=mats.add_texture(ks_img), map_ns=mats.add_texture(0.2 + 0.8 * ns_img))
        ke_img = 0.5 + 0.5 * checker_texture(16, 2)
        ke_img[..., 0] *= 0.8
        lkw = dict(map_ke=mats.add_texture(ke_img))
    m_metal = mats.add("metal", kd=(0.05, 0.05, 0.05), ks=(0.9, 0.9, 0.9), n_s=200.0, **mkw)
    m_glass = mats.add("glass", t_r=1.0, n_i=1.5, n_s=1e6, priority=1)
    m_liquid = mats.add("liquid", tf=(0.8, 0.9, 1.0), t_r=1.0, n_i=1.33, n_s=1e6, priority=2)
    m_light = mats.add("light", ke=(15.0, 15.0, 15.0), **lkw)
    mb = MeshBuilder()
    v, f, nrm, uv = terrain(n, 1984)
    mb.add(v, f, m_terrain, normals=nrm, uvs=uv)
    if metal:
        sv, sf, sn = icosphere((-3.0, 2.0, 0.0), 1.5, 2)
        suv = None
        if all_maps:  # spherical uv so that the specular / exponent maps vary over the sphere
            d = sn / np.linalg.norm(sn, axis=1, keepdims=True)
            suv = np.stack([0.5 + np.arctan2(d[:, 2], d[:, 0]) / (2 * np.pi),
                            0.5 + np.arcsin(np.clip(d[:, 1], -1, 1)) / np.pi], 1).astype(F32)
        mb.add(sv, sf, m_metal, normals=sn, uvs=suv)
    if glass:
        sv, sf, sn = icosphere((3.0, 2.0, 1.0), 1.6, 2)
        mb.add(sv, sf, m_glass, normals=sn)
        sv, sf, sn = icosphere((3.0, 2.0, 1.0), 1.2, 2)
        mb.add(sv, sf, m_liquid, normals=sn)
    if sss_pts:
        # two sub-surface-scattering objects (path_length > 0): point clouds per OBJECT
        m_wax = mats.add("wax", kd=(0.8, 0.55, 0.4), path_length=0.6)
        sv, sf, sn = icosphere((0.0, 3.0, 3.0), 1.2, 2)
        mb.add(sv, sf, m_wax, normals=sn)
        sv, sf, sn = icosphere((-4.0, 3.5, 3.5), 0.8, 1)
        mb.add(sv, sf, m_wax, normals=sn)
    _add_lights(mb, m_light)
    sc = Scene(mb.finish(), mats.finish(), default_camera(width, height),
               default_params(pathtracing_sample_size=spp, pathtracing_level=level,
                              sss_pts_per_object=sss_pts),
               name=name)
    if sky:
        sc.bg = sky_texture()
    return sc


def nested_shells_scene(shells: int = 14, width: int = 128, height: int = 128) -> Scene:
    """`shells` concentric glass spheres of distinct materials whose priority number grows
    inwards: a ray entering them piles every one onto its nested-dielectric list
    (add_new_material, material_list_operations.h:9-24)."""
    mats = MaterialBuilder()
    m_floor = mats.add("floor", kd=(0.6, 0.6, 0.6))
    m_light = mats.add("light", ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    v, f, uv = quad((-12, -2.0, -12), (-12, -2.0, 12), (12, -2.0, 12), (12, -2.0, -12))
    mb.add(v, f, m_floor, uvs=uv)
    for k in range(shells):
        m = mats.add(f"shell{k}", tf=(1, 1, 1), t_r=1.0, n_i=1.0 + 0.02 * (k + 1), n_s=1e6,
                     priority=k + 1)
        sv, sf, sn = icosphere((0.0, 2.0, 0.0), 3.0 - 0.15 * k, 2)
        mb.add(sv, sf, m, normals=sn)
    _add_lights(mb, m_light)
    return Scene(mb.finish(), mats.finish(), default_camera(width, height),
                 default_params(pathtracing_sample_size=1, pathtracing_level=40),
                 name="shells")


def sppm_scene(cells: int = 91, width: int = 1920, height: int = 1080, photons: int = 10_000_000,
               passes: int = 16, fog: bool = True, sphere_subdiv: int = 4, room: float = 10.0,
               name: str = "C4") -> Scene:
    """Config 4: closed diffuse room (6 tessellated walls, inward normals) + a glass sphere +
    a fog box (participating medium) + two area lights under the ceiling; render_mode 2."""
    L = float(room)
    mats = MaterialBuilder()
    m_white = mats.add("white", kd=(0.75, 0.75, 0.75))
    m_red = mats.add("red", kd=(0.75, 0.25, 0.25))
    m_green = mats.add("green", kd=(0.25, 0.75, 0.25))
    m_glass = mats.add("glass", t_r=1.0, n_i=1.5, n_s=1e6, priority=1)
    m_fog = mats.add("fog", t_r=1.0, n_i=1.0, n_s=1e6, priority=3, scattering_coef=0.05,
                     absorption_coef=0.01, g=0.3)
    # the emitter needs a diffuse Kd: with kd = ks = 0 the reference's diffuse/specular split
    # is 0/0 and the light never counts as a diffuse emitter in _compute_direct_radiance
    m_light = mats.add("light", kd=(0.78, 0.78, 0.78), ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    walls = [((-L, -L, -L), (0, 0, 2 * L), (2 * L, 0, 0), m_white),   # floor    (+y)
             ((-L, L, -L), (2 * L, 0, 0), (0, 0, 2 * L), m_white),    # ceiling  (-y)
             ((-L, -L, -L), (2 * L, 0, 0), (0, 2 * L, 0), m_white),   # back     (+z)
             ((-L, -L, L), (0, 2 * L, 0), (2 * L, 0, 0), m_white),    # front    (-z)
             ((-L, -L, -L), (0, 2 * L, 0), (0, 0, 2 * L), m_red),     # left     (+x)
             ((L, -L, -L), (0, 0, 2 * L), (0, 2 * L, 0), m_green)]    # right    (-x)
    for p0, du, dv, m in walls:
        v, f, uv = grid_quad(p0, du, dv, cells, cells)
        mb.add(v, f, m, uvs=uv)
    sv, sf, sn = icosphere((-3.0, -L + 3.0, -1.0), 3.0, sphere_subdiv)
    mb.add(sv, sf, m_glass, normals=sn)
    if fog:
        bv, bf = box((-0.5 * L, -L + 0.5, -0.6 * L), (0.8 * L, 0.4 * L, 0.5 * L))
        mb.add(bv, bf, m_fog)
    _add_lights(mb, m_light, y=L - 0.5, size=0.2 * L, sep=0.4 * L)
    cam = default_camera(width, height, eye=(0.0, 0.0, L - 1.0), center=(0.0, -0.2 * L, 0.0),
                         fovy=60.0)
    par = default_params(render_mode=2, ppm_num_photon_per_pass=photons, ppm_num_pass=passes,
                         ppm_max_bounce=8, ppm_alpha=0.7, ppm_radius_scaling_factor=2.0,
                         pathtracing_sample_size=1, pathtracing_level=8)
    return Scene(mb.finish(), mats.finish(), cam, par, name=name)


def big_scene(n: int = 2237, width: int = 3840, height: int = 2160, spp: int = 256,
              level: int = 5, sss_pts: int = 100000, name: str = "C5") -> Scene:
    """Config 5: terrain-10M (n = 2237 -> 9 999 392 triangles) + a fog box + one sub-surface
    scattering object (path_length .3, sss_pts_per_object 1e5), 4K.  The BVH (1.1 GB) no
    longer fits the L2, so traversal is HBM-bound here."""
    mats = MaterialBuilder()
    m_terrain = mats.add("terrain", kd=(0.7, 0.65, 0.55))
    m_floor = mats.add("floor", kd=(0.5, 0.5, 0.5))
    m_fog = mats.add("fog", t_r=1.0, n_i=1.0, n_s=1e6, priority=3, scattering_coef=0.05,
                     absorption_coef=0.01, g=0.3)
    m_wax = mats.add("wax", kd=(0.8, 0.55, 0.4), path_length=0.3)
    m_light = mats.add("light", ke=(15.0, 15.0, 15.0))
    mb = MeshBuilder()
    v, f, nrm, uv = terrain(n, 1984)
    mb.add(v, f, m_terrain, normals=nrm, uvs=uv)
    v2, f2, uv2 = quad((-12, -2.0, -12), (-12, -2.0, 12), (12, -2.0, 12), (12, -2.0, -12))
    mb.add(v2, f2, m_floor, uvs=uv2)
    bv, bf = box((-6.0, 1.5, -6.0), (5.0, 3.0, 5.0))
    mb.add(bv, bf, m_fog)
    sv, sf, sn = icosphere((3.0, 3.2, 1.0), 1.4, 4)
    mb.add(sv, sf, m_wax, normals=sn)
    _add_lights(mb, m_light)
    return Scene(mb.finish(), mats.finish(), default_camera(width, height),
                 default_params(pathtracing_sample_size=spp, pathtracing_level=level,
                                sss_pts_per_object=sss_pts), name=name)


def make_scene(kind: str, **kw) -> Scene:
    if kind in ("C5", "big"):
        return big_scene(**kw)
    if kind in ("C4", "sppm"):
        return sppm_scene(**kw)
    if kind in ("C1", "terrain"):
        return terrain_scene(**kw)
    if kind in ("C2", "mixed"):
        return mixed_scene(**kw)
    if kind in ("C3", "bottle"):
        return bottle_scene(**kw)
    if kind == "small":
        return small_scene(**kw)
    if kind == "shells":
        return nested_shells_scene(**kw)
    raise ValueError(kind)


# --------------------------------------------------------------------------------------
# file writer: the same scene as OBJ + MTL + raw textures + master file, for the
# reference's own loaders (src/input/read_file_util.h, read_image_util.h, input_param.h)
# --------------------------------------------------------------------------------------
def _write_image(path: str, rgb_u8: np.ndarray):
    """Raw image understood by oracle/jpeg_stub/jpeg.h.  The loader flips rows (it reads
    y = H-1 .. 0 into increasing atlas rows), so atlas row r is stored as file row H-1-r."""
    h, w, _ = rgb_u8.shape
    with open(path, "wb") as f:
        f.write(b"GPTIMG %d %d\n" % (w, h))
        f.write(np.ascontiguousarray(rgb_u8[::-1]).tobytes())


def write_scene_files(sc: Scene, folder: str, prefix: str = "scene") -> str:
    """Writes <folder>/{prefix}.obj/.mtl, textures, bg image and master file; returns the
    master file path.  Floats are printed with %.9g (exact fp32 round trip through stof).
    Loader conventions honoured: material 0 of the loader is its built-in Default_123
    (indices shift by one, values identical); vn/vt slot 0 are the explicit zero entries the
    SoA uses for "absent"; one `usemtl` per run of equal material (faces keep their order)."""
    os.makedirs(folder, exist_ok=True)
    folder = os.path.join(os.path.abspath(folder), "")
    m, mt = sc.mesh, sc.materials
    names = [f"m{i}_{n}" for i, n in enumerate(mt["material_names"])]
    tex_name = {0: None}
    for (h, w, off, img) in mt["images"]:
        tex_name[off] = f"{prefix}_tex{off}.img"
        _write_image(folder + tex_name[off], img)
    _write_image(folder + f"{prefix}_bg.img", np.full((1, 1, 3), 255, np.uint8))
    g9 = lambda v: "%.9g" % float(v)
    with open(folder + prefix + ".mtl", "w") as f:
        for i, n in enumerate(names):
            f.write(f"newmtl {n}\n")
            for key, col in (("Ka", "ka"), ("Kd", "kd"), ("Ks", "ks"), ("Ke", "ke"), ("Tf", "tf")):
                f.write(f"{key} " + " ".join(g9(mt[f"{col}_{a}"][i]) for a in "xyz") + "\n")
            f.write(f"Tr {g9(mt['t_r'][i])}\nNs {g9(mt['n_s'][i])}\nNi {g9(mt['n_i'][i])}\n")
            f.write(f"path_length {g9(mt['path_length'][i])}\n")
            f.write(f"scattering_coef {g9(mt['scattering_coef'][i])}\n")
            f.write(f"absorption_coef {g9(mt['absorption_coef'][i])}\ng {g9(mt['g'][i])}\n")
            if mt["material_priority"][i] != -1:
                f.write(f"priority {int(mt['material_priority'][i])}\n")
            for key, col in (("map_Kd", "diffuse"), ("map_Ks", "specular"), ("map_Ke", "emission"),
                             ("map_Ns", "n_s")):
                off = int(mt["offset_" + col][i])
                if off:
                    f.write(f"{key} {tex_name[off]}\n")
            off = int(mt["offset_bump"][i])
            if off:
                f.write(f"map_Bump -bm {g9(mt['bm'][i])} {tex_name[off]}\n")
            f.write("\n")
    with open(folder + prefix + ".obj", "w") as f:
        f.write(f"mtllib {prefix}.mtl\n")
        f.write("".join("v %.9g %.9g %.9g\n" % t for t in zip(m["x"], m["y"], m["z"])))
        f.write("".join("vn %.9g %.9g %.9g\n" % t
                        for t in zip(m["x_norm"], m["y_norm"], m["z_norm"])))
        f.write("".join("vt %.9g %.9g\n" % t for t in zip(m["x_tex"], m["y_tex"])))
        P = np.stack([m["point_1_idx"], m["point_2_idx"], m["point_3_idx"]], 1) + 1
        T = np.stack([m["tex_1_idx"], m["tex_2_idx"], m["tex_3_idx"]], 1) + 1
        N = np.stack([m["norm_1_idx"], m["norm_2_idx"], m["norm_3_idx"]], 1) + 1
        mat = m["material_idx"]
        lines, cur = [], None
        for i in range(P.shape[0]):
            if mat[i] != cur:
                cur = mat[i]
                lines.append(f"usemtl {names[cur]}\n")
            lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (
                P[i, 0], T[i, 0], N[i, 0], P[i, 1], T[i, 1], N[i, 1], P[i, 2], T[i, 2], N[i, 2]))
        f.write("".join(lines))
    c, p = sc.camera, sc.params
    kv = [("input_folder_path", folder), ("obj_filename", prefix + ".obj"),
          ("texture_bg_path", f"{prefix}_bg.img"), ("image_width", c["width"]),
          ("image_height", c["height"]), ("render_mode", p["render_mode"]),
          ("pathtracing_sample_size", p["pathtracing_sample_size"]),
          ("pathtracing_level", p["pathtracing_level"]),
          ("eye_x", g9(c["eye"][0])), ("eye_y", g9(c["eye"][1])), ("eye_z", g9(c["eye"][2])),
          ("center_x", g9(c["center"][0])), ("center_y", g9(c["center"][1])),
          ("center_z", g9(c["center"][2])), ("up_x", g9(c["up"][0])), ("up_y", g9(c["up"][1])),
          ("up_z", g9(c["up"][2])), ("fovy", g9(c["fovy"])), ("aperture", g9(c["aperture"])),
          ("dof_sample_size", p["dof_sample_size"]), ("focus_dist", g9(c["focus_dist"])),
          ("sky_emission_r", g9(p["sky_emission"][0])), ("sky_emission_g", g9(p["sky_emission"][1])),
          ("sky_emission_b", g9(p["sky_emission"][2])),
          ("sss_pts_per_object", max(1, p["sss_pts_per_object"])),
          ("hittable_pdf_weight", g9(p["hittable_pdf_weight"])),
          ("ppm_num_photon_per_pass", p["ppm_num_photon_per_pass"]),
          ("ppm_num_pass", p["ppm_num_pass"]), ("ppm_max_bounce", p["ppm_max_bounce"]),
          ("ppm_alpha", g9(p["ppm_alpha"])),
          ("ppm_radius_scaling_factor", g9(p["ppm_radius_scaling_factor"])),
          ("ppm_image_output_iteration", 1000000)]
    master = folder + prefix + "_master.txt"
    with open(master, "w") as f:
        for k, v in kv:
            f.write(f"{k}\t{v}\n")
    return master
