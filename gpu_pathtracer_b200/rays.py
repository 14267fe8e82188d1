"""Deterministic ray sets for parity tests and the trace bench (R1-R3).

Rays are (n, 8) float32 rows {ox, oy, oz, tmax, dx, dy, dz, 0}; directions need not be unit
length (gpt_trace normalises them exactly as Ray::Ray does, src/model/ray/ray.h:21-24).
"""
from __future__ import annotations

import numpy as np

from .scene import Scene

F32 = np.float32


def pack(o, d, tmax=np.inf) -> np.ndarray:
    o = np.asarray(o, F32).reshape(-1, 3)
    d = np.asarray(d, F32).reshape(-1, 3)
    r = np.zeros((o.shape[0], 8), F32)
    r[:, 0:3] = o
    r[:, 3] = tmax
    r[:, 4:7] = d
    return r


def camera_rays(scene: Scene) -> np.ndarray:
    """R1 for aperture 0: pinhole rays through pixel centres, written with the same fp32
    operation order as Camera::compute_ray (src/model/camera.h:68-83) except for tan(),
    which is evaluated in float64 here -- use only where bit-exact ray ORIGINATION is not
    under test (both sides of a trace comparison consume the same ray array)."""
    c = scene.camera
    W, H = c["width"], c["height"]
    eye, center, up = (np.asarray(c[k], np.float64) for k in ("eye", "center", "up"))
    w = eye - center
    w /= np.linalg.norm(w)
    u = np.cross(up, w)
    u /= np.linalg.norm(u)
    v = np.cross(w, u)
    fovy = float(c["fovy"])
    fovx = 2.0 * np.degrees(np.arctan(W / H * np.tan(np.radians(fovy) / 2)))
    jj, ii = np.meshgrid(np.arange(W) + 0.5, np.arange(H) + 0.5, indexing="xy")
    a = np.tan(np.radians(fovx) / 2) * (jj - W / 2) / (W / 2)
    b = np.tan(np.radians(fovy) / 2) * (H / 2 - ii) / (H / 2)
    d = a[..., None] * u + b[..., None] * v - w
    o = np.broadcast_to(eye, d.shape)
    return pack(o.reshape(-1, 3), d.reshape(-1, 3))


def random_rays(n: int, seed: int = 7, lo=(-10, -1.5, -10), hi=(10, 6, 10)) -> np.ndarray:
    """R2-like incoherent rays: uniform origins in a box, uniform directions."""
    rng = np.random.Generator(np.random.PCG64(seed))
    o = rng.uniform(lo, hi, size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return pack(o, d)


def bounce_rays(hit_points: np.ndarray, normals: np.ndarray, seed: int = 7) -> np.ndarray:
    """R2: cosine-hemisphere directions around `normals` from `hit_points`."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = hit_points.shape[0]
    r1, r2 = rng.uniform(size=n), rng.uniform(size=n)
    phi = 2 * np.pi * r1
    loc = np.stack([np.cos(phi) * np.sqrt(r2), np.sin(phi) * np.sqrt(r2), np.sqrt(1 - r2)], 1)
    nz = normals / np.linalg.norm(normals, axis=1, keepdims=True)
    a = np.where(np.abs(nz[:, :1]) > 0.9, [[0.0, 1.0, 0.0]], [[1.0, 0.0, 0.0]])
    tx = np.cross(a, nz)
    tx /= np.linalg.norm(tx, axis=1, keepdims=True)
    ty = np.cross(nz, tx)
    d = loc[:, :1] * tx + loc[:, 1:2] * ty + loc[:, 2:3] * nz
    return pack(hit_points + 1e-3 * nz, d)


def edge_case_rays(scene: Scene, n: int = 4096, seed: int = 3) -> np.ndarray:
    """R3: axis-aligned rays (zero direction components -> inf / NaN slabs) and rays aimed
    exactly at mesh vertices and edge midpoints (equal-t ties on shared edges)."""
    m = scene.mesh
    rng = np.random.Generator(np.random.PCG64(seed))
    nv = m["x"].shape[0]
    V = np.stack([m["x"], m["y"], m["z"]], 1).astype(np.float64)
    out = []
    k = n // 4
    # (a) straight down / up / along x / along z through vertices (hits shared vertices)
    vi = rng.integers(0, nv, size=k)
    axes = np.array([[0, -1, 0], [0, 1, 0], [1, 0, 0], [0, 0, -1], [-1, 0, 0], [0, 0, 1]], float)
    ax = axes[rng.integers(0, len(axes), size=k)]
    out.append(pack(V[vi] - 5.0 * ax, ax))
    # (b) axis-aligned through random points (generic position, zero components)
    o = rng.uniform((-9, 5, -9), (9, 7, 9), size=(k, 3))
    out.append(pack(o, np.tile([0.0, -1.0, 0.0], (k, 1))))
    # (c) from a common origin exactly through vertices
    eye = np.array([0.3, 6.0, 0.2])
    vi = rng.integers(0, nv, size=k)
    out.append(pack(np.tile(eye, (k, 1)), V[vi] - eye))
    # (d) through midpoints of triangle edges (shared between two triangles)
    nt = m["point_1_idx"].shape[0]
    ti = rng.integers(0, nt, size=n - 3 * k)
    a, b = V[m["point_1_idx"][ti]], V[m["point_2_idx"][ti]]
    mid = 0.5 * (a + b)
    out.append(pack(np.tile(eye, (len(ti), 1)), mid - eye))
    return np.concatenate(out)
