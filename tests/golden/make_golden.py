"""Generates tests/golden/*.npz from the host replay of the reference's OWN headers
(oracle/_ref, built from /root/reference/src).  Run in the build container (the reference
is not on the GPU box); the fixtures are committed so every test machine can check against
them without the reference sources.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from gpu_pathtracer_b200 import rays, scene as S  # noqa: E402
from tests.refreplay import RefReplay  # noqa: E402


def golden_scene():
    """The fixture scene: terrain + metal + nested glass/liquid + lights, textured."""
    return S.make_scene("small", n=14, width=128, height=128, spp=2, level=5, textured=True)


def golden_rays(sc):
    return np.concatenate([rays.camera_rays(sc)[::11], rays.random_rays(1200, seed=5),
                           rays.edge_case_rays(sc, 800, seed=9)]).astype(np.float32)


WINDOW = (48, 80, 40, 88)  # rows, cols of the rendered fixture window


def main():
    sc = golden_scene()
    ref = RefReplay(sc)
    out = {}
    for k, v in ref.export_triangles().items():
        out["tri_" + k] = v
    for which in (0, 1, 2):
        for k, v in ref.export_bvh(which).items():
            out[f"bvh{which}_{k}"] = v
    r = golden_rays(sc)
    out["rays"] = r
    for which in (0, 1):
        for k, v in ref.trace(r, which).items():
            if k != "ms":
                out[f"trace{which}_{k}"] = v
    img = ref.render_pathtrace(window=WINDOW, nthreads=1)
    r0, r1, c0, c1 = WINDOW
    out["render_window"] = img[r0:r1, c0:c1]
    out["render_rays"] = np.asarray([ref.render_rays], np.int64)
    path = os.path.join(HERE, "small_scene.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
