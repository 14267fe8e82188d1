"""CPU: the C-ABI library loads and exports every symbol include/gpt.h declares (no compute
calls without a GPU), the ctypes mirrors match the C struct layouts, and the host-side scene /
ray generators are deterministic."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

from gpu_pathtracer_b200 import api, cabi, rays, scene as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gpt_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = api.load_library()
    decl = _declared_symbols()
    assert len(decl) >= 25
    missing = [s for s in decl if not hasattr(lib, s)]
    assert not missing, missing
    assert sorted(api.SYMBOLS) == decl
    assert b"sm_100a" in lib.gpt_version()


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(api.GptError, match="no CUDA device|no CPU"):
        api.PathTracer(S.make_scene("small", n=6))


def test_ctypes_struct_layout_matches_header():
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "gpt.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu\n", sizeof(GptMesh), sizeof(GptMaterials), sizeof(GptCamera),
         sizeof(GptParams), sizeof(GptStats));
  printf("%zu %zu %zu %zu %zu\n", offsetof(GptMesh, x), offsetof(GptMesh, object_primitive_offset_idx),
         offsetof(GptMaterials, material_priority), offsetof(GptParams, rng_mode),
         offsetof(GptStats, trace_launches));
  return 0;
}'''
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(src)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"),
                        os.path.join(d, "t.c")], check=True)
        out = subprocess.run([os.path.join(d, "t")], capture_output=True, text=True).stdout.split()
    sizes = [C.sizeof(t) for t in (cabi.GptMesh, cabi.GptMaterials, cabi.GptCamera,
                                   cabi.GptParams, cabi.GptStats)]
    offs = [cabi.GptMesh.x.offset, cabi.GptMesh.object_primitive_offset_idx.offset,
            cabi.GptMaterials.material_priority.offset, cabi.GptParams.rng_mode.offset,
            cabi.GptStats.trace_launches.offset]
    assert [int(v) for v in out] == sizes + offs


def test_scene_generators_are_deterministic_and_wellformed():
    a, b = S.make_scene("small", n=10), S.make_scene("small", n=10)
    for k in a.mesh:
        assert np.array_equal(a.mesh[k], b.mesh[k]), k
    m = a.mesh
    nt = a.num_triangles
    assert m["object_num_primitives"].sum() == nt
    for k in ("point_1_idx", "point_2_idx", "point_3_idx"):
        assert m[k].min() >= 0 and m[k].max() < m["x"].shape[0]
    for k in ("norm_1_idx", "tex_1_idx"):
        assert m[k].min() >= 0
    assert m["x_norm"][0] == 0 and m["x_tex"][0] == 0      # slot 0 = "absent"
    assert a.materials["tex_r"][0] == 1.0                   # 1x1 white default at offset 0
    c1 = S.make_scene("C1")
    assert c1.num_triangles == 2 * 224 * 224 + 2 + 4        # 100 352 terrain + floor + lights


def test_ray_sets():
    sc = S.make_scene("small", n=10)
    r = rays.camera_rays(sc)
    assert r.shape == (128 * 128, 8) and np.isinf(r[:, 3]).all()
    e = rays.edge_case_rays(sc, 512)
    assert (e[:, 4:7] == 0).any()                            # axis-aligned directions present
    assert np.array_equal(rays.random_rays(100, seed=3), rays.random_rays(100, seed=3))


def test_write_ppm_matches_save_image_format(tmp_path):
    """gpt_write_ppm_u8 writes the bytes the reference's save_image writes
    (src/util/image_util.h:50-72): header, then "r g b " per pixel, no line breaks."""
    import numpy as np
    from gpu_pathtracer_b200 import api
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, size=(7, 5, 3), dtype=np.uint8)
    img[0, 0] = (0, 9, 10)
    img[0, 1] = (99, 100, 255)
    path = str(tmp_path / "a.ppm")
    api.write_ppm_u8(path, img)
    want = "P3\n5 7\n255\n" + "".join(f"{int(v)} " for v in img.reshape(-1))
    assert open(path, "rb").read() == want.encode()
