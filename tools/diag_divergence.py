"""Diagnostic: pairwise 8-bit agreement of {reference CUDA build, its -fmad=false build, the host
replay of the reference headers, our CUDA path} on one scene."""
import os, subprocess, sys, itertools
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer
from tests.refreplay import RefReplay
from tools.run_reference_cuda import read_p3, to_u8

kind, spp = sys.argv[1], int(sys.argv[2])
sc = S.make_scene(kind)
sc.params["pathtracing_sample_size"] = spp
folder = f"/tmp/gpt_diag_{kind}"
master = S.write_scene_files(sc, folder, kind)
imgs = {}
for name in ("gpu_path_tracing_100", "gpu_path_tracing_100_nofma"):
    for rep in (0, 1):
        out = os.path.join(folder, f"{name}_{rep}")
        r = subprocess.run([os.path.join(ROOT, "oracle", "_ref", name), master, out], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout[-2000:]
        imgs[f"{name[17:] or 'fma'}#{rep}"] = read_p3(out + ".ppm")
pt = PathTracer(sc); pt.build()
imgs["ours"] = to_u8(pt.render_pathtrace())
imgs["replay"] = to_u8(RefReplay(sc).render_pathtrace())
for a, b in itertools.combinations(imgs, 2):
    d = np.abs(imgs[a] - imgs[b])
    print(f"{a:12s} vs {b:12s}: equal {np.mean(d == 0):.4f}  within1 {np.mean(d <= 1):.4f}  >8 levels {np.mean(d.max(-1) > 8):.4f}")
