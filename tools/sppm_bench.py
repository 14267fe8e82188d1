"""SPPM (config 4) stage timing on one GPU: photon trace + trees, eye pass, gather.
Usage: PYTHONPATH=. python tools/sppm_bench.py [--photons N] [--passes K] [--cells C]"""
import argparse
import json
import time

import numpy as np

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

ap = argparse.ArgumentParser()
ap.add_argument("--photons", type=int, default=10_000_000)
ap.add_argument("--passes", type=int, default=4)
ap.add_argument("--cells", type=int, default=91)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--nofog", action="store_true")
a = ap.parse_args()

sc = S.make_scene("C4", cells=a.cells, width=a.width, height=a.height, photons=a.photons,
                  passes=a.passes, fog=not a.nofog)
pt = PathTracer(sc, device=0)
st = pt.build()
t0 = time.perf_counter()
pt.sppm_init()
pt.synchronize()
out = dict(triangles=sc.num_triangles, photons=a.photons, init_ms=pt.stats()["render_ms"],
           init_rays=pt.stats()["rays_traced"], passes=[])
for p in range(a.passes):
    rec = {}
    for name, mask in (("photon", 1), ("eye", 2), ("gather", 4)):
        pt.sppm_pass_stages(p, mask)
        s = pt.stats()
        rec[name + "_ms"] = s["render_ms"]
        rec[name + "_rays"] = s["rays_traced"]
    pt.sppm_update(p)
    rec.update(pt.sppm_stats())
    out["passes"].append(rec)
    print(json.dumps(rec), flush=True)
img = pt.sppm_resolve(0, a.passes)
out["wall_s"] = time.perf_counter() - t0
out["mean_rgb"] = img.mean(axis=(0, 1)).tolist()
tot_ms = sum(r["photon_ms"] + r["eye_ms"] + r["gather_ms"] for r in out["passes"])
tot_rays = sum(r["photon_rays"] + r["eye_rays"] + r["gather_rays"] for r in out["passes"])
out["mrays_per_s"] = tot_rays / tot_ms / 1e3
print(json.dumps({k: v for k, v in out.items() if k != "passes"}))
