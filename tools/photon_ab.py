"""A/B of the photon pass: wavefront (default) against the one-thread-per-photon megakernel
(GPT_SPPM_MEGAKERNEL=1, separate process) -- identical photon records, stage time.
Usage: PYTHONPATH=. python tools/photon_ab.py [--photons N] (run once per mode; prints a digest)"""
import argparse
import hashlib
import json
import os
import time

import numpy as np

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

ap = argparse.ArgumentParser()
ap.add_argument("--photons", type=int, default=1000000)
ap.add_argument("--passes", type=int, default=3)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
a = ap.parse_args()
sc = S.make_scene("C4", width=a.width, height=a.height)
sc.params["ppm_num_photon_per_pass"] = a.photons
pt = PathTracer(sc)
pt.build()
pt.sppm_init()
res = {"mode": "megakernel" if os.environ.get("GPT_SPPM_MEGAKERNEL") == "1" else "wavefront",
       "photons": a.photons, "passes": []}
h = hashlib.sha256()
for ps in range(a.passes):
    pt.synchronize()
    t0 = time.perf_counter()
    pt.sppm_pass_stages(ps, 1, 0, a.photons)
    pt.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    ph = pt.sppm_export_photons()
    for k in sorted(ph):
        h.update(np.ascontiguousarray(ph[k]).tobytes())
    res["passes"].append({"photon_stage_ms": round(ms, 2), "rays": pt.stats()["rays_traced"],
                          "n_surface": int(ph["n_surface"]) if "n_surface" in ph else None})
res["digest"] = h.hexdigest()[:16]
print(json.dumps(res))
