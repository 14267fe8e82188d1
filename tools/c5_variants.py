"""Which part of C5 costs what: frame time of the C5 scene with / without the SSS object and the
fog box (reduced terrain).  Usage: PYTHONPATH=. python tools/c5_variants.py"""
import json
import time

import numpy as np

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

for label, kw in (("full", {}), ("no_sss_points", dict(sss_pts=0))):
    sc = S.make_scene("C5", n=700, width=1920, height=1080, spp=4, **kw)
    pt = PathTracer(sc)
    pt.build()
    pt.render_pathtrace()
    pt.set_profiling(time_stages=True)
    t0 = time.perf_counter()
    pt.render_pathtrace()
    ms = (time.perf_counter() - t0) * 1e3
    st = pt.stats()
    print(json.dumps({"variant": label, "frame_ms": round(ms, 1), "trace_ms": round(st["trace_kernel_ms"], 1),
                      "shade_ms": round(st["shade_kernel_ms"], 1), "rays": st["rays_traced"]}), flush=True)
    pt.close()
