"""C5's shade stage in small: the C5 scene with a 1 M-triangle terrain at 1080p and 4 spp (the SSS
object, its 10^5 points and the fog box as in the full config, so a lookup costs what it costs
there).  Prints stage times; `--frames 1 --no-stage` is the form to run under ncu.
Usage: PYTHONPATH=. python tools/c5_proxy.py [--spp 4] [--frames 2] [--no-stage] [--sha]"""
import argparse
import hashlib
import json
import time

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=4)
ap.add_argument("--frames", type=int, default=2)
ap.add_argument("--no-stage", action="store_true")
ap.add_argument("--n", type=int, default=700)
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
a = ap.parse_args()
sc = S.make_scene("C5", n=a.n, width=a.width, height=a.height, spp=a.spp)
pt = PathTracer(sc)
st0 = pt.build()
img = pt.render_pathtrace()
if not a.no_stage:
    pt.set_profiling(time_stages=True)
t0 = time.perf_counter()
for _ in range(a.frames):
    img = pt.render_pathtrace()
ms = (time.perf_counter() - t0) / max(1, a.frames) * 1e3
st = pt.stats()
print(json.dumps({"triangles": sc.num_triangles, "spp": a.spp, "build_all_ms": round(st0["build_all_ms"], 1),
                  "frame_ms": round(ms, 1), "trace_ms": round(st["trace_kernel_ms"], 1),
                  "shade_ms": round(st["shade_kernel_ms"], 1), "rays": st["rays_traced"],
                  "frame_sha1": hashlib.sha1(img.tobytes()).hexdigest()[:12], "mean": float(img.mean())}))
